"""CPU: the C-ABI library loads without a GPU and exports every symbol include/b200md.h declares;
compute entry points fail loudly without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "b200md.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200md_[a-z0-9_]+)\s*\(", txt)))


def test_exports_match_header(pkg):
    lib = pkg.b200md.lib()
    syms = declared_symbols()
    assert len(syms) >= 35
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_no_cpu_fallback(pkg):
    B = pkg.b200md
    try:
        ctx = B.Context(0)
    except B.B200MDError as e:
        assert "no CUDA device" in str(e) or "not sm_100" in str(e)
    else:
        ctx.close()
        pytest.skip("GPU present")


def test_product_does_not_import_oracle():
    pk = os.path.join(ROOT, "lammps-kokkos-port_b200")
    for dp, _, fs in os.walk(pk):
        for f in fs:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(dp, f)).read()
                assert "md_oracle" not in src and "from oracle" not in src and "import oracle" not in src, f
