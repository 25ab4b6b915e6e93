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


def test_every_option_is_documented(pkg):
    """the runtime options are the library's whole tuning surface (no environment variables): every keyword the Python
    mirror knows must be in the header's keyword table and in INTEGRATION.md's option table, and the other way round"""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "b200md.h")).read()
    integ = open(os.path.join(root, "INTEGRATION.md")).read()
    table = header[header.index("Keywords (value as text)"):header.index("int b200md_set_option")]
    in_header = set(re.findall(r"^ \*   (\w+) ", table, re.M))
    in_integ = set(re.findall(r"^\| `(\w+)` \|", integ[integ.index("## 7. Runtime options"):], re.M))
    keys = set(pkg.b200md.OPTION_KEYS)
    assert keys == in_header, (keys - in_header, in_header - keys)
    assert keys == in_integ, (keys - in_integ, in_integ - keys)
