"""b200md_ke_sums_static (include/b200md.h, csrc/ke_static.cu): the kinetic sums of thermo in one launch that
allocates nothing — the per-block partial sums live in a static device array of the library.  The entry point has no
counterpart in the reference (it was added); it is checked against the
existing entry point, b200md_thermo_compute, on the golden systems, and against float64 sums on the host.

Collected last on purpose: a failure here must not hide any other test."""
import os

import numpy as np
import pytest

import common as Cm
from test_oracle_golden import CASES

T_TOL = 1e-8          # the thermo tolerance of tests/test_gpu_parity.py
SUM_TOL = 1e-12       # float64 sums of the same numbers in another order


def test_static_scratch_limit_is_enforced(pkg):
    """sizes beyond the static array are refused with a message naming the limit (checked before any CUDA call)"""
    B = pkg.b200md
    for n in (B.KE_STATIC_MAX_ATOMS + 1, -1):
        with pytest.raises(B.B200MDError, match="B200MD_KE_STATIC_MAX_ATOMS"):
            B.ke_sums_static(n, max(n, 0), 0, 0, 0, 0)
    with pytest.raises(B.B200MDError, match="stride"):
        B.ke_sums_static(100, 50, 0, 0, 0, 0)


def device_inputs(torch, v_aos, type_, mass, stride):
    n = v_aos.shape[0]
    v = torch.zeros((3, stride), dtype=torch.float64, device="cuda:0")
    v[:, :n] = torch.from_numpy(np.ascontiguousarray(v_aos.T)).to("cuda:0")
    t = torch.from_numpy(np.ascontiguousarray(type_, dtype=np.int32)).to("cuda:0")
    m = torch.from_numpy(np.ascontiguousarray(mass, dtype=np.float64)).to("cuda:0")
    return v, t, m


def host_sums(v_aos, type_, mass):
    m = np.asarray(mass, dtype=np.float64)[np.asarray(type_)]
    return np.array([np.sum((v_aos * v_aos).sum(axis=1) * m), np.sum(m * v_aos[:, 0]), np.sum(m * v_aos[:, 1]),
                     np.sum(m * v_aos[:, 2]), np.sum(m)])


@pytest.mark.gpu
@pytest.mark.parametrize("name,kind", CASES)
def test_ke_sums_static_equals_thermo_compute(pkg, oracle_mod, name, kind):
    import torch
    from test_gpu_parity import make_ctx
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_HALF_NEWTON, 2)
    ctx.setup()
    ctx.run(10, s["dt"], 0)
    th = ctx.thermo()                                  # the existing entry point: temp, ke
    a = ctx.get_atoms(("v", "type"))
    mvv2e, boltz = ctx.units["mvv2e"], ctx.units["boltz"]
    mass = ctx.mass.copy()
    ctx.close()

    n = a["v"].shape[0]
    stride = n + 37                                    # any stride >= n
    v, t, m = device_inputs(torch, a["v"], a["type"], mass, stride)
    stream = torch.cuda.current_stream().cuda_stream
    want = host_sums(a["v"], a["type"], mass)
    dof = 3.0 * n - 3.0
    outs = []
    for call in range(2):                              # two calls in a row share the static array
        out = torch.full((5,), float("nan"), dtype=torch.float64, device="cuda:0")
        B.ke_sums_static(n, stride, v.data_ptr(), t.data_ptr(), m.data_ptr(), out.data_ptr(), stream)
        got = out.cpu().numpy()
        outs.append(got)
        ke = 0.5 * mvv2e * got[0]                      # Thermo::compute_ke from ComputeTemp's sum
        temp = got[0] * mvv2e / (dof * boltz)
        assert abs(ke - th["ke"]) <= T_TOL * abs(th["ke"]), (call, ke, th["ke"])
        assert abs(temp - th["temp"]) <= T_TOL * abs(th["temp"]), (call, temp, th["temp"])
        scale = np.array([want[0], want[4], want[4], want[4], want[4]])   # momentum sums are ~0: absolute, per unit mass
        assert np.all(np.abs(got - want) <= SUM_TOL * np.abs(scale)), (call, got, want)
    assert np.array_equal(outs[0], outs[1])            # fixed summation order: the same bits


@pytest.mark.gpu
def test_ke_sums_static_many_blocks_two_types(pkg):
    """more blocks than one reduction pass of 256 threads, a ragged last block, two masses, no type array, and a
    smaller call after a larger one (the array keeps the larger call's rows)"""
    import torch
    B = pkg.b200md
    rng = np.random.default_rng(20261105)
    n = 1_500_003
    v_aos = rng.normal(size=(n, 3))
    type_ = rng.integers(1, 3, size=n).astype(np.int32)
    mass = np.array([0.0, 1.0, 63.55])
    v, t, m = device_inputs(torch, v_aos, type_, mass, n)
    stream = torch.cuda.current_stream().cuda_stream
    out = torch.empty((5,), dtype=torch.float64, device="cuda:0")

    def check(count, types):
        want = host_sums(v_aos[:count], type_[:count] if types else np.ones(count, dtype=np.int32), mass)
        scale = np.array([want[0], want[4], want[4], want[4], want[4]])
        res = []
        for _ in range(2):
            out.fill_(float("nan"))
            B.ke_sums_static(count, n, v.data_ptr(), t.data_ptr() if types else 0, m.data_ptr(), out.data_ptr(), stream)
            res.append(out.cpu().numpy())
            assert np.all(np.abs(res[-1] - want) <= SUM_TOL * np.abs(scale)), (count, types, res[-1], want)
        assert np.array_equal(res[0], res[1])

    check(n, True)
    check(4097, True)          # two blocks after 367: the stale rows beyond the second must not be read
    check(1, False)
    check(n, False)
    B.ke_sums_static(0, 0, 0, 0, 0, 0)                 # n = 0: set-up only, no launch, no pointer needed
