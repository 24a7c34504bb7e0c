"""b200_compute_thermo_static: the allocation-free variant of b200_compute_thermo (scratch in a file-scope __device__ array,
include/hoomd_b200.h) against the existing entry point, through the C ABI, two calls in a row.  Collected last on purpose
(pytest -x: a failure here must not hide other tests); passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(N, seed, pitch):
    rng = np.random.default_rng(seed)
    vel = rng.normal(0, 1.1, (N, 4)).astype(np.float32)
    vel[:, 3] = rng.uniform(0.5, 2.0, N).astype(np.float32)          # mass
    force = rng.normal(0, 5.0, (N, 4)).astype(np.float32)            # .w = per-particle potential energy
    virial = rng.normal(0, 3.0, (6, pitch)).astype(np.float32)
    return vel, force, virial


def test_thermo_static_matches_thermo_two_calls_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from hoomd_b200 import _lib
    L = _lib.lib()
    # (N, partial blocks, group): the second call uses FEWER blocks and other data than the first, so the rows the first
    # call left in the shared array (stale partial sums) must not leak into it; the third is the maximum size
    cases = [(50000, 196, False), (7777, 31, True), (300000, 1024, False)]
    for k, (N, nb, use_group) in enumerate(cases):
        pitch = N + (16 - (N & 15))
        vel, force, virial = _inputs(N, 100 + k, pitch)
        d_vel = torch.from_numpy(vel).cuda()
        d_f = torch.from_numpy(force).cuda()
        d_vir = torch.from_numpy(virial).cuda()
        group, gsize = None, N
        if use_group:
            members = np.arange(0, N, 3, dtype=np.uint32)             # sparse group: every third particle
            d_grp = torch.from_numpy(members.astype(np.int32)).cuda()
            group, gsize = d_grp.data_ptr(), len(members)
        d_partial = torch.full((nb * 9,), float("nan"), dtype=torch.float64, device="cuda")
        out_ref = torch.zeros(9, dtype=torch.float64, device="cuda")
        out_new = torch.full((9,), float("nan"), dtype=torch.float64, device="cuda")
        assert L.b200_compute_thermo(out_ref.data_ptr(), d_partial.data_ptr(), nb, d_vel.data_ptr(), d_f.data_ptr(),
                                     d_vir.data_ptr(), pitch, group, gsize, None) == 0
        assert L.b200_compute_thermo_static(out_new.data_ptr(), nb, d_vel.data_ptr(), d_f.data_ptr(), d_vir.data_ptr(),
                                            pitch, group, gsize, None) == 0
        torch.cuda.synchronize()
        a, b = out_ref.cpu().numpy(), out_new.cpu().numpy()
        # same kernels, same fixed two-pass summation order: bit-identical (the tolerance of the thermo tests is 1e-12)
        assert np.array_equal(a, b), (k, a, b)
        # and both agree with a float64 numpy sum
        sel = np.arange(0, N, 3) if use_group else np.arange(N)
        v = vel[sel].astype(np.float64)
        ke = 0.5 * (v[:, 3] * (v[:, :3] ** 2).sum(1)).sum()
        assert abs(b[0] - ke) <= 1e-10 * abs(ke)
        assert abs(b[1] - force[sel, 3].astype(np.float64).sum()) <= 1e-9 * N


def test_thermo_static_rejects_sizes_beyond_the_maximum(capfd):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from hoomd_b200 import _lib
    L = _lib.lib()
    out = torch.zeros(9, dtype=torch.float64, device="cuda")
    x = torch.zeros((16, 4), dtype=torch.float32, device="cuda")
    assert L.b200_compute_thermo_static(out.data_ptr(), 1025, x.data_ptr(), x.data_ptr(), None, 0, None, 16, None) == 1
    assert L.b200_compute_thermo_static(out.data_ptr(), 0, x.data_ptr(), x.data_ptr(), None, 0, None, 16, None) == 1
    assert "outside [1, 1024]" in capfd.readouterr().err
    torch.cuda.synchronize()
