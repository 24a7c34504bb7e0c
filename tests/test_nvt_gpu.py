"""GPU parity of md.integrate.nvt (the thermostat half: TwoStepNVTMTK + its ComputeThermo) through the
C ABI: the engine in its three stepping modes against the golden vectors generated from the reference, the driver-level
entry points (step one / step two bit for bit, advanceThermostat to the last bit of xi on exactly representable sums),
the chunk replay after an Nmax overflow, and a long run's temperature control."""
import ctypes as C

import numpy as np
import pytest

import golden_util as G
from hoomd_b200 import workloads

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hb():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hoomd_b200
    return hoomd_b200


@pytest.fixture(scope="module")
def oc():
    from oracle import oracle_c
    return oracle_c


@pytest.mark.parametrize("use_graph", [0, 2, 3])
def test_golden(hb, use_graph):
    """host-driven (0), conditional graph (2, the default) and stream launches (3) against tests/golden/nvt.npz"""
    def make(pos, box, ntypes, vel=None, image=None):
        s = hb.System(pos, box, ntypes=ntypes, vel4=vel, image=image)
        s.set_tuning(use_graph=use_graph)
        return s
    G.check_nvt(make, bit_exact=False)


def test_stepping_modes_agree(hb):
    """The thermostat inside k_md_step (last CTA reduces the kinetic energy) and the stand-alone kernels of the host-driven
    loop sum the same doubles in different orders: the FP32 temperature, hence xi, may differ in the last bit, nothing more."""
    pos, vel, L = workloads.sc_lattice(12, (1.0 / 0.5) ** (1.0 / 3.0), kT=2.0, seed=5)
    out = []
    for mode in (0, 2):
        s = hb.System(pos, (L, L, L), vel4=vel)
        s.setup_lj(r_cut=2.5, r_buff=0.4)
        s.setup_integrator(0.005)
        s.add_nvt(1.2, 0.5)
        s.set_tuning(use_graph=mode)
        s.run(50)
        out.append((s.particles(), s.nvt_state()))
    (pa, xa), (pb, xb) = out
    assert np.allclose(xa, xb, rtol=1e-5, atol=1e-8), (xa, xb)
    assert np.abs(pa["pos"][:, :3] - pb["pos"][:, :3]).max() < 1e-4
    assert np.abs(pa["vel"][:, :3] - pb["vel"][:, :3]).max() < 1e-3


def test_vs_oracle_fresh_system_with_overflow(hb, oc):
    """dense lattice start: Nmax grows during the first chunks (device-driven replay restores the thermostat state from
    its checkpoint), several rebuilds; state against the C restatement (pinned bit-exact against the reference)"""
    pos, vel, L = workloads.sc_lattice(10, (1.0 / 0.7) ** (1.0 / 3.0), kT=3.0, seed=9)
    ref = oc.OracleSystem(pos, (L, L, L), ntypes=1, vel4=vel)
    gpu = hb.System(pos, (L, L, L), vel4=vel)
    for s in (ref, gpu):
        s.setup_lj(r_cut=2.5, r_buff=0.3)
        s.setup_integrator(0.004)
        s.add_nvt(1.5, 0.3)
        s.run(120)
    assert gpu.stats()["overflows"] >= 1
    xa, xb = ref.nvt_state(), gpu.nvt_state()
    assert np.allclose(xa, xb, rtol=2e-5, atol=1e-7), (xa, xb)
    pa, pb = ref.particles(), gpu.particles()
    assert np.abs(pa["pos"][:, :3] - pb["pos"][:, :3]).max() < 5e-4



def test_drivers(hb, oc):
    """b200_nvt_mtk_step_one / _step_two on raw device buffers: bit for bit against the restatement for a thermostat factor
    != 1, wrap and image flags included; b200_nvt_mtk_advance: xi, eta, exp_fac from the device's own two-pass sum"""
    import torch
    from hoomd_b200 import _lib
    L_ = _lib.lib()
    N, L = 5000, 9.0
    rng = np.random.default_rng(3)
    pos = workloads.random_system(N, L, seed=8)
    pos[:40, 0] = np.float32(L / 2 - 1e-5)
    pos[40:80, 2] = np.float32(-L / 2 + 1e-6)
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = rng.normal(0, 1.3, (N, 3))
    vel[:, 3] = rng.uniform(0.5, 3.0, N)
    accel = rng.normal(0, 30.0, (N, 3)).astype(np.float32)
    fnet = rng.normal(0, 25.0, (N, 4)).astype(np.float32)
    dt, kT, tau = 0.005, 1.1, 0.4

    # the restatement, driven through a few thermostat updates so that exp_fac != 1
    o = oc.OracleSystem(pos, (L, L, L), ntypes=1, vel4=vel)
    o.setup_lj(r_cut=1.0, r_buff=0.1)
    o.setup_integrator(dt)
    o.add_nvt(kT, tau)
    o.set_accel(accel)

    state = torch.tensor([0.0, 0.0, 1.0, 0.0], dtype=torch.float32, device="cuda")
    d_pos = torch.from_numpy(pos).cuda()
    d_vel = torch.from_numpy(vel).cuda()
    d_acc = torch.from_numpy(accel).cuda()
    d_img = torch.zeros((N, 3), dtype=torch.int32, device="cuda")
    d_f = torch.from_numpy(fnet).cuda()
    nb = 64
    d_part = torch.zeros(nb, dtype=torch.float64, device="cuda")
    box = _lib.make_box(L, L, L)
    ndof = 3 * N - 3
    for it in range(3):
        o.step_one(it)
        assert L_.b200_nvt_mtk_step_one(d_pos.data_ptr(), d_vel.data_ptr(), d_acc.data_ptr(), d_img.data_ptr(), None, N,
                                        C.byref(box), state.data_ptr(), C.c_float(dt), None) == 0
        assert L_.b200_nvt_mtk_advance(state.data_ptr(), d_part.data_ptr(), nb, d_vel.data_ptr(), None, N, ndof,
                                       C.c_float(kT), C.c_float(tau), C.c_float(dt), None) == 0
        torch.cuda.synchronize()
        p = o.particles()
        assert np.array_equal(d_pos.cpu().numpy().view(np.uint32), p["pos"].view(np.uint32)), it
        assert np.array_equal(d_vel.cpu().numpy().view(np.uint32), p["vel"].view(np.uint32)), it
        assert np.array_equal(d_img.cpu().numpy(), p["image"]), it
        st = state.cpu().numpy()
        assert np.allclose(st[:2], o.nvt_state(), rtol=3e-7, atol=1e-9), (it, st, o.nvt_state())
        assert st[2] == np.float32(np.exp(np.float64(np.float32(np.float32(-0.5) * st[0]) * np.float32(dt))))
        # keep the two thermostats on exactly the same factor for the bit-exact comparison of the next half steps
        state[:2] = torch.from_numpy(o.nvt_state()).cuda()
        xi = o.nvt_state()[0]
        state[2] = float(np.float32(np.exp(np.float64(np.float32(np.float32(-0.5) * xi) * np.float32(dt)))))
        o.step_two(it, fnet)
        assert L_.b200_nvt_mtk_step_two(d_vel.data_ptr(), d_acc.data_ptr(), None, N, d_f.data_ptr(), state.data_ptr(),
                                        C.c_float(dt), None) == 0
        torch.cuda.synchronize()
        p = o.particles()
        assert np.array_equal(d_vel.cpu().numpy().view(np.uint32), p["vel"].view(np.uint32)), it
        assert np.array_equal(d_acc.cpu().numpy().view(np.uint32), p["accel"].view(np.uint32)), it


def test_temperature_control(hb):
    """4000 steps from kT 2.5 towards kT 1.0: the running mean temperature of the second half sits on the set point"""
    pos, vel, L = workloads.sc_lattice(14, (1.0 / 0.5) ** (1.0 / 3.0), kT=2.5, seed=2)
    s = hb.System(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=2.5, r_buff=0.4)
    s.setup_integrator(0.005)
    s.add_nvt(1.0, 0.5)
    s.run(2000)
    temps = []
    for _ in range(20):
        s.run(100)
        temps.append(s.thermo()["T"])
    assert abs(np.mean(temps) - 1.0) < 0.03, temps
    xi, eta = s.nvt_state()
    assert np.isfinite(xi) and np.isfinite(eta)


def test_errors(hb):
    pos, vel, L = workloads.sc_lattice(6, 1.3, kT=1.0, seed=1)
    s = hb.System(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=2.5, r_buff=0.4)
    s.setup_integrator(0.005)
    with pytest.raises(RuntimeError):
        s.add_nvt(0.0, 0.5)
    with pytest.raises(RuntimeError):
        s.nvt_state()


@pytest.mark.parametrize("method", ["langevin", "nvt"])
def test_temperature_ramp(hb, oc, method):
    """b200_sys_set_temperature once per step (what a hoomd.variant set point amounts to) against the restatement"""
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.4) ** (1.0 / 3.0), kT=2.0, seed=5)
    o = oc.OracleSystem(pos, (L, L, L), ntypes=1, vel4=vel)
    g = hb.System(pos, (L, L, L), vel4=vel)
    for s in (o, g):
        s.setup_lj(r_cut=2.5, r_buff=0.4)
        s.setup_integrator(0.005)
        s.add_langevin(2.0, 11) if method == "langevin" else s.add_nvt(2.0, 0.4)
        for ts in range(60):
            s.set_kT(np.float32(2.0 - 1.2 * ts / 60.0))
            s.run(1)
    a, b = o.particles(), g.particles()
    assert np.abs(a["pos"][:, :3] - b["pos"][:, :3]).max() < 2e-4
    assert np.abs(a["vel"][:, :3] - b["vel"][:, :3]).max() < 2e-3
    if method == "nvt":
        assert np.allclose(o.nvt_state(), g.nvt_state(), rtol=2e-5, atol=1e-7)
    with pytest.raises(RuntimeError):
        n = hb.System(pos, (L, L, L), vel4=vel)
        n.setup_lj(r_cut=2.5, r_buff=0.4)
        n.setup_integrator(0.005)
        n.add_nve()
        n.set_kT(1.0)
