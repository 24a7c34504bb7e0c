"""md.pair.dpd -- the pairwise DPD thermostat (PotentialPairDPDThermo<EvaluatorPairDPDThermo>) -- against the UNMODIFIED reference
CPU class: forces incl. the per-pair random numbers (Saru keyed by the two tags and seed + timestep) at several time steps and
seeds, energy / virial of the conservative part, a two-type mixture, trajectories under md.integrate.nve with list rebuilds, the
stand-alone driver on caller-owned buffers, and the job-script path."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref():
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    return R


@pytest.fixture(scope="module")
def hb():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hoomd_b200
    return hoomd_b200


def _state(hb, n_side=10, rho=3.0, ntypes=1, seed=3):
    kw = dict(ntypes=2, type_fractions=[0.7, 0.3]) if ntypes == 2 else {}
    pos, vel, L = hb.workloads.sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=1.0, seed=seed, **kw)
    pos[:, :3] += np.random.default_rng(seed).normal(0, 0.12, (len(pos), 3)).astype(np.float32)
    pos[:, :3] = np.clip(pos[:, :3], -L / 2, np.nextafter(np.float32(L / 2), np.float32(0)))
    return pos, vel, L


def _coeffs(ntypes):
    if ntypes == 1:
        return dict(r_cut=1.0, A=25.0, gamma=4.5)
    return dict(r_cut=[[1.0, 0.9], [0.9, 1.1]], A=[[25.0, 40.0], [40.0, 15.0]], gamma=[[4.5, 1.2], [1.2, 3.0]])


def _systems(ref, hb, pos, vel, L, ntypes, kT, seed, dt=0.01):
    c = _coeffs(ntypes)
    r = ref.RefSystem(pos, (L, L, L), ntypes=ntypes, vel4=vel)
    r.setup_dpd_thermostat(kT=kT, seed=seed, r_buff=0.3, **c)
    r.set_virial(True)
    r.setup_integrator(dt)
    r.add_nve()
    g = hb.System(pos, (L, L, L), ntypes=ntypes, vel4=vel)
    g.setup_pair("dpd", r_buff=0.3, **c)
    g.set_dpd_thermostat(kT, seed)
    g.set_virial(True)
    g.setup_integrator(dt)
    g.add_nve()
    return r, g


@pytest.mark.parametrize("ntypes", [1, 2])
def test_dpd_forces_vs_reference(ref, hb, ntypes):
    pos, vel, L = _state(hb, ntypes=ntypes)
    for seed, ts in ((7, 0), (7, 3), (12345, 1000), (0xffffffff, 77)):
        r, g = _systems(ref, hb, pos, vel, L, ntypes, kT=1.3, seed=seed)
        r.compute_forces(ts)
        g.compute_forces(ts)
        fr, vr = r.forces(True)
        fg, vg = g.forces(True)
        # the sum of absolute pair terms is not available from the reference class; |f_i| (>= 1) bounds it from below, so
        # this is STRICTER than north_star's 1e-5 of sum_j |f_ij|
        scale = np.maximum(np.linalg.norm(fr[:, :3], axis=1), 1.0)
        d = np.linalg.norm(fr[:, :3].astype(np.float64) - fg[:, :3].astype(np.float64), axis=1)
        assert (d / scale).max() < 2e-5, (seed, ts, (d / scale).max())
        assert np.abs(fr[:, 3] - fg[:, 3]).max() <= 1e-5 * np.abs(fr[:, 3]).max()
        assert np.abs(vr - vg).max() <= 1e-5 * np.abs(vr).max()
        assert np.array_equal(r.nlist()[0], g.nlist(counts_only=True)[0])
    # the random force depends on the time step and on the seed; energy and virial (conservative part) do not
    r, g = _systems(ref, hb, pos, vel, L, ntypes, kT=1.3, seed=7)
    g.compute_forces(0)
    f0, v0 = g.forces(True)
    g.compute_forces(1)
    f1, v1 = g.forces(True)
    assert np.abs(f0[:, :3] - f1[:, :3]).max() > 1.0
    assert np.array_equal(f0[:, 3], f1[:, 3]) and np.array_equal(v0, v1)
    # momentum conservation: every pair force is antisymmetric incl. drag and noise
    assert np.abs(f0[:, :3].astype(np.float64).sum(axis=0)).max() < 1e-2


def test_dpd_trajectory_and_temperature(ref, hb):
    pos, vel, L = _state(hb, n_side=12)
    r, g = _systems(ref, hb, pos, vel, L, 1, kT=1.0, seed=42)
    for s in (r, g):
        s.set_sorter(40)
    r.run(120)
    g.run(60)
    g.run(60)
    a, b = r.particles(), g.particles()
    assert np.array_equal(a["tag"], b["tag"])
    d = a["pos"][:, :3] - b["pos"][:, :3]
    d -= np.round(d / L) * L
    assert np.abs(d).max() < 2e-3
    gs = g.stats()
    assert r.nlist_stats()["updates"] == gs["updates"] + gs["forced"] and gs["updates"] >= 3
    assert g.stepping()["multistep"] is False
    # the thermostat holds the set point: start cold, end near kT
    vel0 = vel.copy()
    vel0[:, :3] *= 0.2
    g2 = hb.System(pos, (L, L, L), vel4=vel0)
    g2.setup_pair("dpd", r_buff=0.3, **_coeffs(1))
    g2.set_dpd_thermostat(1.0, 5)
    g2.setup_integrator(0.01)
    g2.add_nve()
    t0 = g2.thermo()["T"]
    g2.run(1500)
    t1 = g2.thermo()["T"]
    assert t0 < 0.2 and 0.9 < t1 < 1.12, (t0, t1)
    # switching the thermostat off gives the conservative force again
    g2.set_dpd_thermostat(0.0, 0)
    g2.compute_forces(g2.timestep)
    fa = g2.forces(False)[0]
    g2.compute_forces(g2.timestep + 1)
    assert np.array_equal(fa, g2.forces(False)[0])


def test_dpd_driver_on_caller_buffers(ref, hb):
    """b200_compute_dpd_thermo_forces with caller-owned buffers: the neighbor list comes from the reference"""
    import torch
    L_ = hb._lib.lib()
    pos, vel, L = _state(hb, n_side=8, ntypes=2)
    N = len(pos)
    r, _g = _systems(ref, hb, pos, vel, L, 2, kT=0.8, seed=99)
    r.compute_forces(5)
    fr, vr = r.forces(True)
    nn, head, nl = r.nlist()
    c = _coeffs(2)
    prm = hb._lib.pair_params("dpd", 2, A=c["A"], gamma=c["gamma"])
    prm2 = np.ascontiguousarray(prm.reshape(4, -1)[:, :2])
    rcutsq = (np.asarray(c["r_cut"], np.float32) ** 2).astype(np.float32).reshape(-1)
    tag = np.arange(N, dtype=np.uint32)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    ptr = lambda t: C.c_void_p(t.data_ptr())
    pitch = N + 5
    dpos, dvel, dtag = d(pos), d(vel), d(tag.view(np.int32))
    dnn, dhead, dnl = d(nn.view(np.int32)), d(head.view(np.int32)), d(nl.view(np.int32))
    dprm, drc = d(prm2), d(rcutsq)
    dforce = torch.zeros((N, 4), dtype=torch.float32, device="cuda")
    dvir = torch.zeros((6, pitch), dtype=torch.float32, device="cuda")
    a = hb._lib.PairArgs()
    a.d_force, a.d_virial, a.virial_pitch, a.N, a.d_pos = ptr(dforce), ptr(dvir), pitch, N, ptr(dpos)
    a.box = hb._lib.make_box(L, L, L)
    a.d_n_neigh, a.d_nlist, a.d_head_list, a.d_rcutsq, a.ntypes = ptr(dnn), ptr(dnl), ptr(dhead), ptr(drc), 2
    a.compute_virial = 1
    t = hb._lib.DpdThermoArgs()
    L_.b200_langevin_hash_seed.restype = C.c_uint
    hashed = L_.b200_langevin_hash_seed(C.c_uint(99))    # the driver takes m_seed as setSeed leaves it
    t.d_vel, t.d_tag, t.seed, t.timestep, t.T, t.deltaT = ptr(dvel), ptr(dtag), hashed, 5, 0.8, 0.01
    for tpp in (0, 1, 2, 8):
        a.threads_per_particle = tpp
        dforce.zero_()
        assert L_.b200_compute_dpd_thermo_forces(C.byref(a), ptr(dprm), C.byref(t), None) == 0
        torch.cuda.synchronize()
        f = dforce.cpu().numpy()
        scale = np.maximum(np.linalg.norm(fr[:, :3], axis=1), 1.0)
        assert (np.linalg.norm(fr[:, :3] - f[:, :3], axis=1) / scale).max() < 2e-5, tpp
        assert np.abs(dvir.cpu().numpy()[:, :N] - vr).max() <= 1e-5 * np.abs(vr).max()
    a.threads_per_particle = 3
    assert L_.b200_compute_dpd_thermo_forces(C.byref(a), ptr(dprm), C.byref(t), None) != 0
    t.T = 0.0
    a.threads_per_particle = 0
    assert L_.b200_compute_dpd_thermo_forces(C.byref(a), ptr(dprm), C.byref(t), None) != 0


def test_dpd_in_job_script(ref, hb):
    from hoomd_b200 import script as hoomd
    hoomd.install()
    hoomd.context.initialize()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=0.7, type_name="A"), n=10)
    N = 1000
    snap = hoomd.context.current.system.take_snapshot()
    snap.particles.velocity[:] = np.random.default_rng(8).normal(0, 1.0, (N, 3))
    snap.particles.position[:] += np.random.default_rng(9).normal(0, 0.05, (N, 3))
    Lb = hoomd.context.current.system.box.Lx
    snap.particles.position[:] = np.clip(snap.particles.position, -Lb / 2, np.nextafter(np.float32(Lb / 2), np.float32(0)))
    hoomd.context.current.system.restore_snapshot(snap)
    p0, v0 = snap.particles.position.copy(), snap.particles.velocity.copy()
    L = hoomd.context.current.system.box.Lx
    nl = hoomd.md.nlist.cell(r_buff=0.3)
    dpd = hoomd.md.pair.dpd(r_cut=1.0, nlist=nl, kT=1.0, seed=3)
    dpd.pair_coeff.set("A", "A", A=25.0, gamma=4.5)
    with pytest.raises(RuntimeError):
        dpd.set_params(mode="shift")
    hoomd.md.integrate.mode_standard(dt=0.01)
    hoomd.md.integrate.nve(group=hoomd.group.all())
    hoomd.run(50)
    p1 = hoomd.context.current.system.take_snapshot().particles.position
    pos = np.zeros((N, 4), np.float32)
    pos[:, :3] = p0
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = v0
    vel[:, 3] = 1.0
    r = ref.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_dpd_thermostat(r_cut=1.0, A=25.0, gamma=4.5, kT=1.0, seed=3, r_buff=0.3)
    r.setup_integrator(0.01)
    r.add_nve()
    r.set_sorter(300)
    r.run(50)
    a = r.particles()
    d = a["pos"][np.argsort(a["tag"])][:, :3] - p1
    d -= np.round(d / L) * L
    assert np.abs(d).max() < 1e-3
