"""md.integrate.npt (TwoStepNPTMTK, Martyna-Tobias-Klein barostat + thermostat) against the UNMODIFIED reference CPU class
(hoomd/md/TwoStepNPTMTK.cc, compiled into oracle/_ref): box lengths, barostat / thermostat variables, log quantities and
trajectories over a run with neighbor-list rebuilds in a changing box; every coupling mode the orthorhombic box allows, NPH,
rescale_all with a sparse group, the damping factor, a second run() continuing the first, and the stand-alone drivers
(b200_npt_mtk_step_one / _wrap / _step_two / _rescale, b200_compute_thermo_tensor) on caller-owned buffers."""
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


def _state(hb, n_side=12, rho=0.5, seed=3):
    pos, vel, L = hb.workloads.sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=1.0, seed=seed)
    pos[:, :3] += np.random.default_rng(seed).normal(0, 0.05, (len(pos), 3)).astype(np.float32)
    return pos, vel, L


def _build(cls, pos, vel, L, members=None, **npt):
    is_ref = cls.__name__ == "RefSystem"
    s = cls(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=2.5, r_buff=0.4, full=True) if is_ref else s.setup_lj(r_cut=2.5, r_buff=0.4)
    s.setup_integrator(0.005)
    if members is not None and is_ref:
        s.set_group(members)
    s.add_npt(**npt)
    if members is not None and not is_ref:
        s.set_group(members)
    return s


CASES = [
    dict(kT=1.0, tau=0.5, P=1.5, tauP=0.8),                                         # isotropic
    dict(kT=1.2, tau=0.4, P=0.7, tauP=1.0, couple="none"),                          # three independent lengths
    dict(kT=1.0, tau=0.5, P=2.0, tauP=0.8, couple="xy"),                            # xy coupled, z free
    dict(kT=1.0, tau=0.5, P=1.0, tauP=0.8, couple="xyz", z=False),                  # degrades to xy, z fixed
    dict(kT=1.0, tau=0.5, P=1.0, tauP=0.8, couple="yz", x=False, gamma=0.05),       # damping of the box momenta
    dict(kT=1.0, tau=1.0, P=1.2, tauP=0.6, nph=True),                               # no thermostat
]


@pytest.mark.parametrize("case", range(len(CASES)))
def test_npt_run_vs_reference(ref, hb, case):
    kw = CASES[case]
    pos, vel, L = _state(hb)
    N = len(pos)
    nsteps = 120
    r = _build(ref.RefSystem, pos, vel, L, **kw)
    g = _build(hb.System, pos, vel, L, **kw)
    for s in (r, g):
        s.set_sorter(50)
    # two runs on our side: the barostat / thermostat variables and the cached full-step thermo carry over
    r.run(nsteps)
    g.run(nsteps // 3)
    g.run(nsteps - nsteps // 3)
    Lr, Lg = np.array(r.box_lengths(), np.float64), np.array(g.box_lengths(), np.float64)
    assert np.abs(Lr - L).max() > 1e-3, "the box did not move: the barostat was not exercised"
    assert np.abs(Lg - Lr).max() <= 2e-5 * L, (Lg, Lr)
    if not kw.get("z", True):
        assert Lg[2] == np.float32(L) and Lr[2] == np.float32(L)
    if not kw.get("x", True):
        assert Lg[0] == np.float32(L)
    sr, sg = r.npt_state(), g.npt_state()
    # {eta, xi, nu xx..zz}: sums of O(100) FP32 increments along trajectories that differ by force round-off
    assert np.allclose(sg[:10], sr[:10], rtol=2e-3, atol=2e-5), (sg, sr)
    assert np.allclose(sg[10:], sr[10:], rtol=3e-3, atol=1e-2), (sg[10:], sr[10:])
    if kw.get("nph"):
        assert sg[0] == 0.0 and sg[1] == 0.0
    a, b = r.particles(), g.particles()
    assert np.array_equal(a["tag"], b["tag"])
    d = a["pos"][:, :3] - b["pos"][:, :3]
    d -= np.round(d / Lr) * Lr
    assert np.abs(d).max() < 2e-3
    assert np.abs(a["vel"][:, :3] - b["vel"][:, :3]).max() < 1e-2
    gs = g.stats()
    assert r.nlist_stats()["updates"] == gs["updates"] + gs["forced"], (r.nlist_stats(), gs)
    assert gs["updates"] + gs["forced"] >= 3
    # pressure tensor of the final state (full-step ComputeThermo)
    pr, pg = r.pressure_tensor(), g.pressure_tensor()
    assert np.allclose(pg, pr, rtol=5e-3, atol=5e-3), (pg, pr)


def test_npt_first_steps_tight(ref, hb):
    """a handful of steps from identical state: before round-off has time to grow the variables agree to FP32 resolution"""
    pos, vel, L = _state(hb, n_side=10, rho=0.6, seed=8)
    kw = dict(kT=1.1, tau=0.3, P=2.5, tauP=0.5, couple="none")
    r = _build(ref.RefSystem, pos, vel, L, **kw)
    g = _build(hb.System, pos, vel, L, **kw)
    for k in range(4):
        r.run(1)
        g.run(1)
        sr, sg = r.npt_state(), g.npt_state()
        assert np.allclose(sg[:10], sr[:10], rtol=3e-5, atol=1e-7), (k, sg, sr)
        assert np.allclose(np.array(g.box_lengths()), np.array(r.box_lengths()), rtol=3e-7, atol=0)
        a, b = r.particles(), g.particles()
        assert np.abs(a["pos"][:, :3] - b["pos"][:, :3]).max() < 5e-6
        assert np.abs(a["vel"][:, :3] - b["vel"][:, :3]).max() < 5e-6


def test_npt_rescale_all_with_sparse_group(ref, hb):
    pos, vel, L = _state(hb, n_side=10, rho=0.5, seed=5)
    N = len(pos)
    members = np.arange(0, N, 2, dtype=np.uint32)
    mask = np.zeros(N, bool)
    mask[members] = True
    for rescale_all in (False, True):
        kw = dict(kT=1.0, tau=0.5, P=1.5, tauP=0.8, rescale_all=rescale_all)
        r = _build(ref.RefSystem, pos, vel, L, members=members, **kw)
        g = _build(hb.System, pos, vel, L, members=members, **kw)
        r.run(60)
        g.run(60)
        a, b = r.particles(), g.particles()
        oa, ob = np.argsort(a["tag"]), np.argsort(b["tag"])
        Lr = np.array(r.box_lengths(), np.float64)
        assert np.abs(np.array(g.box_lengths()) - Lr).max() <= 2e-5 * L
        d = a["pos"][oa][:, :3] - b["pos"][ob][:, :3]
        d -= np.round(d / Lr) * Lr
        assert np.abs(d).max() < 1e-3
        # non-members keep their velocities; their positions are scaled with the box only under rescale_all
        assert np.array_equal(b["vel"][ob][~mask].view(np.uint32), vel[~mask].view(np.uint32))
        moved = np.abs(b["pos"][ob][~mask][:, :3] - pos[~mask][:, :3]).max()
        assert (moved > 1e-4) == rescale_all, (rescale_all, moved)
        assert np.allclose(g.npt_state()[:10], r.npt_state()[:10], rtol=2e-3, atol=2e-5)


def test_npt_rejects_what_the_box_cannot_do(hb):
    pos, vel, L = _state(hb, n_side=6)
    s = hb.System(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=2.5, r_buff=0.4)
    s.setup_integrator(0.005)
    L_ = hb._lib.lib()
    assert L_.b200_sys_add_npt(s.h, C.c_float(1.0), C.c_float(0.5), C.c_float(1.0), C.c_float(0.5), 4, 7 | 8, 0, 0,
                               C.c_float(0.0)) != 0      # baro_xy needs a triclinic box
    assert b"triclinic" in L_.b200_sys_last_error(s.h)
    assert L_.b200_sys_add_npt(s.h, C.c_float(1.0), C.c_float(0.5), C.c_float(1.0), C.c_float(0.5), 9, 7, 0, 0,
                               C.c_float(0.0)) != 0
    assert L_.b200_sys_add_npt(s.h, C.c_float(-1.0), C.c_float(0.5), C.c_float(1.0), C.c_float(0.5), 4, 7, 0, 0,
                               C.c_float(0.0)) != 0


def test_npt_drivers_on_caller_buffers(hb):
    """the four stand-alone drivers + the tensor ComputeThermo against an FP32 numpy restatement of the reference loops
    (TwoStepNPTMTK.cc:271-317, 497-530; ComputeThermo.cc:224-240): bit-exact positions / velocities / images."""
    import torch
    L_ = hb._lib.lib()
    rng = np.random.default_rng(12)
    N = 5000
    f32 = np.float32
    Lbox = np.array([11.0, 12.5, 9.75], f32)
    pos = np.zeros((N, 4), f32)
    pos[:, :3] = rng.uniform(-0.5, 0.5, (N, 3)).astype(f32) * Lbox
    vel = rng.normal(0, 1, (N, 4)).astype(f32)
    vel[:, 3] = rng.uniform(0.5, 2.0, N).astype(f32)
    accel = rng.normal(0, 3, (N, 3)).astype(f32)
    force = rng.normal(0, 3, (N, 4)).astype(f32)
    image = rng.integers(-2, 3, (N, 3)).astype(np.int32)
    group = np.sort(rng.choice(N, N // 2, replace=False)).astype(np.uint32)
    mv = np.array([1.0004, 0, 0, 0.9991, 0, 1.0002], f32)
    mr = np.array([1.013, 0, 0, 0.985, 0, 1.021], f32)     # grows x and z enough for some particles to leave the box
    mri = np.array([0.00501, 0, 0, 0.00499, 0, 0.005003], f32)
    fac, dt = f32(0.9987), f32(0.005)

    d = lambda a: torch.from_numpy(a).cuda()
    ptr = lambda t: C.c_void_p(t.data_ptr())
    for rescale_all in (0, 1):
        dp, dv, da, di, dg = d(pos.copy()), d(vel.copy()), d(accel), d(image.copy()), d(group.view(np.int32))
        if rescale_all:
            assert L_.b200_npt_mtk_rescale(N, ptr(dp), *[C.c_float(float(x)) for x in mr], None) == 0
        assert L_.b200_npt_mtk_step_one(ptr(dp), ptr(dv), ptr(da), ptr(dg), len(group), C.c_float(float(fac)),
                                        mv.ctypes.data_as(C.c_void_p), mr.ctypes.data_as(C.c_void_p),
                                        mri.ctypes.data_as(C.c_void_p), C.c_float(float(dt)), rescale_all, None) == 0
        newL = (mr[[0, 3, 5]] * Lbox).astype(f32)
        nb = hb._lib.make_box(*[float(x) for x in newL])
        assert L_.b200_npt_mtk_wrap(N, ptr(dp), ptr(di), C.byref(nb), None) == 0
        torch.cuda.synchronize()
        # restatement
        p, v, im = pos.copy(), vel.copy(), image.copy()
        if rescale_all:
            p[:, 0] = mr[0] * p[:, 0] + mr[1] * p[:, 1] + mr[2] * p[:, 2]
            p[:, 1] = mr[3] * p[:, 1] + mr[4] * p[:, 2]
            p[:, 2] = mr[5] * p[:, 2]
        g = group
        hdt = dt / f32(2.0)
        vg = v[g, :3] + hdt * accel[g]
        vx = mv[0] * vg[:, 0] + mv[1] * vg[:, 1] + mv[2] * vg[:, 2]
        vy = mv[3] * vg[:, 1] + mv[4] * vg[:, 2]
        vz = mv[5] * vg[:, 2]
        vg = np.stack([vx, vy, vz], 1) * fac
        pg = p[g, :3].copy()
        if not rescale_all:
            px = mr[0] * pg[:, 0] + mr[1] * pg[:, 1] + mr[2] * pg[:, 2]
            py = mr[3] * pg[:, 1] + mr[4] * pg[:, 2]
            pz = mr[5] * pg[:, 2]
            pg = np.stack([px, py, pz], 1)
        pg[:, 0] += mri[0] * vg[:, 0] + mri[1] * vg[:, 1] + mri[2] * vg[:, 2]
        pg[:, 1] += mri[3] * vg[:, 1] + mri[4] * vg[:, 2]
        pg[:, 2] += mri[5] * vg[:, 2]
        v[g, :3] = vg
        p[g, :3] = pg
        hi, lo = newL / f32(2.0), -(newL / f32(2.0))
        for a in range(3):
            up, dn = p[:, a] >= hi[a], p[:, a] < lo[a]
            p[up, a] -= newL[a]
            im[up, a] += 1
            p[dn, a] += newL[a]
            im[dn, a] -= 1
        assert (im != image).any(), "no particle crossed the new box: the wrap was not exercised"
        assert np.array_equal(dp.cpu().numpy().view(np.uint32), p.view(np.uint32))
        assert np.array_equal(dv.cpu().numpy().view(np.uint32), v.view(np.uint32))
        assert np.array_equal(di.cpu().numpy(), im)

    # step two
    dv, da, df, dg = d(vel.copy()), d(np.zeros((N, 3), f32)), d(force), d(group.view(np.int32))
    assert L_.b200_npt_mtk_step_two(ptr(dv), ptr(da), ptr(dg), len(group), ptr(df), mv.ctypes.data_as(C.c_void_p),
                                    C.c_float(float(dt)), C.c_float(float(fac)), None) == 0
    torch.cuda.synchronize()
    v, a = vel.copy(), np.zeros((N, 3), f32)
    g = group
    minv = f32(1.0) / v[g, 3]
    a[g] = force[g, :3] * minv[:, None]
    vg = v[g, :3] * fac
    vx = mv[0] * vg[:, 0] + mv[1] * vg[:, 1] + mv[2] * vg[:, 2]
    vy = mv[3] * vg[:, 1] + mv[4] * vg[:, 2]
    vz = mv[5] * vg[:, 2]
    v[g, :3] = np.stack([vx, vy, vz], 1) + (dt / f32(2.0)) * a[g]
    assert np.array_equal(dv.cpu().numpy().view(np.uint32), v.view(np.uint32))
    assert np.array_equal(da.cpu().numpy().view(np.uint32), a.view(np.uint32))

    # tensor thermo: double sums, any order
    pitch = N + 7
    virial = rng.normal(0, 1, (6, pitch)).astype(f32)
    dvel, dfor, dvir, dg = d(vel), d(force), d(virial), d(group.view(np.int32))
    nblocks = 40
    out, part = torch.zeros(13, dtype=torch.float64, device="cuda"), torch.zeros(13 * nblocks, dtype=torch.float64,
                                                                                 device="cuda")
    assert L_.b200_compute_thermo_tensor(ptr(out), ptr(part), nblocks, ptr(dvel), ptr(dfor), ptr(dvir), pitch, ptr(dg),
                                         len(group), None) == 0
    torch.cuda.synchronize()
    vv = vel[group].astype(np.float64)
    m = vv[:, 3]
    exp = [np.sum(m * vv[:, i] * vv[:, j]) for i, j in ((0, 0), (0, 1), (0, 2), (1, 1), (1, 2), (2, 2))]
    exp.append(np.sum(force[group, 3].astype(np.float64)))
    exp += [np.sum(virial[k, group].astype(np.float64)) for k in range(6)]
    assert np.allclose(out.cpu().numpy(), np.array(exp), rtol=1e-12, atol=1e-9)
    assert L_.b200_compute_thermo_tensor(ptr(out), ptr(part), 0, ptr(dvel), ptr(dfor), ptr(dvir), pitch, ptr(dg),
                                         len(group), None) != 0


def test_npt_in_job_script(ref, hb):
    """md.integrate.npt on the job-script surface: the system box follows the barostat, the two log quantities and the box
    quantities are served, a re-configuration in the middle of the job keeps the IntegratorVariables; the whole job agrees
    with the same commands on the reference's CPU classes."""
    from hoomd_b200 import script as hoomd
    hoomd.install()
    hoomd.context.initialize()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.3, type_name="A"), n=9)
    N = 729
    snap = hoomd.context.current.system.take_snapshot()
    snap.particles.velocity[:] = np.random.default_rng(6).normal(0, 1.0, (N, 3))
    hoomd.context.current.system.restore_snapshot(snap)
    p0, v0 = snap.particles.position.copy(), snap.particles.velocity.copy()
    L0 = hoomd.context.current.system.box.Lx
    nl = hoomd.md.nlist.cell()
    lj = hoomd.md.pair.lj(r_cut=2.5, nlist=nl)
    lj.pair_coeff.set("A", "A", epsilon=1.0, sigma=1.0)
    hoomd.md.integrate.mode_standard(dt=0.005)
    with pytest.raises(RuntimeError):
        hoomd.md.integrate.npt(group=hoomd.group.all(), kT=1.0, tau=0.5, P=1.0)            # no tauP
    with pytest.raises(RuntimeError):
        hoomd.md.integrate.npt(group=hoomd.group.all(), kT=1.0, tau=0.5, P=1.0, tauP=1.0, xy=True)
    hoomd.context.current.integration_methods.clear()
    npt = hoomd.md.integrate.npt(group=hoomd.group.all(), kT=1.0, tau=0.5, P=2.0, tauP=0.7, couple="xy")
    log = hoomd.analyze.log(filename=None, period=20,
                            quantities=["npt_thermostat_energy", "npt_barostat_energy", "volume", "lx", "lz", "pressure"])
    hoomd.run(60)
    log.disable()            # re-configures nothing; then a parameter change that does: the variables must survive
    npt.set_params(tau=0.5)
    hoomd.run(40)
    box = hoomd.context.current.system.box
    assert abs(box.Lx - L0) > 1e-3 and box.Lx == box.Ly and box.Lx != box.Lz
    assert len(log.rows) == 3 and log.rows[0][3] == pytest.approx(L0 ** 3, rel=1e-6)
    # the same job on the reference classes
    pos = np.zeros((N, 4), np.float32)
    pos[:, :3] = p0
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = v0
    vel[:, 3] = 1.0
    r = ref.RefSystem(pos, (L0, L0, L0), vel4=vel)
    r.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
    r.setup_integrator(0.005)
    r.add_npt(kT=1.0, tau=0.5, P=2.0, tauP=0.7, couple="xy")
    r.set_sorter(300)
    r.run(40)
    st40 = r.npt_state()
    r.run(60)
    Lr = r.box_lengths()
    assert np.allclose([box.Lx, box.Ly, box.Lz], Lr, rtol=2e-5)
    assert np.allclose(hoomd.context.current.engine.npt_state()[:10], r.npt_state()[:10], rtol=2e-3, atol=2e-5)
    # row at step 40: {thermostat energy, barostat energy}
    assert np.allclose(log.rows[2][1:3], st40[10:], rtol=3e-3, atol=1e-2)
