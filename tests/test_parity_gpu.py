"""GPU parity tests: the CUDA path (through the C ABI) against the UNMODIFIED reference CPU classes
(oracle/_ref/libhoomd_ref.so, HOOMD-blue 2.3.0 SINGLE_PRECISION) on identical snapshots and seeds.

Bars (BASELINE.json north_star): neighbor lists equal as sorted per-particle sets (here: equal arrays, order included),
forces / energies / virials within 1e-5 relative in FP32, integration half steps and the Saru stream bit-exact.
"""
import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu

# north_star: 1e-5 relative in FP32, against the sums of absolute pair terms sum_j |f_ij| / |V_ij/2| / |virial term| that a
# particle's force / energy / virial is made of (golden_util.assert_parity; scales from oracle.c:orc_get_force_scales)
FORCE_RTOL = G.PARITY_RTOL


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="module")
def ref():
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    return R


@pytest.fixture(scope="module")
def hb():
    if not _have_gpu():
        pytest.skip("no CUDA device")
    import hoomd_b200
    return hoomd_b200


def _pair(ref, hb, pos, L, ntypes=1, vel=None, pair_mode=0, **lj):
    """pair_mode 0 = the reference-layout path (same particle order, head list and neighbor order as the reference, so
    the comparisons below are index by index); the cell-tile path is compared by tag in test_tile_gpu.py"""
    box = (L, L, L) if np.isscalar(L) else L
    r = ref.RefSystem(pos, box, ntypes=ntypes, vel4=vel)
    g = hb.System(pos, box, ntypes=ntypes, vel4=vel)
    g.set_pair_mode(pair_mode)
    r.setup_lj(full=True, **lj)
    g.setup_lj(**lj)
    return r, g


def _assert_nlist_equal(r, g, exact_order=True):
    rn, rh, rl = r.nlist()
    gn, gh, gl = g.nlist()
    assert np.array_equal(rn, gn), "n_neigh differs"
    assert np.array_equal(rh, gh), "head list differs"
    for i in range(len(rn)):
        a = rl[rh[i]:rh[i] + rn[i]]
        b = gl[gh[i]:gh[i] + gn[i]]
        if exact_order:
            assert np.array_equal(a, b), "neighbor order of particle %d differs" % i
        else:
            assert np.array_equal(np.sort(a), np.sort(b)), "neighbor set of particle %d differs" % i


def _force_close(fr, fg, scale):
    d = np.abs(fr.astype(np.float64) - fg.astype(np.float64))
    return d.max(), (d / scale).max()


def test_lj_known_answer_3_particles(ref, hb):
    """hoomd/md/test/test_lj_force.cc:41-178 (lj_force_particle_test): F = -+93.09822608552962 etc."""
    pos = np.zeros((3, 4), np.float32)
    pos[1, 0] = 2.0 ** (1.0 / 6.0)
    pos[2, 0] = 2.0 * 2.0 ** (1.0 / 6.0)
    tol = 1e-2 * 1e-2  # the reference's `tol` is 1e-2 PERCENT (hoomd/test/upp11_config.h:19-37)
    # first part: sigma = 1 puts every pair at the LJ minimum => zero force, pe = -eps/2 per bond end
    r, g = _pair(ref, hb, pos, 1000.0, r_cut=1.3, epsilon=1.15, sigma=1.0, alpha=1.0, r_buff=3.0)
    for s in (r, g):
        s.set_virial(True)
        s.compute_forces(0)
    fg, vg = g.forces()
    assert np.abs(fg[:, :3]).max() < 1e-3
    np.testing.assert_allclose(fg[:, 3], [-0.575, -1.15, -0.575], rtol=tol)
    assert np.abs(vg[0] + vg[3] + vg[5]).max() < 1e-3
    # second part: sigma = 1.2, alpha = 0.45
    r, g = _pair(ref, hb, pos, 1000.0, r_cut=1.3, epsilon=1.15, sigma=1.2, alpha=0.45, r_buff=3.0)
    for s in (r, g):
        s.set_virial(True)
        s.compute_forces(0)
    fr, vr = r.forces()
    fg, vg = g.forces()
    np.testing.assert_allclose(fg[0, 0], -93.09822608552962, rtol=tol)
    np.testing.assert_allclose(fg[2, 0], 93.09822608552962, rtol=tol)
    assert np.abs(fg[1, :3]).max() < 1e-3 and np.abs(fg[:, 1:3]).max() < 1e-3
    np.testing.assert_allclose(fg[:, 3], [3.5815110377468, 7.1630220754935, 3.581511037746], rtol=tol)
    np.testing.assert_allclose((vg[0] + vg[3] + vg[5]) / 3.0, [17.416537590989, 34.833075181975, 17.416537590989],
                               rtol=tol)
    np.testing.assert_allclose(fg, fr, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(vg, vr, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("N,L,seed", [(1000, 12.0, 1), (5000, 20.0, 2), (2000, 9.0, 3)])
def test_nlist_and_forces_random(ref, hb, N, L, seed):
    """neighborlist_comparison_test protocol (hoomd/md/test/test_neighborlist.cc:873-937): random particles,
    r_cut 3.0, r_buff 0.4, full mode; plus forces / energy / virial on the same list."""
    pos = hb.workloads.random_system(N, L, seed=seed)
    r, g = _pair(ref, hb, pos, L, r_cut=3.0, r_buff=0.4)
    r.compute_nlist(0)
    g.compute_nlist(0)
    _assert_nlist_equal(r, g)


def test_nlist_two_types_per_pair_rcut(ref, hb):
    """per-type-pair r_cut incl. a disabled pair (r_cut <= 0 skips, hoomd/md/NeighborListBinned.cc:193)."""
    N, L = 3000, 14.0
    pos = hb.workloads.random_system(N, L, seed=7, ntypes=2)
    rc = [[2.5, 2.0], [2.0, -1.0]]
    eps = [[1.0, 1.5], [1.5, 0.5]]
    sig = [[1.0, 0.8], [0.8, 0.88]]
    r, g = _pair(ref, hb, pos, L, ntypes=2, r_cut=rc, epsilon=eps, sigma=sig, r_buff=0.3)
    r.compute_nlist(0)
    g.compute_nlist(0)
    _assert_nlist_equal(r, g)


def _liquid(hb, n_side=12, rho=0.2, kT=1.2, seed=5):
    pos, vel, L = hb.workloads.sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=kT, seed=seed)
    return pos, vel, L


def test_forces_liquid_1e5(ref, hb):
    """Equilibrated-ish LJ liquid: forces, per-particle energies and virials within 1e-5 relative (FP32)."""
    pos, vel, L = _liquid(hb)
    lj = dict(r_cut=3.0, r_buff=0.4)
    r, g = _pair(ref, hb, pos, L, vel=vel, **lj)
    # let the reference melt the lattice, then hand the SAME state to both
    r.setup_integrator(0.005)
    r.add_langevin(1.2, 42)
    r.run(300)
    st = r.particles()
    r2, g = _pair(ref, hb, st["pos"], L, vel=st["vel"], **lj)
    for s in (r2, g):
        s.set_virial(True)
        s.compute_forces(0)
    _assert_nlist_equal(r2, g)
    fr, vr = r2.forces()
    sc = G.abs_scales(lambda o: o.setup_lj(**lj), st["pos"], (L, L, L))
    for tpp in (1, 2, 4, 8, 16, 32):
        g.set_tuning(pair_tpp=tpp)
        g.compute_forces(0)
        fg, vg = g.forces()
        G.assert_parity(fr, vr, fg, vg, sc, "liquid tpp %d" % tpp, FORCE_RTOL)


@pytest.mark.parametrize("mode", [1, 2])
def test_shift_modes(ref, hb, mode):
    """energy shift / xplor (hoomd/md/test/test_lj_force.cc:369-533)"""
    pos, vel, L = _liquid(hb, n_side=8)
    pos[:, :3] += np.random.default_rng(3).normal(0, 0.15, (len(pos), 3)).astype(np.float32)
    lj = dict(r_cut=3.0, r_buff=0.4, shift_mode=mode, r_on=2.0 if mode == 2 else None)
    r, g = _pair(ref, hb, pos, L, **lj)
    for s in (r, g):
        s.compute_forces(0)
    fr, _ = r.forces(False)
    fg, _ = g.forces(False)
    G.assert_parity(fr, None, fg, None, G.abs_scales(lambda o: o.setup_lj(**lj), pos, (L, L, L)), "shift mode %d" % mode,
                    FORCE_RTOL)


def test_nve_half_steps_bit_exact(ref, hb):
    """TwoStepNVE step one / step two on identical inputs are bit-identical, including wrap + image flags
    (hoomd/md/test/test_nve_integrator.cc:47-252)."""
    N, L = 4096, 10.0
    rng = np.random.default_rng(11)
    pos = hb.workloads.random_system(N, L, seed=4)
    pos[:64, 0] = np.float32(L / 2 - 1e-4)   # about to cross +x
    pos[64:128, 1] = np.float32(-L / 2 + 1e-5)  # about to cross -y
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = rng.normal(0, 1.5, (N, 3))
    vel[:, 3] = rng.uniform(0.5, 2.0, N)
    accel = rng.normal(0, 10.0, (N, 3)).astype(np.float32)
    fnet = rng.normal(0, 20.0, (N, 4)).astype(np.float32)
    for limit in (None, 0.01):
        r, g = _pair(ref, hb, pos, L, vel=vel, r_cut=2.0, r_buff=0.3)
        for s in (r, g):
            s.setup_integrator(0.005)
            s.add_nve(limit=limit)
            s.set_accel(accel)
            s.step_one(0)
        a, b = r.particles(), g.particles()
        for k in ("pos", "vel", "image"):
            assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), (k, limit)
        assert np.abs(b["image"]).sum() > 0
        r.step_two(0, fnet)
        g.step_two(0, fnet)
        a, b = r.particles(), g.particles()
        for k in ("vel", "accel"):
            assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), (k, limit)


def test_langevin_step_two_bit_exact(ref, hb):
    """TwoStepLangevin::integrateStepTwo incl. the Saru(tag, timestep, seed) stream is bit-identical."""
    N, L = 4096, 10.0
    rng = np.random.default_rng(12)
    pos = hb.workloads.random_system(N, L, seed=5, ntypes=2)
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = rng.normal(0, 1.0, (N, 3))
    vel[:, 3] = rng.uniform(0.5, 2.0, N)
    fnet = rng.normal(0, 20.0, (N, 4)).astype(np.float32)
    r, g = _pair(ref, hb, pos, L, ntypes=2, vel=vel, r_cut=2.0, r_buff=0.3)
    for ts in (0, 17, 123456):
        for s in (r, g):
            s.setup_integrator(0.005)
            s.add_langevin(1.2, 4, gamma=[1.0, 2.5])
            s.set_particles(vel4=vel)
            s.step_two(ts, fnet)
        a, b = r.particles(), g.particles()
        for k in ("vel", "accel"):
            assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), (k, ts)


def test_saru_stream(ref, hb):
    """hoomd/Saru.h on a grid of (tag, timestep, seed) triples: u32 and s(-1,1) bit-exact (device vs reference header)."""
    import ctypes as C
    import torch
    from hoomd_b200 import _lib
    rng = np.random.default_rng(0)
    triples = np.concatenate([rng.integers(0, 2 ** 32, (500, 3), dtype=np.uint64).astype(np.uint32),
                              np.array([[0, 0, 0], [1, 2, 3], [0xffffffff, 0xffffffff, 0xffffffff], [5, 0, 7]],
                                       np.uint32)])
    nd = 4
    d_seeds = torch.from_numpy(triples.astype(np.int64)).to(torch.int32).cuda().contiguous()
    d_u = torch.zeros(len(triples) * nd, dtype=torch.int32, device="cuda")
    d_f = torch.zeros(len(triples) * nd, dtype=torch.float32, device="cuda")
    rc = _lib.lib().b200_saru_draw(d_u.data_ptr(), d_f.data_ptr(), d_seeds.data_ptr(), len(triples), nd, None)
    assert rc == 0
    torch.cuda.synchronize()
    gu = d_u.cpu().numpy().view(np.uint32).reshape(-1, nd)
    gf = d_f.cpu().numpy().reshape(-1, nd)
    for k, (a, b, c) in enumerate(triples):
        ru, rf = ref.saru(int(a), int(b), int(c), nd)
        assert np.array_equal(ru, gu[k])
        assert np.array_equal(rf.view(np.uint32), gf[k].view(np.uint32))


def _run_pair(ref, hb, nsteps, method, use_graph, n_side=10, sorter=0, virial=False):
    pos, vel, L = _liquid(hb, n_side=n_side)
    lj = dict(r_cut=3.0, r_buff=0.4)
    r, g = _pair(ref, hb, pos, L, vel=vel, **lj)
    for s in (r, g):
        s.set_virial(virial)
        s.setup_integrator(0.005)
        if method == "nve":
            s.add_nve()
        else:
            s.add_langevin(1.2, 42)
        if sorter:
            s.set_sorter(sorter)
    g.set_tuning(pair_tpp=1, use_graph=use_graph)
    r.run(nsteps)
    g.run(nsteps)
    return r, g


@pytest.mark.parametrize("method", ["nve", "langevin"])
@pytest.mark.parametrize("use_graph", [0, 1, 2])
def test_short_trajectory_vs_reference(ref, hb, method, use_graph):
    """50 steps from a lattice: same rebuild count, trajectories agree to FP32 round-off growth."""
    r, g = _run_pair(ref, hb, 50, method, use_graph)
    a, b = r.particles(), g.particles()
    assert np.array_equal(a["tag"], b["tag"])
    assert np.abs(a["pos"][:, :3] - b["pos"][:, :3]).max() < 2e-4
    assert np.abs(a["vel"][:, :3] - b["vel"][:, :3]).max() < 2e-3
    assert np.array_equal(a["image"], b["image"])
    assert r.nlist_stats()["updates"] == g.stats()["updates"]


@pytest.mark.parametrize("pair_mode", [0, 1])
@pytest.mark.parametrize("method", ["nve", "langevin"])
def test_graph_modes_bit_identical_to_host_driven(ref, hb, method, pair_mode):
    """device-driven (plain graph, conditional graph) == host-driven control flow, bit for bit, over rebuilds,
    Nmax overflows (the lattice start overflows Nmax=8 at once) and sorter events."""
    outs = []
    for ug in (0, 1, 2):
        pos, vel, L = _liquid(hb, n_side=12)
        g = hb.System(pos, (L, L, L), vel4=vel)
        g.set_pair_mode(pair_mode)
        g.setup_lj(r_cut=3.0, r_buff=0.4)
        assert g.pair_mode()["tile"] == bool(pair_mode)
        g.setup_integrator(0.005)
        g.add_nve() if method == "nve" else g.add_langevin(1.2, 7)
        g.set_sorter(100)
        g.set_tuning(pair_tpp=4, use_graph=ug)
        g.run(450)
        p = g.particles()
        outs.append((p, g.stats()))
    for p, st in outs[1:]:
        for k in ("pos", "vel", "accel", "image", "tag"):
            assert np.array_equal(outs[0][0][k].view(np.uint32), p[k].view(np.uint32)), k
        assert st["updates"] == outs[0][1]["updates"]
    assert outs[0][1]["updates"] > 5


def test_sorter_order_matches_reference(ref, hb):
    """SFCPackUpdater: the Hilbert order computed per particle on the device equals the reference's table-based order."""
    pos, vel, L = _liquid(hb, n_side=10)
    pos[:, :3] += np.random.default_rng(9).normal(0, 0.3, (len(pos), 3)).astype(np.float32)
    pos[:, :3] = np.clip(pos[:, :3], -L / 2, L / 2 - 1e-3)
    r, g = _pair(ref, hb, pos, L, vel=vel, r_cut=3.0, r_buff=0.4)
    r.sort_now()
    g.sort_now()
    a, b = r.particles(), g.particles()
    assert np.array_equal(a["tag"], b["tag"])
    assert np.array_equal(a["rtag"], b["rtag"])
    assert np.array_equal(a["pos"].view(np.uint32), b["pos"].view(np.uint32))


def test_thermo(ref, hb):
    r, g = _run_pair(ref, hb, 20, "nve", 2, virial=True)
    tr, tg = r.thermo(), g.thermo()
    assert abs(tr["ke"] - tg["ke"]) < 1e-4 * abs(tr["ke"])
    assert abs(tr["pe"] - tg["pe"]) < 1e-4 * abs(tr["pe"])
    assert abs(tr["T"] - tg["T"]) < 1e-4 * abs(tr["T"])
    assert abs(tr["P"] - tg["P"]) < 2e-4 * abs(tr["P"])
    assert tr["ndof"] == tg["ndof"]


def test_nve_energy_drift_matches_reference(ref, hb):
    """BASELINE north_star: NVE energy drift over 1e5 steps must match the reference's.  1000-particle LJ liquid
    (rho 0.2, r_cut 3.0 unshifted, r_buff 0.4, dt 0.005), equilibrated by the reference, then 1e5 NVE steps on both;
    total energy per particle sampled every 2000 steps.  Trajectories decorrelate (chaos), so the comparison is between
    the drift statistics: end-to-end drift and RMS fluctuation of E/N agree, and both stay small."""
    import os
    nsteps = int(os.environ.get("HOOMD_B200_DRIFT_STEPS", "100000"))
    pos, vel, L = _liquid(hb, n_side=10)
    lj = dict(r_cut=3.0, r_buff=0.4)
    r0 = ref.RefSystem(pos, (L, L, L), vel4=vel)
    r0.setup_lj(full=True, **lj)
    r0.setup_integrator(0.005)
    r0.add_langevin(1.2, 3)
    r0.run(3000)
    st = r0.particles()
    order = np.argsort(st["tag"])
    p0, v0, i0 = st["pos"][order], st["vel"][order], st["image"][order]

    def trace(s):
        s.setup_lj(**lj) if not isinstance(s, ref.RefSystem) else s.setup_lj(full=False, **lj)
        s.setup_integrator(0.005)
        s.add_nve()
        s.set_sorter(300)
        e = []
        for _ in range(nsteps // 2000):
            s.run(2000)
            th = s.thermo()
            e.append((th["ke"] + th["pe"]) / len(p0))
        return np.array(e)

    eg = trace(hb.System(p0, (L, L, L), vel4=v0, image=i0))
    er = trace(ref.RefSystem(p0, (L, L, L), vel4=v0, image=i0))
    scale = abs(er.mean())
    drift_g, drift_r = (eg[-1] - eg[0]) / scale, (er[-1] - er[0]) / scale
    slope_g = np.polyfit(np.arange(len(eg)), eg, 1)[0] * len(eg) / scale
    slope_r = np.polyfit(np.arange(len(er)), er, 1)[0] * len(er) / scale
    print("E/N ref %.5f gpu %.5f | drift ref %.2e gpu %.2e | fitted ref %.2e gpu %.2e | rms ref %.2e gpu %.2e" % (
        er.mean(), eg.mean(), drift_r, drift_g, slope_r, slope_g, er.std() / scale, eg.std() / scale))
    assert abs(eg.mean() - er.mean()) < 3e-3 * scale   # same energy shell (samples fluctuate by the cutoff jumps)
    assert abs(slope_g) < 5e-3 and abs(slope_r) < 5e-3  # both conserve energy to FP32 + cutoff-jump level
    assert abs(slope_g - slope_r) < 5e-3
    assert abs(eg.std() - er.std()) < 3e-3 * scale


def _ka_system(hb, n_side, seed=42):
    """Kob-Andersen 80:20 (BASELINE config 3): rho 1.2, types by permutation, per-pair eps/sigma/r_cut, shifted LJ"""
    rho = 1.2
    pos, vel, L = hb.workloads.sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=1.0, seed=seed, ntypes=2,
                                          type_fractions=[0.8, 0.2])
    return pos, vel, L


def test_kob_andersen_small_vs_reference(ref, hb):
    """two-type parameter table, per-pair r_cut, mode=shift, per-type gamma Langevin: nlist exact, forces 1e-5,
    60-step trajectory with the same rebuild count (BASELINE config 3 at N = 1728)"""
    pos, vel, L = _ka_system(hb, 12)
    pos[:, :3] += np.random.default_rng(1).normal(0, 0.05, (len(pos), 3)).astype(np.float32)
    lj = dict(shift_mode=1, r_buff=0.3, **hb.workloads.KA_PARAMS)
    r, g = _pair(ref, hb, pos, L, ntypes=2, vel=vel, **lj)
    for s in (r, g):
        s.set_virial(True)
        s.compute_forces(0)
    _assert_nlist_equal(r, g)
    fr, vr = r.forces()
    fg, vg = g.forces()
    G.assert_parity(fr, vr, fg, vg, G.abs_scales(lambda o: o.setup_lj(**lj), pos, (L, L, L), ntypes=2), "Kob-Andersen 1728",
                    FORCE_RTOL)
    for s in (r, g):
        s.setup_integrator(0.002)
        s.add_langevin(1.0, 7, gamma=[1.0, 1.5])
        s.run(60)
    a, b = r.particles(), g.particles()
    assert np.array_equal(a["tag"], b["tag"])
    assert np.abs(a["pos"][:, :3] - b["pos"][:, :3]).max() < 5e-4
    assert r.nlist_stats()["updates"] == g.stats()["updates"]


def _full_size_properties(g, N, ntypes):
    """size-independent properties of a full list + its forces: the list is symmetric (j in N(i) <=> i in N(j)),
    entries are in range and unique per row, and Newton's third law makes the net force vanish"""
    g.compute_forces(0)
    n, head, nl = g.nlist()
    assert n.sum() % 2 == 0
    rows = np.repeat(np.arange(N, dtype=np.int64), n)
    idx = (head.astype(np.int64)[:, None] + np.arange(n.max(), dtype=np.int64)[None, :])
    mask = np.arange(n.max())[None, :] < n[:, None]
    cols = nl[idx[mask]].astype(np.int64)
    assert cols.min() >= 0 and cols.max() < N and (cols != rows).all()
    fwd = np.sort(rows * N + cols)
    bwd = np.sort(cols * N + rows)
    assert np.array_equal(fwd, bwd), "neighbor list is not symmetric"
    assert len(np.unique(fwd)) == len(fwd), "duplicate neighbor entries"
    f = g.forces(False)[0].astype(np.float64)
    fsum = np.abs(f[:, :3].sum(axis=0)).max()
    assert fsum < 1e-3 * np.abs(f[:, :3]).sum() / N * np.sqrt(N), fsum


def test_full_size_lj64k_properties(hb):
    """BASELINE config 2 at full size (N = 64,000)"""
    pos, vel, L = hb.workloads.lj_liquid(64000)
    pos[:, :3] += np.random.default_rng(2).normal(0, 0.08, (len(pos), 3)).astype(np.float32)
    pos[:, :3] = np.clip(pos[:, :3], -L / 2, np.nextafter(np.float32(L / 2), np.float32(0)))
    g = hb.System(pos, (L, L, L), vel4=vel)
    g.setup_lj(r_cut=3.0, r_buff=0.4)
    _full_size_properties(g, len(pos), 1)
    g.setup_integrator(0.005)
    g.add_nve()
    g.set_sorter(300)
    p0 = (g.particles()["vel"][:, :3].astype(np.float64)).sum(axis=0)
    g.run(400)
    p1 = (g.particles()["vel"][:, :3].astype(np.float64)).sum(axis=0)
    assert np.abs(p1 - p0).max() < 0.05, (p0, p1)  # total momentum conserved (mass 1)


def test_full_size_ka256k_properties(hb):
    """BASELINE config 3 at full size (N = 256,000: fcc 40^3 start), Langevin kT = 1.0"""
    pos, vel, L = hb.workloads.fcc_lattice(40, 1.2, kT=1.0, seed=42, ntypes=2, type_fractions=[0.8, 0.2])
    pos[:, :3] += np.random.default_rng(3).normal(0, 0.03, (len(pos), 3)).astype(np.float32)
    g = hb.System(pos, (L, L, L), ntypes=2, vel4=vel)
    g.setup_lj(shift_mode=1, r_buff=0.3, **hb.workloads.KA_PARAMS)
    _full_size_properties(g, len(pos), 2)
    g.setup_integrator(0.002)
    g.add_langevin(1.0, 7)
    g.set_sorter(300)
    g.run(300)
    th = g.thermo()
    # the sc start at rho 1.2 is far from equilibrium: the melt heats up before the thermostat (gamma 1) pulls it back
    assert 0.5 < th["T"] < 8.0 and np.isfinite(th["pe"]), th
    assert g.stats()["updates"] > 0


@pytest.mark.parametrize("n_side,method", [(24, "nve"), (30, "langevin"), (38, "nve"), (40, "langevin")])
def test_multistep_kernel_bit_identical_to_one_step_kernel(hb, n_side, method):
    """k_md_steps (several MD steps per launch, grid-wide barrier instead of a kernel boundary; 8 / 4 / 2 lanes per particle
    at these sizes) against the one-step kernel in the same conditional graph: identical bits after 260 steps with list
    rebuilds, a particle sort and two run() calls, equal rebuild counts; the switch and the stepping query work."""
    pos, vel, L = hb.workloads.sc_lattice(n_side, (1.0 / 0.2) ** (1.0 / 3.0), kT=1.2, seed=21)
    pos[:, :3] += np.random.default_rng(3).normal(0, 0.1, (len(pos), 3)).astype(np.float32)
    out = []
    for multi in (True, False):
        s = hb.System(pos, (L, L, L), vel4=vel)
        s.setup_lj(r_cut=3.0, r_buff=0.4)
        s.setup_integrator(0.005)
        if method == "nve":
            s.add_nve()
        else:
            s.add_langevin(1.2, 5)
        s.set_sorter(100)
        s.set_multistep(multi)
        s.run(130)
        s.run(130)
        st = s.stepping()
        assert st["multistep"] == multi and st["use_graph"] == 2, st
        assert (st["steps_per_unit"] == 0) == multi
        p = s.particles()
        out.append((p, s.stats(), s.forces(False)[0]))
        s.close()
    (pa, sa, fa), (pb, sb, fb) = out
    for k in ("pos", "vel", "image", "tag", "accel"):
        assert np.array_equal(pa[k].view(np.uint32), pb[k].view(np.uint32)), k
    assert np.array_equal(fa.view(np.uint32), fb.view(np.uint32))
    assert (sa["updates"], sa["forced"]) == (sb["updates"], sb["forced"]) and sa["updates"] >= 5
    assert sa["kernels_last_run"] < sb["kernels_last_run"]    # fewer launches is the point
