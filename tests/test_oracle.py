"""CPU tests of the oracle (test infrastructure): the plain-C restatement oracle/oracle.c is pinned
(a) against golden vectors generated from the unmodified reference (tests/golden/, make_golden.py),
(b) against the reference's own known-answer constants (hoomd/md/test/test_lj_force.cc), and
(c) when oracle/_ref/libhoomd_ref.so is present, bit for bit against the reference CPU classes on fresh inputs."""
import numpy as np
import pytest

import golden_util as G
from hoomd_b200 import workloads
from oracle import oracle_c, ref

pytestmark = pytest.mark.skipif(not oracle_c.available(), reason="oracle/_ref/liboracle.so not built (run build())")


def make(pos, box, ntypes, vel=None, image=None, n_bond_types=0):
    return oracle_c.OracleSystem(pos, box, ntypes=ntypes, vel4=vel, image=image, n_bond_types=n_bond_types)


def test_saru_golden():
    g = G.load("saru")
    for t, u, f in zip(g["triples"], g["u32"], g["f"]):
        uu, ff = oracle_c.saru(int(t[0]), int(t[1]), int(t[2]), 4)
        assert np.array_equal(uu, u) and np.array_equal(ff.view(np.uint32), f.view(np.uint32))


def test_nlist_golden():
    G.check_nlist(make, exact_order=True)


def test_forces_golden_bit_exact():
    G.check_forces(make, rtol=0.0, bit_exact=True)


def test_half_steps_golden_bit_exact():
    G.check_half_steps(make)


def test_trajectories_golden_bit_exact():
    G.check_trajectories(make, bit_exact=True)


def test_lj_known_answers():
    """hoomd/md/test/test_lj_force.cc:41-149 (3 particles on the x axis, r_cut 1.3, buffer 3.0)"""
    pos = np.zeros((3, 4), np.float32)
    pos[1, 0] = 2.0 ** (1.0 / 6.0)
    pos[2, 0] = 2.0 * 2.0 ** (1.0 / 6.0)
    tol = 1e-4
    s = make(pos, (1000.0,) * 3, 1)
    s.setup_lj(r_cut=1.3, epsilon=1.15, sigma=1.0, alpha=1.0, r_buff=3.0)
    s.set_virial(True)
    s.compute_forces(0)
    f, v = s.forces()
    assert np.abs(f[:, :3]).max() < 1e-3
    np.testing.assert_allclose(f[:, 3], [-0.575, -1.15, -0.575], rtol=tol)
    s = make(pos, (1000.0,) * 3, 1)
    s.setup_lj(r_cut=1.3, epsilon=1.15, sigma=1.2, alpha=0.45, r_buff=3.0)
    s.set_virial(True)
    s.compute_forces(0)
    f, v = s.forces()
    np.testing.assert_allclose(f[[0, 2], 0], [-93.09822608552962, 93.09822608552962], rtol=tol)
    np.testing.assert_allclose(f[:, 3], [3.5815110377468, 7.1630220754935, 3.581511037746], rtol=tol)
    np.testing.assert_allclose((v[0] + v[3] + v[5]) / 3.0, [17.416537590989, 34.833075181975, 17.416537590989], rtol=tol)


def test_lj_periodic_known_answers():
    """hoomd/md/test/test_lj_force.cc:180-288 (lj_force_periodic_test): 6 particles, 3 types, each pair interacts
    through a different face of the box (20 x 40 x 60)."""
    pos = np.zeros((6, 4), np.float32)
    xyz = [(-9.6, 0, 0), (9.6, 0, 0), (0, -19.6, 0), (0, 19.6, 0), (0, 0, -29.6), (0, 0, 29.6)]
    types = np.array([0, 1, 2, 0, 2, 1], np.int32)
    pos[:, :3] = xyz
    pos[:, 3] = types.view(np.float32)
    s = make(pos, (20.0, 40.0, 60.0), 3)
    eps = np.array([[1.0, 2.0, 3.0], [2.0, 4.0, 5.0], [3.0, 5.0, 6.0]])  # "just change epsilon for the different pairs"
    s.setup_lj(r_cut=1.3, epsilon=eps, sigma=0.5, alpha=0.45, r_buff=3.0)
    s.set_virial(True)
    s.compute_forces(0)
    f, v = s.forces()
    tol = 1e-4
    W = (v[0] + v[3] + v[5]) / 3.0
    np.testing.assert_allclose([f[0, 0], f[1, 0]], [-1.18299976747949, 1.18299976747949], rtol=tol)
    np.testing.assert_allclose([f[2, 1], f[3, 1]], [-1.77449965121923, 1.77449965121923], rtol=tol)
    np.testing.assert_allclose([f[4, 2], f[5, 2]], [-2.95749941869871, 2.95749941869871], rtol=tol)
    np.testing.assert_allclose(W, [-0.15773330233059] * 2 + [-0.23659995349591] * 2 + [-0.39433325582651] * 2, rtol=tol)
    off_axis = f[:, :3].copy()
    off_axis[[0, 1], 0] = off_axis[[2, 3], 1] = off_axis[[4, 5], 2] = 0.0
    assert np.abs(off_axis).max() < 1e-3


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libhoomd_ref.so not built")
@pytest.mark.parametrize("method,ntypes", [("nve", 1), ("langevin", 1), ("langevin", 2)])
def test_restatement_matches_reference_bitwise(method, ntypes):
    """fresh inputs, longer run, including Nmax overflow and many rebuilds"""
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.4) ** (1.0 / 3.0), kT=1.0, seed=77, ntypes=ntypes,
                                       type_fractions=[0.8, 0.2][:ntypes] if ntypes > 1 else None)
    lj = dict(r_cut=2.5, r_buff=0.4)
    if ntypes == 2:
        lj = dict(r_cut=[[2.5, 2.0], [2.0, 2.2]], epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]],
                  r_buff=0.3, shift_mode=1)
    outs = []
    for cls in (ref.RefSystem, oracle_c.OracleSystem):
        s = cls(pos, (L, L, L), ntypes=ntypes, vel4=vel)
        s.setup_lj(full=True, **lj)
        s.setup_integrator(0.005)
        s.add_nve() if method == "nve" else s.add_langevin(1.0, 7, gamma=[1.0, 2.0][:ntypes])
        s.run(150)
        outs.append((s.particles(), s.nlist(), s.nlist_stats()["updates"]))
    (pa, na, ua), (pb, nb, ub) = outs
    for k in ("pos", "vel", "accel", "image"):
        assert np.array_equal(pa[k].view(np.uint32), pb[k].view(np.uint32)), k
    assert ua == ub and ua > 3
    assert np.array_equal(na[0], nb[0]) and np.array_equal(na[1], nb[1])
    for i in range(len(na[0])):
        assert np.array_equal(na[2][na[1][i]:na[1][i] + na[0][i]], nb[2][nb[1][i]:nb[1][i] + nb[0][i]])


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libhoomd_ref.so not built")
def test_cell_list_matches_reference():
    pos = workloads.random_system(2000, 15.0, seed=3)
    outs = []
    for cls in (ref.RefSystem, oracle_c.OracleSystem):
        s = cls(pos, (15.0,) * 3)
        s.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
        s.compute_nlist(0)
        outs.append(s.cell_list())
    a, b = outs
    assert np.array_equal(a["dim"], b["dim"]) and a["Nmax"] == b["Nmax"]
    assert np.array_equal(a["cell_size"], b["cell_size"])
    for c in range(len(a["cell_size"])):
        n = a["cell_size"][c]
        assert np.array_equal(a["xyzf"][c, :n].view(np.uint32), b["xyzf"][c, :n].view(np.uint32))


def test_kremer_grest_golden_bit_exact():
    """WCA + FENE + bond exclusions: filtered list, per-ForceCompute forces, trajectories"""
    G.check_kremer_grest(make, 0.0, bit_exact=True, traj_bit_exact=True)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libhoomd_ref.so not built")
@pytest.mark.parametrize("kind,excl", [("fene", True), ("harmonic", True), ("harmonic", False)])
def test_bonded_restatement_matches_reference_bitwise(kind, excl):
    """fresh inputs: two bond types, branched bonds (up to 3 per bead), exclusions on / off, 120 Langevin steps"""
    pos, vel, L, bonds = workloads.polymer_lattice(8, 16, rho=0.85, kT=1.0, seed=9)
    rng = np.random.default_rng(5)
    types = rng.integers(0, 2, len(bonds)).astype(np.uint32)
    # a few cross links between lattice neighbors of different chains (third bond on some beads)
    extra = np.array([[i, i + 31] for i in range(0, 400, 40)], np.uint32)
    d = pos[extra[:, 1], :3] - pos[extra[:, 0], :3]
    d -= L * np.round(d / L)
    extra = extra[np.linalg.norm(d, axis=1) < 1.2]
    assert len(extra) > 0
    bonds2 = np.concatenate([bonds, extra])
    types2 = np.concatenate([types, np.ones(len(extra), np.uint32)])
    outs = []
    for cls in (ref.RefSystem, oracle_c.OracleSystem):
        s = cls(pos, (L, L, L), vel4=vel, n_bond_types=2)
        s.set_bonds(bonds2, types2)
        s.setup_lj(r_cut=2.0 ** (1.0 / 6.0), r_buff=0.4, full=True, shift_mode=1)
        if kind == "fene":
            s.setup_fene(k=[30.0, 25.0], r0=[1.5, 1.6])
        else:
            s.setup_harmonic(k=[300.0, 200.0], r0=[1.0, 1.05])
        s.exclusions_from_bonds(excl)
        s.set_virial(True)
        s.setup_integrator(0.004)
        s.add_langevin(1.0, 3)
        s.run(120)
        s.compute_forces(s.timestep)
        outs.append((s.particles(), s.nlist(), s.nlist_stats()["updates"], s.force_component(1, True),
                     s.force_component(0, True)))
    (pa, na, ua, ba, fa), (pb, nb, ub, bb, fb) = outs
    for k in ("pos", "vel", "accel", "image"):
        assert np.array_equal(pa[k].view(np.uint32), pb[k].view(np.uint32)), k
    assert ua == ub and ua > 2
    assert np.array_equal(na[0], nb[0]) and np.array_equal(na[1], nb[1])
    for i in range(len(na[0])):
        assert np.array_equal(na[2][na[1][i]:na[1][i] + na[0][i]], nb[2][nb[1][i]:nb[1][i] + nb[0][i]])
    assert np.array_equal(ba[0][:, :3].view(np.uint32), bb[0][:, :3].view(np.uint32))
    assert np.array_equal(ba[1].view(np.uint32), bb[1].view(np.uint32))
    assert np.array_equal(fa[0][:, :3].view(np.uint32), fb[0][:, :3].view(np.uint32))


def test_pair_evaluators_golden_bit_exact():
    """Gauss, Yukawa, Mie, force-shifted LJ, Morse in all energy-shift modes + trajectories"""
    G.check_pair_evaluators(make, 0.0, bit_exact=True)


def test_nvt_golden_bit_exact():
    """TwoStepNVTMTK + ComputeThermo restated: trajectory AND thermostat variables bit for bit (tests/golden/nvt.npz)"""
    G.check_nvt(make, bit_exact=True)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libhoomd_ref.so not built")
def test_nvt_restatement_matches_reference_bitwise():
    """fresh two-type input, another tau / set point, 250 steps incl. Nmax growth"""
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.6) ** (1.0 / 3.0), kT=2.0, seed=41, ntypes=2, type_fractions=[0.8, 0.2])
    outs = []
    for cls in (ref.RefSystem, oracle_c.OracleSystem):
        s = cls(pos, (L, L, L), ntypes=2, vel4=vel)
        s.setup_lj(full=True, r_cut=[[2.5, 2.0], [2.0, 2.2]], epsilon=[[1.0, 1.5], [1.5, 0.5]],
                   sigma=[[1.0, 0.8], [0.8, 0.88]], r_buff=0.3, shift_mode=1)
        s.setup_integrator(0.004)
        s.add_nvt(0.8, 0.25)
        s.run(250)
        outs.append((s.particles(), s.nvt_state(), s.nvt_reservoir_energy()))
    (pa, xa, _), (pb, xb, _) = outs
    for k in ("pos", "vel", "accel", "image"):
        assert np.array_equal(pa[k].view(np.uint32), pb[k].view(np.uint32)), k
    assert np.array_equal(xa.view(np.uint32), xb.view(np.uint32)), (xa, xb)
    assert abs(xa[0]) > 1e-3
    assert outs[0][2] == outs[1][2] and outs[0][2] != 0.0  # log quantity nvt_mtk_reservoir_energy


def test_box_resize_golden_bit_exact():
    """BoxResizeUpdater + lambda-scaled distance check + cell-list re-dimensioning restated (tests/golden/box_resize.npz)"""
    G.check_box_resize(make, bit_exact=True, updates=lambda s: s.nlist_stats()["updates"])


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libhoomd_ref.so not built")
@pytest.mark.parametrize("every", [0, 1, 5])
def test_dist_check_off_matches_reference_bitwise(every):
    """md.nlist.cell(dist_check=False): period 0 rebuilds every step, period 1 still runs the distance check, period 5
    rebuilds when the period is hit (hoomd/md/NeighborList.cc:1126-1135).  (Switching it BETWEEN two runs forces a build:
    the restatement and the engine do it at the current step, in prepRun's force evaluation, the reference one step later
    because its ForceCompute skips a repeated compute(ts) -- so the switch is made before the first run here.)"""
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.4) ** (1.0 / 3.0), kT=1.5, seed=77)
    outs = []
    for cls in (ref.RefSystem, oracle_c.OracleSystem):
        s = cls(pos, (L, L, L), ntypes=1, vel4=vel)
        s.setup_lj(full=True, r_cut=2.5, r_buff=0.4, check_period=every)
        s.setup_integrator(0.005)
        s.add_nve()
        s.set_dist_check(False)
        s.run(80)
        outs.append((s.particles(), s.nlist_stats()["updates"]))
    (pa, ua), (pb, ub) = outs
    for k in ("pos", "vel", "accel", "image"):
        assert np.array_equal(pa[k].view(np.uint32), pb[k].view(np.uint32)), k
    assert ua == ub


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libhoomd_ref.so not built")
@pytest.mark.parametrize("method", ["langevin", "nvt"])
def test_temperature_ramp_matches_reference_bitwise(method):
    """a temperature variant = setT with the value of every step (Scalar T = m_T->getValue(timestep))"""
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.4) ** (1.0 / 3.0), kT=2.0, seed=5)
    outs = []
    for cls in (ref.RefSystem, oracle_c.OracleSystem):
        s = cls(pos, (L, L, L), ntypes=1, vel4=vel)
        s.setup_lj(full=True, r_cut=2.5, r_buff=0.4)
        s.setup_integrator(0.005)
        s.add_langevin(2.0, 11) if method == "langevin" else s.add_nvt(2.0, 0.4)
        for ts in range(60):
            s.set_kT(np.float32(2.0 - 1.2 * ts / 60.0))
            s.run(1)
        outs.append(s.particles())
    for k in ("pos", "vel", "accel", "image"):
        assert np.array_equal(outs[0][k].view(np.uint32), outs[1][k].view(np.uint32)), k
