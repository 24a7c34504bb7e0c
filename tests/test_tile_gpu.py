"""GPU parity tests of the cell-tile pair path (hoomd_b200/csrc/tile.cu, pair_mode 1) against the UNMODIFIED reference
CPU classes (oracle/_ref).  The tile path keeps the particles in cell order, so every comparison goes through the tags:
same neighbor SETS per particle (BASELINE north_star bar), forces / energies / virials within 1e-5 relative in FP32,
same rebuild decisions, trajectories within FP32 round-off growth.
"""
import numpy as np
import pytest

import golden_util as G
from test_parity_gpu import FORCE_RTOL, _ka_system, _liquid, hb, ref  # noqa: F401 (fixtures)

pytestmark = pytest.mark.gpu


def _pair(ref, hb, pos, L, ntypes=1, vel=None, tile=(0, 0, 0), tpp=0, **lj):
    box = (L, L, L)
    r = ref.RefSystem(pos, box, ntypes=ntypes, vel4=vel)
    g = hb.System(pos, box, ntypes=ntypes, vel4=vel)
    g.set_pair_mode(1, tile=tile, tpp=tpp)
    r.setup_lj(full=True, **lj)
    g.setup_lj(**lj)
    assert g.pair_mode()["tile"]
    return r, g


def _melted(ref, hb, n_side=12, steps=300):
    pos, vel, L = _liquid(hb, n_side=n_side)
    r = ref.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_lj(full=True, r_cut=3.0, r_buff=0.4)
    r.setup_integrator(0.005)
    r.add_langevin(1.2, 42)
    r.run(steps)
    st = r.particles()
    o = np.argsort(st["tag"])
    return st["pos"][o], st["vel"][o], L


def _assert_forces(r, g, L, lj, ntypes=1, virial=True, what="tile"):
    """by tag, against the sums of absolute pair terms (north_star: 1e-5 relative, golden_util.assert_parity)"""
    fr, vr = r.forces(virial)
    fg, vg = g.forces(virial)
    pr = r.particles()
    tg, tr = g.particles()["tag"], pr["tag"]
    fr2 = np.empty_like(fr); fr2[tr] = fr
    fg2 = np.empty_like(fg); fg2[tg] = fg
    vr2 = vg2 = None
    if virial:
        vr2 = np.empty_like(vr); vr2[:, tr] = vr
        vg2 = np.empty_like(vg); vg2[:, tg] = vg
    pos_by_tag = np.empty_like(pr["pos"]); pos_by_tag[tr] = pr["pos"]
    sc = G.abs_scales(lambda o: o.setup_lj(**lj), pos_by_tag, (L, L, L), ntypes=ntypes)
    return G.assert_parity(fr2, vr2, fg2, vg2, sc, what, FORCE_RTOL)


def _assert_nlist_sets(r, g):
    rn, rh, rl = r.nlist()
    gn, gh, gl = g.nlist()
    tr, tg = r.particles()["tag"], g.particles()["tag"]
    ref_sets = {}
    for i in range(len(rn)):
        ref_sets[int(tr[i])] = np.sort(tr[rl[rh[i]:rh[i] + rn[i]]])
    for i in range(len(gn)):
        mine = np.sort(tg[gl[gh[i]:gh[i] + gn[i]]])
        assert np.array_equal(mine, ref_sets[int(tg[i])]), "neighbor set of tag %d differs" % tg[i]


@pytest.mark.parametrize("tile", [(0, 0, 0), (1, 1, 1), (2, 2, 2), (4, 2, 2), (3, 1, 2), (1, 4, 4)])
def test_tile_nlist_sets_and_forces(ref, hb, tile, tpp=0):
    pos, vel, L = _melted(ref, hb)
    r, g = _pair(ref, hb, pos, L, vel=vel, tile=tile, tpp=tpp, r_cut=3.0, r_buff=0.4)
    for s in (r, g):
        s.set_virial(True)
        s.compute_forces(0)
    pm = g.pair_mode()
    if tile[0]:
        assert pm["cells_per_tile"] == tile
    _assert_nlist_sets(r, g)
    _assert_forces(r, g, L, dict(r_cut=3.0, r_buff=0.4))
    # the particles are stored in cell order now: positions are a permutation of the input, tags say which
    p = g.particles()
    assert np.array_equal(p["pos"][np.argsort(p["tag"])].view(np.uint32), pos.view(np.uint32))
    assert np.array_equal(p["rtag"][p["tag"]], np.arange(len(pos)))


@pytest.mark.parametrize("mode", [1, 2])
def test_tile_shift_modes(ref, hb, mode):
    pos, vel, L = _liquid(hb, n_side=8)
    pos[:, :3] += np.random.default_rng(3).normal(0, 0.15, (len(pos), 3)).astype(np.float32)
    lj = dict(r_cut=3.0, r_buff=0.4, shift_mode=mode, r_on=2.0 if mode == 2 else None)
    r, g = _pair(ref, hb, pos, L, **lj)
    for s in (r, g):
        s.compute_forces(0)
    _assert_forces(r, g, L, lj, virial=False, what="tile shift mode %d" % mode)


def _compare_state(r, g, ptol, vtol):
    a, b = r.particles(), g.particles()
    oa, ob = np.argsort(a["tag"]), np.argsort(b["tag"])
    assert np.abs(a["pos"][oa, :3] - b["pos"][ob, :3]).max() < ptol
    assert np.abs(a["vel"][oa, :3] - b["vel"][ob, :3]).max() < vtol
    assert np.array_equal(a["image"][oa], b["image"][ob])
    assert np.array_equal(a["pos"][oa, 3].view(np.uint32), b["pos"][ob, 3].view(np.uint32))  # types travel with the tag


@pytest.mark.parametrize("method", ["nve", "langevin"])
@pytest.mark.parametrize("use_graph", [0, 1, 2])
def test_tile_short_trajectory_vs_reference(ref, hb, method, use_graph):
    """50 steps from a lattice (several rebuilds, each one a re-ordering): same rebuild count as the reference,
    trajectories agree to FP32 round-off growth"""
    pos, vel, L = _liquid(hb, n_side=10)
    r, g = _pair(ref, hb, pos, L, vel=vel, r_cut=3.0, r_buff=0.4)
    for s in (r, g):
        s.setup_integrator(0.005)
        s.add_nve() if method == "nve" else s.add_langevin(1.2, 42)
    g.set_tuning(use_graph=use_graph)
    r.run(50)
    g.run(50)
    _compare_state(r, g, 2e-4, 2e-3)
    assert r.nlist_stats()["updates"] == g.stats()["updates"]
    assert g.stats()["updates"] > 0


def test_tile_kob_andersen_vs_reference(ref, hb):
    """two types, per-pair r_cut / eps / sigma, mode=shift, per-type gamma: sets, forces, 60-step trajectory"""
    pos, vel, L = _ka_system(hb, 12)
    pos[:, :3] += np.random.default_rng(1).normal(0, 0.05, (len(pos), 3)).astype(np.float32)
    lj = dict(shift_mode=1, r_buff=0.3, **hb.workloads.KA_PARAMS)
    r, g = _pair(ref, hb, pos, L, ntypes=2, vel=vel, **lj)
    for s in (r, g):
        s.set_virial(True)
        s.compute_forces(0)
    _assert_nlist_sets(r, g)
    _assert_forces(r, g, L, lj, ntypes=2, what="tile Kob-Andersen")
    for s in (r, g):
        s.setup_integrator(0.002)
        s.add_langevin(1.0, 7, gamma=[1.0, 1.5])
        s.run(60)
    _compare_state(r, g, 5e-4, 5e-3)
    assert r.nlist_stats()["updates"] == g.stats()["updates"]


def test_tile_row_overflow_grows(ref, hb, monkeypatch):
    """a row capacity that is too small is grown by the host protocol (NeighborList::checkConditions) -- in the
    host-driven build and, from the checkpoint, when a device-driven rebuild overflows inside a chunk"""
    monkeypatch.setenv("HOOMD_B200_TILE_NMAX0", "8")
    pos, vel, L = _liquid(hb, n_side=10)
    r, g = _pair(ref, hb, pos, L, vel=vel, r_cut=3.0, r_buff=0.4)
    assert g.pair_mode()["row_capacity"] == 8
    for s in (r, g):
        s.setup_integrator(0.005)
        s.add_nve()
    r.run(120)
    g.run(120)
    assert g.pair_mode()["row_capacity"] > 32
    assert g.stats()["overflows"] >= 2
    _compare_state(r, g, 1e-3, 1e-2)
    assert r.nlist_stats()["updates"] == g.stats()["updates"]


def test_tile_clustered_density(ref, hb):
    """very uneven cell populations (a dense slab in a dilute box): staging capacity and row capacity both grow"""
    rng = np.random.default_rng(12)
    L = 24.0
    n_dense, n_gas = 3000, 600
    dense = rng.uniform([-L / 2, -L / 2, -3.0], [L / 2, L / 2, 3.0], (n_dense, 3))
    gas = rng.uniform(-L / 2, L / 2, (n_gas, 3))
    pos = np.zeros((n_dense + n_gas, 4), np.float32)
    pos[:, :3] = np.concatenate([dense, gas])
    # soft enough that random overlaps do not overflow FP32: epsilon small
    lj = dict(r_cut=2.5, r_buff=0.4, epsilon=1e-3)
    r, g = _pair(ref, hb, pos, L, **lj)
    for s in (r, g):
        s.compute_forces(0)
    _assert_nlist_sets(r, g)
    _assert_forces(r, g, L, lj, virial=False, what="tile clustered")  # overlapping pairs: huge cancelling terms


def test_tile_mode_refuses_small_boxes(hb):
    """fewer than 3 cells per axis: the default stays on the reference-layout path; forcing tiles is a loud error"""
    pos = hb.workloads.random_system(200, 8.0, seed=2)
    s = hb.System(pos, (8.0,) * 3)
    s.setup_lj(r_cut=3.0, r_buff=0.4)  # dim = 2 per axis
    assert not s.pair_mode()["tile"]
    s.compute_forces(0)
    with pytest.raises(RuntimeError, match="cell-tile path needs"):
        s.set_pair_mode(1)
