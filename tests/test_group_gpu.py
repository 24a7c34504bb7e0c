"""Integration groups and Langevin options in the ENGINE (the single-kernel MD step, all three stepping modes) against the
UNMODIFIED reference CPU classes: the reference integrates m_group->getIndexArray() (hoomd/md/TwoStepNVEGPU.cc:48-70), so
md.integrate.nve / langevin / nvt over a sparse group move the members only; md.integrate.langevin(dscale=lambda) uses
gamma = lambda * diameter and tally=True keeps the reservoir energy (hoomd/md/TwoStepLangevin.cc:287-318,416-417)."""
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


def _state(hb, n_side=10, ntypes=2):
    pos, vel, L = hb.workloads.sc_lattice(n_side, (1.0 / 0.3) ** (1.0 / 3.0), kT=1.2, seed=9, ntypes=ntypes,
                                          type_fractions=[0.6, 0.4])
    pos[:, :3] += np.random.default_rng(4).normal(0, 0.08, (len(pos), 3)).astype(np.float32)
    return pos, vel, L


def _build(cls, pos, vel, L, members, method, diam=None, **kw):
    is_ref = cls.__name__ == "RefSystem"
    s = cls(pos, (L, L, L), ntypes=2, vel4=vel)
    lj = dict(r_cut=2.5, r_buff=0.4, epsilon=[[1.0, 1.2], [1.2, 0.8]], sigma=[[1.0, 0.9], [0.9, 0.95]])
    s.setup_lj(full=True, **lj) if is_ref else s.setup_lj(**lj)
    if diam is not None:
        s.set_diameters(diam)
    s.setup_integrator(0.005)
    if is_ref:
        s.set_group(members)          # the methods created afterwards act on this group
    if method == "nve":
        s.add_nve()
    elif method == "nvt":
        s.add_nvt(1.1, 0.4)
    else:
        s.add_langevin_ex(1.3, 7, **kw)
    if not is_ref:
        s.set_group(members)
    return s


@pytest.mark.parametrize("method", ["nve", "langevin", "nvt"])
def test_sparse_group_trajectory_vs_reference_all_modes(ref, hb, method):
    pos, vel, L = _state(hb)
    N = len(pos)
    members = np.arange(1, N, 3, dtype=np.uint32)
    mask = np.zeros(N, bool)
    mask[members] = True
    kw = dict(gamma=[1.0, 2.0]) if method == "langevin" else {}
    r = _build(ref.RefSystem, pos, vel, L, members, method, **kw)
    r.set_sorter(20)
    r.run(80)
    a = r.particles()
    oa = np.argsort(a["tag"])
    states = []
    for use_graph in (0, 1, 2, 3):
        g = _build(hb.System, pos, vel, L, members, method, **kw)
        g.set_tuning(use_graph=use_graph)
        g.set_sorter(20)
        g.run(80)
        b = g.particles()
        ob = np.argsort(b["tag"])
        st = {k: b[k][ob] for k in ("pos", "vel", "image")}
        states.append(st)
        # members follow the reference to FP32 round-off growth; non-members have not moved at all
        assert np.abs(st["pos"][mask][:, :3] - a["pos"][oa][mask][:, :3]).max() < 3e-4, (method, use_graph)
        assert np.abs(st["vel"][mask][:, :3] - a["vel"][oa][mask][:, :3]).max() < 3e-3, (method, use_graph)
        assert np.array_equal(st["pos"][~mask].view(np.uint32), pos[~mask].view(np.uint32)), "a non-member moved"
        assert np.array_equal(st["vel"][~mask].view(np.uint32), vel[~mask].view(np.uint32)), "a non-member accelerated"
        assert np.array_equal(st["image"], a["image"][oa])
        # NeighborList::getNumUpdates = normal + forced updates (hoomd/md/NeighborList.h:423-426); the sorter forces some
        gs = g.stats()
        assert r.nlist_stats()["updates"] == gs["updates"] + gs["forced"], (method, use_graph, gs)
        th, tr = g.thermo(), r.thermo()
        assert th["ndof"] == tr["ndof"] == 3 * len(members) - 3
        assert abs(th["T"] - tr["T"]) <= 2e-3 * tr["T"]
        if method == "nvt":
            assert np.allclose(g.nvt_state(), r.nvt_state(), rtol=5e-4, atol=1e-6)
    if method != "nvt":   # NVT sums its kinetic energy in another order per mode (fixed point vs two-pass double)
        for st in states[1:]:
            for k in ("pos", "vel", "image"):
                assert np.array_equal(states[0][k].view(np.uint32), st[k].view(np.uint32)), "stepping modes differ in " + k


@pytest.mark.parametrize("dscale", [None, 0.7])
def test_langevin_dscale_and_tally_vs_reference(ref, hb, dscale):
    pos, vel, L = _state(hb)
    N = len(pos)
    diam = np.random.default_rng(2).uniform(0.8, 1.6, N).astype(np.float32)
    kw = dict(tally=True, dscale=dscale) if dscale is not None else dict(tally=True, gamma=[1.0, 2.0])
    r = _build(ref.RefSystem, pos, vel, L, None, "langevin", diam=diam, **kw)
    nsteps = 60
    r.run(nsteps)
    e_ref = r.langevin_reservoir_energy()
    a = r.particles()
    vals = []
    for use_graph in (0, 2, 3):
        g = _build(hb.System, pos, vel, L, None, "langevin", diam=diam, **kw)
        g.set_tuning(use_graph=use_graph)
        g.run(nsteps // 2)            # two runs: the tally carries over, the last-step term moves with it
        g.run(nsteps - nsteps // 2)
        b = g.particles()
        assert np.array_equal(a["tag"], b["tag"])
        assert np.abs(a["pos"][:, :3] - b["pos"][:, :3]).max() < 3e-4
        e = g.langevin_reservoir_energy()
        vals.append(e)
        # sums of ~1e5 FP32 terms along slightly different trajectories
        assert abs(e - e_ref) <= 2e-3 * abs(e_ref) + 0.05, (e, e_ref, use_graph)
    assert abs(e_ref) > 1.0          # the bath did exchange energy in this run
    assert max(vals) - min(vals) <= 1e-4 * abs(e_ref) + 1e-3


def test_group_surface_in_job_script(ref, hb):
    """hoomd.group.tags / tag_list / type on the job-script surface: md.integrate.nve(group=...) moves the members only"""
    from hoomd_b200 import script as hoomd
    hoomd.install()
    hoomd.context.initialize()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.6, type_name="A"), n=8)
    N = 512
    snap = hoomd.context.current.system.take_snapshot()
    snap.particles.velocity[:] = np.random.default_rng(3).normal(0, 1.0, (N, 3))
    hoomd.context.current.system.restore_snapshot(snap)
    nl = hoomd.md.nlist.cell()
    lj = hoomd.md.pair.lj(r_cut=2.5, nlist=nl)
    lj.pair_coeff.set("A", "A", epsilon=1.0, sigma=1.0)
    hoomd.md.integrate.mode_standard(dt=0.005)
    grp = hoomd.group.tags(100, 299)
    assert len(grp) == 200 and len(hoomd.group.tag_list("x", [1, 5, 5, 9])) == 3 and len(hoomd.group.type("A")) == N
    with pytest.raises(RuntimeError):
        hoomd.group.tag_list("bad", [N])
    hoomd.md.integrate.nve(group=grp)
    p0 = snap.particles.position.copy()
    hoomd.run(40)
    p1 = hoomd.context.current.system.take_snapshot().particles.position
    moved = np.abs(p1 - p0).max(axis=1) > 0
    assert moved[100:300].all() and not moved[:100].any() and not moved[300:].any()
    # the same on the reference's CPU classes
    pos = np.zeros((N, 4), np.float32)
    pos[:, :3] = p0
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = snap.particles.velocity
    vel[:, 3] = 1.0
    L = hoomd.context.current.system.box.Lx
    r = ref.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
    r.setup_integrator(0.005)
    r.set_group(np.arange(100, 300, dtype=np.uint32))
    r.add_nve()
    r.set_sorter(300)
    r.run(40)
    a = r.particles()
    assert np.abs(a["pos"][np.argsort(a["tag"])][:, :3] - p1).max() < 2e-4
