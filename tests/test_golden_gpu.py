"""GPU parity against the committed golden vectors (tests/golden/, generated from the unmodified reference): these do
not need oracle/_ref at run time.  Bars as in test_parity_gpu.py."""
import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hb():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hoomd_b200
    return hoomd_b200


def _make(hb, use_graph=2, tpp=0):
    def make(pos, box, ntypes, vel=None, image=None):
        s = hb.System(pos, box, ntypes=ntypes, vel4=vel, image=image)
        s.set_pair_mode(0)  # golden vectors are index by index in the reference's particle order
        s.set_tuning(pair_tpp=tpp, use_graph=use_graph)
        return s
    return make


def test_nlist_golden_exact_order(hb):
    G.check_nlist(_make(hb), exact_order=True)


@pytest.mark.parametrize("tpp", [0, 1, 8, 32])
def test_forces_golden(hb, tpp):
    G.check_forces(_make(hb, tpp=tpp), rtol=1e-4)


def test_half_steps_golden_bit_exact(hb):
    G.check_half_steps(_make(hb))


@pytest.mark.parametrize("use_graph", [0, 1, 2])
def test_trajectories_golden(hb, use_graph):
    G.check_trajectories(_make(hb, use_graph=use_graph), bit_exact=False)


def test_sfc_order_golden(hb):
    g = G.load("sfc_order")
    L = float(g["L"])
    s = hb.System(g["pos"], (L, L, L))
    s.setup_lj(r_cut=3.0, r_buff=0.4)
    s.sort_now()
    assert np.array_equal(s.particles()["tag"], g["tag"])


def test_errors_are_loud(hb):
    """reference error behaviour: box too small for r_cut (NeighborListBinned.cc:109-115), particle out of the box
    (CellList.cc:589-618), NaN position"""
    pos = hb.workloads.random_system(100, 6.0, seed=1)
    s = hb.System(pos, (6.0,) * 3)
    s.setup_lj(r_cut=3.0, r_buff=0.4)
    with pytest.raises(RuntimeError, match="too small"):
        s.compute_nlist(0)
    pos = hb.workloads.random_system(100, 10.0, seed=1)
    pos[5, 0] = 7.0
    s = hb.System(pos, (10.0,) * 3)
    s.setup_lj(r_cut=2.0, r_buff=0.4)
    with pytest.raises(RuntimeError, match="no longer in the simulation box"):
        s.compute_nlist(0)
    pos[5, 0] = np.nan
    s = hb.System(pos, (10.0,) * 3)
    s.setup_lj(r_cut=2.0, r_buff=0.4)
    with pytest.raises(RuntimeError, match="NaN"):
        s.compute_nlist(0)


def test_empty_and_tiny_systems(hb):
    """N = 0, 1, 2 and a box with fewer than 3 cells per axis (adjacency de-duplication, CellList.cc:283-296)"""
    for N in (0, 1, 2):
        pos = hb.workloads.random_system(N, 8.0, seed=2)
        s = hb.System(pos, (8.0,) * 3)
        s.setup_lj(r_cut=3.0, r_buff=0.4)
        s.compute_forces(0)
        f, _ = s.forces(False)
        assert f.shape == (N, 4) and np.isfinite(f).all()
    pos = np.zeros((2, 4), np.float32)
    pos[1, 0] = 1.1
    s = hb.System(pos, (8.0,) * 3)   # dim = 2 per axis
    s.setup_lj(r_cut=3.0, r_buff=0.4)
    s.compute_forces(0)
    n, h, nl = s.nlist()
    assert list(n) == [1, 1] and nl[h[0]] == 1 and nl[h[1]] == 0
