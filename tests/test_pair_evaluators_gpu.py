"""GPU parity of the other PotentialPair<evaluator> instantiations (Gauss, Yukawa, Mie, force-shifted LJ,
Morse) through the C ABI: against the golden vectors generated from the reference, against the reference CPU classes on a
fresh two-type system, the three stepping modes against each other, and the driver-level entry point."""
import ctypes as C

import numpy as np
import pytest

import golden_util as G
from hoomd_b200 import workloads

pytestmark = pytest.mark.gpu
KINDS = list(G.PAIR_CASES)


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


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("tpp", [0, 8])
def test_golden(hb, kind, tpp):
    def make(pos, box, ntypes, vel=None, image=None):
        s = hb.System(pos, box, ntypes=ntypes, vel4=vel, image=image)
        s.set_tuning(pair_tpp=tpp)
        return s
    G.check_pair_evaluators(make, rtol=1e-5, kinds=[kind])


TWO_TYPE = {  # per type pair (AA, AB, BB) coefficients
    "gauss": dict(epsilon=[[2.0, 1.0], [1.0, 0.5]], sigma=[[0.9, 1.0], [1.0, 1.1]]),
    "yukawa": dict(epsilon=[[1.5, 1.0], [1.0, 2.0]], kappa=[[1.2, 1.0], [1.0, 0.8]]),
    "mie": dict(epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]], n=[[13.0, 12.0], [12.0, 14.0]],
                m=[[7.0, 6.0], [6.0, 7.0]]),
    "force_shifted_lj": dict(epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]]),
    "morse": dict(D0=[[1.2, 1.0], [1.0, 0.7]], alpha=[[3.0, 2.5], [2.5, 2.0]], r0=[[1.1, 1.0], [1.0, 1.2]]),
    "dpd_conservative": dict(A=[[25.0, 40.0], [40.0, 15.0]]),
}


@pytest.mark.parametrize("kind", KINDS)
def test_two_types_vs_reference_and_modes(hb, ref, kind):
    """80:20 mixture, per-pair r_cut: forces vs the reference, then 120 Langevin steps in the three stepping modes
    (bit-identical to each other) vs the reference trajectory"""
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.6) ** (1.0 / 3.0), kT=1.0, seed=31, ntypes=2, type_fractions=[0.8, 0.2])
    pos[:, :3] += np.random.default_rng(4).normal(0, 0.08, (len(pos), 3)).astype(np.float32)
    rc = [[2.5, 2.0], [2.0, 2.2]]
    co = TWO_TYPE[kind]

    def build(cls, **kw):
        s = cls(pos, (L, L, L), ntypes=2, vel4=vel)
        s.setup_pair(kind, r_cut=rc, r_buff=0.3, shift_mode=1, **co, **kw)
        s.set_virial(True)
        return s

    r, g = build(ref.RefSystem, full=True), build(hb.System)
    r.compute_forces(0)
    g.compute_forces(0)
    fr, vr = r.forces()
    fg, vg = g.forces()
    sc = G.abs_scales(lambda o: o.setup_pair(kind, r_cut=rc, r_buff=0.3, shift_mode=1, **co), pos, (L, L, L), ntypes=2)
    G.assert_parity(fr, vr, fg, vg, sc, "two-type " + kind)
    r.setup_integrator(0.004)
    r.add_langevin(1.0, 9, gamma=[1.0, 2.0])
    r.run(120)
    pr = r.particles()
    outs = []
    for mode in (0, 1, 2):
        s = build(hb.System)
        s.set_tuning(use_graph=mode)
        s.setup_integrator(0.004)
        s.add_langevin(1.0, 9, gamma=[1.0, 2.0])
        s.run(120)
        p = s.particles()
        o = np.argsort(p["tag"])
        outs.append({k: p[k][o] for k in ("pos", "vel", "image")})
        st = s.stats()
        assert st["updates"] + st["forced"] == r.nlist_stats()["updates"], (mode, st)
    for k in ("pos", "vel", "image"):
        assert np.array_equal(outs[0][k].view(np.uint32), outs[1][k].view(np.uint32)), k
        assert np.array_equal(outs[0][k].view(np.uint32), outs[2][k].view(np.uint32)), k
    assert np.abs(outs[2]["pos"][:, :3] - pr["pos"][:, :3]).max() < 5e-4
    assert np.abs(outs[2]["vel"][:, :3] - pr["vel"][:, :3]).max() < 5e-3


def test_driver_entry_point_matches_engine(hb):
    """b200_compute_pair_forces on raw device buffers (the reference's param_type layout) == the engine's force"""
    import torch
    from hoomd_b200 import _lib
    Lb = _lib.lib()
    g = G.load("pair_evaluators")
    L = float(g["L"])
    for kind in ("yukawa", "mie"):
        co, rc, _ = G.PAIR_CASES[kind]
        s = hb.System(g["pos"], (L, L, L), vel4=g["vel"])
        s.setup_pair(kind, r_cut=rc, r_buff=0.4, shift_mode=1, **co)
        s.compute_forces(0)
        f_eng = s.forces(False)[0]
        n, h, nl = s.nlist()
        N = s.N
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        d_pos, d_n, d_h, d_nl = dev(g["pos"]), dev(n.view(np.int32)), dev(h.view(np.int32)), dev(nl.view(np.int32))
        prm = s.pair_params[0, 0]
        d_prm = dev(prm[:2].copy() if kind == "yukawa" else prm.copy())  # Scalar2 vs Scalar4 param_type
        d_rcsq = dev(np.array([rc * rc], np.float32))
        d_f = torch.zeros((N, 4), dtype=torch.float32, device="cuda")
        a = _lib.PairArgs()
        a.d_force, a.d_virial, a.virial_pitch, a.N, a.d_pos, a.box = d_f.data_ptr(), None, 0, N, d_pos.data_ptr(), s.box
        a.d_n_neigh, a.d_nlist, a.d_head_list = d_n.data_ptr(), d_nl.data_ptr(), d_h.data_ptr()
        a.d_rcutsq, a.d_ronsq, a.ntypes, a.shift_mode, a.compute_virial = d_rcsq.data_ptr(), None, 1, 1, 0
        a.threads_per_particle, a.block_size, a.wrap_skip_margin = 0, 0, 0.0
        assert Lb.b200_compute_pair_forces(_lib.PAIR_KINDS[kind], C.byref(a), d_prm.data_ptr(), None) == 0
        torch.cuda.synchronize()
        fd = d_f.cpu().numpy()
        fscale = np.maximum(np.linalg.norm(f_eng[:, :3], axis=1, keepdims=True), np.abs(f_eng[:, :3]).mean())
        assert (np.abs(fd[:, :3] - f_eng[:, :3]) / fscale).max() < 1e-6, kind
    assert Lb.b200_compute_pair_forces(99, C.byref(a), d_prm.data_ptr(), None) != 0


def test_pair_script_surface(hb, ref):
    """md.pair.yukawa through the job-script surface against the reference"""
    from hoomd_b200 import script as hoomd
    hoomd.install()
    hoomd.context.initialize()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.3, type_name="A"), n=8)
    nl = hoomd.md.nlist.cell()
    yk = hoomd.md.pair.yukawa(r_cut=3.0, nlist=nl)
    yk.pair_coeff.set("A", "A", epsilon=2.0, kappa=1.5)
    hoomd.md.integrate.mode_standard(dt=0.005)
    hoomd.md.integrate.langevin(group=hoomd.group.all(), kT=1.0, seed=4)
    hoomd.run(50)
    snap = hoomd.context.current.system.take_snapshot()
    pos, vel, L = workloads.sc_lattice(8, 1.3)
    r = ref.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_pair("yukawa", r_cut=3.0, r_buff=0.4, full=True, epsilon=2.0, kappa=1.5)
    r.setup_integrator(0.005)
    r.add_langevin(1.0, 4)
    r.set_sorter(300)
    r.run(50)
    a = r.particles()
    o = np.argsort(a["tag"])
    assert np.abs(a["pos"][o][:, :3] - snap.particles.position).max() < 2e-4
