"""GPU parity of the bonded path (Kremer-Grest = WCA pair + FENE bonds + bond exclusions) through
the C ABI: against the committed golden vectors, against the reference CPU classes on fresh inputs, and through
size-independent properties at BASELINE's full size (1M beads).

Bars: filtered neighbor lists equal as arrays incl. entry order; bond / pair / net forces, energies and virials within
1e-5 relative (FP32 summation order); all stepping modes of the engine bit-identical to each other; same rebuild counts
as the reference."""
import ctypes as C

import numpy as np
import pytest

import golden_util as G
from hoomd_b200 import workloads

pytestmark = pytest.mark.gpu
RC = 2.0 ** (1.0 / 6.0)


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


def _make(hb, use_graph=2):
    def make(pos, box, ntypes, vel=None, image=None, n_bond_types=0):
        s = hb.System(pos, box, ntypes=ntypes, vel4=vel, image=image, n_bond_types=n_bond_types)
        s.set_tuning(use_graph=use_graph)
        return s
    return make


@pytest.mark.parametrize("use_graph", [0, 1, 2])
def test_kremer_grest_golden(hb, use_graph):
    G.check_kremer_grest(_make(hb, use_graph), rtol=1e-5)


def _melt(cls, kind, excl, steps, **kw):
    pos, vel, L, bonds = workloads.polymer_lattice(10, 20, rho=0.85, kT=1.0, seed=13)
    rng = np.random.default_rng(2)
    types = rng.integers(0, 2, len(bonds)).astype(np.uint32)
    s = cls(pos, (L, L, L), vel4=vel, n_bond_types=2, **kw)
    s.set_bonds(bonds, types)
    lj = dict(r_cut=RC, r_buff=0.4, shift_mode=1)
    if cls.__name__ != "System":  # the CPU reference defaults to half storage; every GPU class uses full lists
        lj["full"] = True
    s.setup_lj(**lj)
    if kind == "fene":
        s.setup_fene(k=[30.0, 25.0], r0=[1.5, 1.6])
    else:
        s.setup_harmonic(k=[300.0, 200.0], r0=[1.0, 1.05])
    s.exclusions_from_bonds(excl)
    s.set_virial(True)
    s.setup_integrator(0.004)
    s.add_langevin(1.0, 3)
    if steps:
        s.run(steps)
    return s, L


def _fscale(f):
    return np.maximum(np.linalg.norm(f[:, :3], axis=1, keepdims=True), np.abs(f[:, :3]).mean())


@pytest.mark.parametrize("kind,excl", [("fene", True), ("harmonic", True), ("fene", False)])
def test_forces_and_list_vs_reference(hb, ref, kind, excl):
    """melt with the reference, then the same state on both: filtered list (exact), bond / pair forces and virials"""
    r, L = _melt(ref.RefSystem, kind, excl, 200)
    st = r.particles()
    g, _ = _melt(hb.System, kind, excl, 0)
    g.set_particles(st["pos"], st["vel"], st["image"])
    r2, _ = _melt(ref.RefSystem, kind, excl, 0)
    r2.set_particles(st["pos"], st["vel"], st["image"])
    r2.compute_forces(0)
    g.compute_forces(0)
    rn, rh, rl = r2.nlist()
    gn, gh, gl = g.nlist()
    assert np.array_equal(rn, gn) and np.array_equal(rh, gh)
    for i in range(len(rn)):
        assert np.array_equal(rl[rh[i]:rh[i] + rn[i]], gl[gh[i]:gh[i] + gn[i]]), i
    from oracle import oracle_c
    o, _ = _melt(oracle_c.OracleSystem, kind, excl, 0)
    o.set_particles(st["pos"], st["vel"], st["image"])
    o.compute_forces(0)
    for which in (0, 1):
        fr, vr = r2.force_component(which, True)
        fg, vg = g.force_component(which, True)
        G.assert_parity(fr, vr, fg, vg, o.force_scales(which + 1), "%s excl=%s %s force" % (kind, excl, ("pair", "bond")[which]))
    # Integrator::computeNetForce: net = pair + bond in FP32, in the order the forces were added
    (fp, vp), (fb, vb) = r2.force_component(0, True), r2.force_component(1, True)
    G.assert_parity(fp + fb, vp + vb, g.forces()[0], g.forces()[1], o.force_scales(3), "net force")
    # Newton's third law over the bond force and over the pair force: the totals vanish
    fb = g.force_component(1)[0][:, :3].astype(np.float64)
    assert np.abs(fb.sum(axis=0)).max() < 1e-3 * np.abs(fb).sum(axis=0).max() / len(fb) ** 0.5


@pytest.mark.parametrize("method", ["nve", "langevin"])
def test_trajectory_vs_reference_and_modes(hb, ref, method):
    """150 steps incl. Nmax overflows, rebuilds and a sorter event: reference rebuild count, host-driven == plain graph ==
    conditional graph bit for bit"""
    r, L = _melt(ref.RefSystem, "fene", True, 100)
    st = r.particles()

    def run(cls, use_graph=None):
        s, _ = _melt(cls, "fene", True, 0)
        s.set_particles(st["pos"], st["vel"], st["image"])
        if use_graph is not None:
            s.set_tuning(use_graph=use_graph)
        s.setup_integrator(0.004)
        s.add_nve() if method == "nve" else s.add_langevin(1.0, 8)
        s.set_sorter(70)
        s.run(150)
        return s

    rr = run(ref.RefSystem)
    pr = rr.particles()
    outs = []
    for mode in (0, 1, 2):
        s = run(hb.System, mode)
        p = s.particles()
        o = np.argsort(p["tag"])
        outs.append({k: p[k][o] for k in ("pos", "vel", "accel", "image")})
        st_g = s.stats()
        # NeighborList::getNumUpdates = normal + forced updates (hoomd/md/NeighborList.h:423-426)
        assert st_g["updates"] + st_g["forced"] == rr.nlist_stats()["updates"], (mode, st_g)
        assert st_g["overflows"] > 0  # the run crosses the Nmax overflow / chunk replay path
    for k in ("pos", "vel", "accel", "image"):
        assert np.array_equal(outs[0][k].view(np.uint32), outs[1][k].view(np.uint32)), k
        assert np.array_equal(outs[0][k].view(np.uint32), outs[2][k].view(np.uint32)), k
    o = np.argsort(pr["tag"])
    assert np.abs(outs[2]["pos"][:, :3] - pr["pos"][o][:, :3]).max() < 5e-4
    assert np.abs(outs[2]["vel"][:, :3] - pr["vel"][o][:, :3]).max() < 5e-3
    assert np.array_equal(outs[2]["image"], pr["image"][o])


def test_forces_follow_the_sorter(hb):
    """tag-keyed bond / exclusion tables: forces by tag are unchanged by a particle sort"""
    g, L = _melt(hb.System, "fene", True, 60)
    p0 = g.particles()
    g.set_particles(p0["pos"], p0["vel"], p0["image"])  # forces a list rebuild at the current positions
    g.compute_forces(g.timestep)
    f0, b0 = g.forces(False)[0], g.force_component(1)[0]
    n0 = g.nlist(counts_only=True)[0]
    g.sort_now()
    g.compute_forces(g.timestep)
    p1 = g.particles()
    assert not np.array_equal(p0["tag"], p1["tag"])
    f1, b1 = g.forces(False)[0], g.force_component(1)[0]
    n1 = g.nlist(counts_only=True)[0]
    o0, o1 = np.argsort(p0["tag"]), np.argsort(p1["tag"])
    assert np.array_equal(n0[o0], n1[o1])
    assert np.array_equal(b0[o0].view(np.uint32), b1[o1].view(np.uint32))  # same bond rows, same order per tag
    assert (np.abs(f0[o0][:, :3] - f1[o1][:, :3]) / _fscale(f0[o0])).max() < 1e-5


def test_bond_out_of_bounds_raises(hb):
    """EvaluatorBondFENE returns false for r >= r0: the reference throws 'bond out of bounds' (PotentialBond.h:288-292)"""
    pos, vel, L, bonds = workloads.polymer_lattice(6, 8, rho=0.85, kT=1.0, seed=1)
    pos[1, 0] += 0.6  # stretches bond (0,1) beyond r0 = 1.5 and squeezes (1,2)
    for use_graph in (0, 2):
        s = hb.System(pos, (L, L, L), vel4=vel)
        s.set_bonds(bonds)
        s.setup_lj(r_cut=RC, r_buff=0.4, shift_mode=1)
        s.setup_fene()
        s.set_tuning(use_graph=use_graph)
        with pytest.raises(RuntimeError, match="out of bounds"):
            s.compute_forces(0)


def test_driver_level_index_keyed_tables(hb):
    """b200_compute_fene_bond_forces / b200_nlist_filter with the reference's INDEX-keyed tables (gpu_table_indexer /
    exli layouts) on raw device buffers == the engine's tag-keyed path"""
    import torch
    from hoomd_b200 import _lib
    Lb = _lib.lib()
    g, L = _melt(hb.System, "fene", False, 40)
    g.compute_forces(g.timestep)
    p = g.particles()
    N = g.N
    pos, vel, Lx, bonds = workloads.polymer_lattice(10, 20, rho=0.85, kT=1.0, seed=13)
    types = np.random.default_rng(2).integers(0, 2, len(bonds)).astype(np.uint32)
    rtag = p["rtag"]
    # index-keyed bond table, pitch N, rows in bond order (BondedGroupData::rebuildGPUTable)
    cnt = np.zeros(N, np.uint32)
    table = np.zeros((2, N, 2), np.uint32)
    for (a, b), t in zip(bonds, types):
        ia, ib = rtag[a], rtag[b]
        table[cnt[ia], ia] = (ib, t); cnt[ia] += 1
        table[cnt[ib], ib] = (ia, t); cnt[ib] += 1
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_pos, d_tab, d_cnt = dev(p["pos"]), dev(table.view(np.int32)), dev(cnt.view(np.int32))
    d_force = torch.zeros((N, 4), dtype=torch.float32, device="cuda")
    d_flags = torch.zeros(1, dtype=torch.int32, device="cuda")
    K, R, lj1, lj2 = g.bond_params
    d_par = dev(np.stack([K, R, lj1, lj2], axis=1).astype(np.float32))
    a = _lib.BondArgs()
    a.d_force, a.d_virial, a.virial_pitch, a.N = d_force.data_ptr(), None, 0, N
    a.d_pos, a.d_diameter, a.box = d_pos.data_ptr(), None, g.box
    a.d_bond_list, a.pitch, a.d_n_bonds = d_tab.data_ptr(), N, d_cnt.data_ptr()
    a.n_bond_types, a.compute_virial, a.d_tag, a.d_rtag = 2, 0, None, None
    assert Lb.b200_compute_fene_bond_forces(C.byref(a), d_par.data_ptr(), d_flags.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert int(d_flags.item()) == 0
    fb = g.force_component(1)[0]
    assert np.array_equal(d_force.cpu().numpy().view(np.uint32), fb.view(np.uint32))

    # filter: index-keyed exclusion table vs a numpy restatement of NeighborList::filterNlist
    n, h, nl = g.nlist()
    ex_n = cnt.copy()
    ex = table[:, :, 0].copy()  # partner indices, exli(idx, k) = k * N + idx
    d_n, d_h, d_nl = dev(n.view(np.int32)), dev(h.view(np.int32)), dev(nl.view(np.int32))
    d_exn, d_ex = dev(ex_n.view(np.int32)), dev(ex.view(np.int32))
    assert Lb.b200_nlist_filter(d_n.data_ptr(), d_nl.data_ptr(), d_h.data_ptr(), d_exn.data_ptr(), d_ex.data_ptr(), N, 2,
                                None, None, N, None) == 0
    torch.cuda.synchronize()
    n2, nl2 = d_n.cpu().numpy().view(np.uint32), d_nl.cpu().numpy().view(np.uint32)
    removed = 0
    for i in range(N):
        row = nl[h[i]:h[i] + n[i]]
        keep = row[~np.isin(row, ex[:ex_n[i], i])]
        removed += len(row) - len(keep)
        assert n2[i] == len(keep) and np.array_equal(nl2[h[i]:h[i] + n2[i]], keep), i
    assert removed == 2 * len(bonds)  # every bonded pair was inside r_list


@pytest.mark.timeout(900)
def test_full_size_kremer_grest_properties(hb):
    """BASELINE configs[3]: 1M beads (10,000 chains x 100), WCA + FENE, NVE.  Size-independent properties: bond and pair
    forces sum to zero (Newton's third law), no bonded pair survives in the list, every row sorted-unique, NVE conserves
    energy, bonds stay inside r0."""
    kg = workloads.KG_PARAMS
    pos, vel, L, bonds = workloads.polymer_lattice(100, 100, rho=kg["rho"], kT=kg["kT"], seed=42)
    s = hb.System(pos, (L, L, L), vel4=vel)
    s.set_bonds(bonds)
    s.setup_lj(r_cut=kg["r_cut"], r_buff=0.4, shift_mode=1)
    s.setup_fene(k=kg["fene_k"], r0=kg["fene_r0"])
    s.exclusions_from_bonds(True)
    s.setup_integrator(0.005)
    s.add_langevin(kg["kT"], 42)
    s.set_sorter(300)
    s.run(1500)
    s.add_nve()
    s.run(100)
    e0 = s.thermo()
    s.run(400)
    e1 = s.thermo()
    etot0, etot1 = e0["ke"] + e0["pe"], e1["ke"] + e1["pe"]
    assert abs(etot1 - etot0) < 2e-3 * abs(e0["ke"]), (e0, e1)
    assert 0.8 < e1["T"] < 1.2
    s.compute_forces(s.timestep)
    for which in (0, 1):
        f = s.force_component(which)[0][:, :3].astype(np.float64)
        assert np.abs(f.sum(axis=0)).max() < 1e-4 * np.abs(f).sum(axis=0).max() / len(f) ** 0.5 * 30, which
    p = s.particles()
    rtag = p["rtag"]
    d = p["pos"][rtag[bonds[:, 1]], :3].astype(np.float64) - p["pos"][rtag[bonds[:, 0]], :3]
    d -= L * np.round(d / L)
    rb = np.linalg.norm(d, axis=1)
    assert rb.max() < 1.5 and 0.9 < rb.mean() < 1.05
    n, h, nl = s.nlist()
    assert 3.0 < n.mean() < 20.0
    # no bonded partner in any row (sampled rows: first 20,000 particles)
    partner = {}
    for a, b in bonds:
        partner.setdefault(int(a), []).append(int(b))
        partner.setdefault(int(b), []).append(int(a))
    tag = p["tag"]
    for i in range(20000):
        row = nl[h[i]:h[i] + n[i]]
        assert len(np.unique(row)) == len(row)
        ex = rtag[partner.get(int(tag[i]), [])]
        assert not np.isin(row, ex).any(), i
