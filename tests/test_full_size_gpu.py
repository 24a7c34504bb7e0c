"""Full-size parity: every single-GPU BASELINE config at the size BASELINE.json names, CUDA path (through the C ABI) against
the UNMODIFIED reference CPU classes (oracle/_ref) on the same state -- ONE reference evaluation each (seconds):

  * configs[1]  LJ liquid 64,000 (the equilibrated bench snapshot tests/golden/lj64k_equil.npz, made by the reference),
  * configs[2]  Kob-Andersen 80:20, 256,000 particles (fcc 40^3 start, melted on the GPU engine first),
  * configs[3]  Kremer-Grest 1,000,000 beads (WCA + FENE + bond exclusions, melted on the GPU engine first).

Protocol of hoomd/md/test/test_neighborlist.cc:873-937 and test_lj_force.cc:232-286 at full size: neighbor lists equal as
ARRAYS (row lengths, head list and every entry, order included); forces, energies and virials within 1e-5 of the sums of
absolute pair / bond terms (golden_util.assert_parity).
"""
import os

import numpy as np
import pytest

import golden_util as G

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


def assert_nlist_arrays_equal(r, g, chunk=65536):
    """row lengths, head list and every entry (order included) equal; vectorised in chunks of rows"""
    rn, rh, rl = r.nlist()
    gn, gh, gl = g.nlist()
    assert np.array_equal(rn, gn), "n_neigh differs"
    assert np.array_equal(rh, gh), "head list differs"
    for a in range(0, len(rn), chunk):
        n, h = rn[a:a + chunk].astype(np.int64), rh[a:a + chunk].astype(np.int64)
        k = np.arange(int(n.max()) if len(n) else 0, dtype=np.int64)
        flat = (h[:, None] + k[None, :])[k[None, :] < n[:, None]]
        assert np.array_equal(rl[flat], gl[flat]), "neighbor entries differ in rows %d..%d" % (a, a + chunk)
    return int(rn.astype(np.int64).sum())


def test_lj64k_bench_snapshot_vs_reference(ref, hb):
    """BASELINE configs[1] on the equilibrated snapshot both bench arms start from"""
    g = G.load("lj64k_equil")
    L = float(g["L"])
    pos, vel, image = g["pos"], g["vel"], g["image"].astype(np.int32)
    assert len(pos) == 64000
    lj = dict(r_cut=3.0, r_buff=0.4)
    r = ref.RefSystem(pos, (L, L, L), vel4=vel, image=image)
    s = hb.System(pos, (L, L, L), vel4=vel, image=image)
    r.setup_lj(full=True, **lj)
    s.setup_lj(**lj)
    for q in (r, s):
        q.set_virial(True)
        q.setup_integrator(0.005)
        q.add_nve()
        q.compute_forces(0)
    entries = assert_nlist_arrays_equal(r, s)
    assert np.array_equal(r.nlist()[0].astype(np.uint16), g["n_neigh"])   # the fixture's own record of that state
    fr, vr = r.forces()
    fg, vg = s.forces()
    sc = G.abs_scales(lambda o: o.setup_lj(**lj), pos, (L, L, L))
    out = G.assert_parity(fr, vr, fg, vg, sc, "LJ 64,000 (bench snapshot)")
    r.run(0)                                      # prepRun: Integrator::computeNetForce, which ComputeThermo sums
    tr, tg = r.thermo(), s.thermo()
    for k in ("ke", "pe", "T", "P"):
        assert abs(tr[k] - tg[k]) <= 1e-5 * max(abs(tr[k]), 1e-30), (k, tr[k], tg[k])
    print("64k: %d list entries equal, %s" % (entries, out))


def test_ka256k_vs_reference(ref, hb):
    """BASELINE configs[2]: 256,000 particles, two types, per-pair r_cut, shifted LJ"""
    pos, vel, L = hb.workloads.fcc_lattice(40, 1.2, kT=1.0, seed=42, ntypes=2, type_fractions=[0.8, 0.2])
    assert len(pos) == 256000
    lj = dict(shift_mode=1, r_buff=0.4, **hb.workloads.KA_PARAMS)
    m = hb.System(pos, (L, L, L), ntypes=2, vel4=vel)
    m.setup_lj(**lj)
    m.setup_integrator(0.002)
    m.add_langevin(1.0, 7)
    m.set_sorter(300)
    m.run(600)                                    # melt the lattice (untimed, GPU engine)
    p = m.particles()
    m.close()
    o = np.argsort(p["tag"], kind="stable")
    pos, vel, image = p["pos"][o], p["vel"][o], p["image"][o]
    r = ref.RefSystem(pos, (L, L, L), ntypes=2, vel4=vel, image=image)
    s = hb.System(pos, (L, L, L), ntypes=2, vel4=vel, image=image)
    r.setup_lj(full=True, **lj)
    s.setup_lj(**lj)
    for q in (r, s):
        q.set_virial(True)
        q.compute_forces(0)
    entries = assert_nlist_arrays_equal(r, s)
    fr, vr = r.forces()
    fg, vg = s.forces()
    sc = G.abs_scales(lambda q: q.setup_lj(**lj), pos, (L, L, L), ntypes=2)
    out = G.assert_parity(fr, vr, fg, vg, sc, "Kob-Andersen 256,000")
    print("KA 256k: %d list entries equal, %s" % (entries, out))


def test_kremer_grest_1m_vs_reference(ref, hb):
    """BASELINE configs[3]: 10,000 chains x 100 beads, WCA + FENE, bond exclusions: filtered list + pair / bond / net forces"""
    pos, vel, L, bonds = hb.workloads.polymer_lattice(100, 100, rho=0.85, kT=1.0, seed=42)
    assert len(pos) == 1000000

    def fields(q, **kw):
        q.set_bonds(bonds)
        q.setup_lj(r_cut=2.0 ** (1.0 / 6.0), r_buff=0.4, shift_mode=1, **kw)
        q.setup_fene(k=30.0, r0=1.5)
        q.exclusions_from_bonds(True)

    m = hb.System(pos, (L, L, L), vel4=vel, n_bond_types=1)
    fields(m)
    m.setup_integrator(0.005)
    m.add_langevin(1.0, 42)
    m.set_sorter(300)
    m.run(400)                                    # melt (untimed, GPU engine)
    p = m.particles()
    m.close()
    o = np.argsort(p["tag"], kind="stable")
    pos, vel, image = p["pos"][o], p["vel"][o], p["image"][o]
    r = ref.RefSystem(pos, (L, L, L), vel4=vel, image=image, n_bond_types=1)
    s = hb.System(pos, (L, L, L), vel4=vel, image=image, n_bond_types=1)
    fields(r, full=True)
    fields(s)
    for q in (r, s):
        q.set_virial(True)
        q.compute_forces(0)
    entries = assert_nlist_arrays_equal(r, s)
    from oracle import oracle_c
    oc = oracle_c.OracleSystem(pos, (L, L, L), n_bond_types=1, threads=os.cpu_count() or 1)
    fields(oc)
    oc.set_virial(True)
    oc.compute_forces(0)
    out = {}
    for which, name in ((0, "pair"), (1, "bond")):
        fr, vr = r.force_component(which, True)
        fg, vg = s.force_component(which, True)
        out[name] = G.assert_parity(fr, vr, fg, vg, oc.force_scales(which + 1), "Kremer-Grest 1M " + name)
    # Integrator::computeNetForce: net = pair + bond in FP32, in the order the forces were added
    (fp, vp), (fb, vb) = r.force_component(0, True), r.force_component(1, True)
    fg, vg = s.forces()
    out["net"] = G.assert_parity(fp + fb, vp + vb, fg, vg, oc.force_scales(3), "Kremer-Grest 1M net")
    print("KG 1M: %d list entries equal, %s" % (entries, out))
