"""Shared checks of an implementation (oracle restatement or CUDA engine) against tests/golden/*.npz, the vectors
generated from the unmodified reference by tests/golden/make_golden.py."""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def flat_nlist(s):
    n, h, nl = s.nlist()
    return n, h, [nl[h[i]:h[i] + n[i]] for i in range(len(n))]


def check_nlist(make, exact_order=True):
    """make(pos, box, ntypes) -> system with setup_lj/compute_nlist/nlist.  Sorted-set equality (north_star bar); with
    exact_order also entry order (CPU reference order)."""
    g = load("nlist_random")
    L = float(g["L"])
    s = make(g["pos"], (L, L, L), 1)
    s.setup_lj(r_cut=3.0, r_buff=0.4)
    s.compute_nlist(0)
    _cmp_nlist(s, g, exact_order)
    g = load("nlist_two_types")
    L = float(g["L"])
    s = make(g["pos"], (L, L, L), 2)
    s.setup_lj(r_cut=g["rcut"], epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]], r_buff=0.3)
    s.compute_nlist(0)
    _cmp_nlist(s, g, exact_order)


def _cmp_nlist(s, g, exact_order):
    n, h, segs = flat_nlist(s)
    assert np.array_equal(n, g["n_neigh"])
    assert np.array_equal(h, g["head"])
    off = np.concatenate([[0], np.cumsum(g["n_neigh"].astype(np.int64))]).astype(np.int64)
    for i, seg in enumerate(segs):
        want = g["nlist"][off[i]:off[i + 1]]
        if exact_order:
            assert np.array_equal(seg, want), i
        else:
            assert np.array_equal(np.sort(seg), np.sort(want)), i



# ---- the north_star bar: "forces, energies and virials must agree within 1e-5 relative in FP32" ----------------------------
# A per-particle force is a sum of ~30 pair terms that largely cancel, so "relative" is taken against what the sum is made
# of: |a_i - b_i| <= 1e-5 * sum_j |term_ij| (force: vector norms |f_ij|; energy: |V_ij / 2|; virial component c: the
# |virial_c term|).  The sums of absolute terms come from the C restatement (oracle.c:orc_get_force_scales), evaluated on the
# same state with the same force field.
PARITY_RTOL = 1e-5


def abs_scales(configure, pos, box, ntypes=1, n_bond_types=0, which=3):
    """configure(system) applies the force field (setup_lj / setup_pair / set_bonds / ...) to an OracleSystem built on
    (pos, box); returns (sum|f_ij| [N], sum|V_ij/2| [N], sum|virial term| [6, N]) in index order."""
    from oracle import oracle_c
    o = oracle_c.OracleSystem(pos, box, ntypes=ntypes, n_bond_types=n_bond_types, threads=os.cpu_count() or 1)
    configure(o)
    o.set_virial(True)
    o.compute_forces(0)
    return o.force_scales(which)


def assert_parity(fr, vr, f, v, scales, what="", rtol=PARITY_RTOL, verbose=True):
    """|f - fr| <= rtol * scale per particle for force vectors, energies and (when given) virial components; totals of the
    energy and of the virial trace within rtol of the sum of absolute terms as well.  Returns the measured maxima."""
    sf, se, sv = scales
    f64 = np.float64
    df = np.linalg.norm(fr[:, :3].astype(f64) - f[:, :3].astype(f64), axis=1)
    de = np.abs(fr[:, 3].astype(f64) - f[:, 3].astype(f64))
    out = {}

    def ratio(d, sc, name):
        assert np.all(d[sc == 0.0] == 0.0), "%s %s: difference where nothing contributes" % (what, name)
        m = sc > 0.0
        r = float((d[m] / sc[m]).max()) if m.any() else 0.0
        out[name] = r
        assert r <= rtol, "%s: %s differs by %.3g of sum|terms| (bar %.1g)" % (what, name, r, rtol)

    ratio(df, sf, "force")
    ratio(de, se, "energy")
    out["energy_total"] = abs(float(fr[:, 3].astype(f64).sum() - f[:, 3].astype(f64).sum())) / max(float(se.sum()), 1e-300)
    assert out["energy_total"] <= rtol, (what, out)
    if v is not None and vr is not None:
        dv = np.abs(vr.astype(f64) - v.astype(f64))
        ratio(dv.reshape(-1), sv.reshape(-1), "virial")
        tr = lambda x: (x[0].astype(f64) + x[3] + x[5]).sum()
        out["virial_total"] = abs(float(tr(vr) - tr(v))) / max(float((sv[0] + sv[3] + sv[5]).sum()), 1e-300)
        assert out["virial_total"] <= rtol, (what, out)
    if verbose:
        print("parity %s: max |d| / sum|terms|: %s (bar %.0e)" % (what, ", ".join("%s %.2e" % kv for kv in out.items()), rtol))
    return out

def check_forces(make, rtol, bit_exact=False):
    g = load("liquid_forces")
    L = float(g["L"])
    for mode, name in ((0, "none"), (1, "shift"), (2, "xplor")):
        s = make(g["pos"], (L, L, L), 1)
        s.setup_lj(r_cut=3.0, r_buff=0.4, shift_mode=mode, r_on=2.0 if mode == 2 else None)
        s.set_virial(True)
        s.compute_forces(0)
        f, v = s.forces()
        fr, vr = g["force_" + name], g["virial_" + name]
        if bit_exact:
            assert np.array_equal(f.view(np.uint32), fr.view(np.uint32)), name
            assert np.array_equal(v.view(np.uint32), vr.view(np.uint32)), name
            continue
        sc = abs_scales(lambda o: o.setup_lj(r_cut=3.0, r_buff=0.4, shift_mode=mode, r_on=2.0 if mode == 2 else None),
                        g["pos"], (L, L, L))
        assert_parity(fr, vr, f, v, sc, "liquid_forces/" + name, rtol)


def check_half_steps(make):
    g = load("half_steps")
    L = float(g["L"])
    eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))
    for limit, tag in ((None, "nolimit"), (0.01, "limit")):
        s = make(g["pos"], (L, L, L), 1, vel=g["vel"])
        s.setup_lj(r_cut=2.0, r_buff=0.3)
        s.setup_integrator(0.005)
        s.add_nve(limit=limit)
        s.set_accel(g["accel"])
        s.step_one(0)
        p = s.particles()
        assert eq(p["pos"], g["s1_pos_" + tag]) and eq(p["vel"], g["s1_vel_" + tag])
        assert np.array_equal(p["image"], g["s1_image_" + tag])
        s.step_two(0, g["fnet"])
        p = s.particles()
        assert eq(p["vel"], g["s2_vel_" + tag]) and eq(p["accel"], g["s2_accel_" + tag])
    s = make(g["pos"], (L, L, L), 1, vel=g["vel"])
    s.setup_lj(r_cut=2.0, r_buff=0.3)
    s.setup_integrator(0.005)
    s.add_langevin(1.2, 4)
    s.step_two(123456, g["fnet"])
    p = s.particles()
    assert eq(p["vel"], g["lang_vel"]) and eq(p["accel"], g["lang_accel"])


def check_trajectories(make, bit_exact, pos_tol=2e-4, vel_tol=2e-3):
    for method in ("nve", "langevin"):
        g = load("traj_" + method)
        L = float(g["L"])
        s = make(g["pos0"], (L, L, L), 1, vel=g["vel0"], image=g["image0"])
        s.setup_lj(r_cut=3.0, r_buff=0.4)
        s.set_virial(True)
        s.setup_integrator(0.005)
        s.add_nve() if method == "nve" else s.add_langevin(1.2, 4)
        s.run(40)
        p = s.particles()
        if bit_exact:
            for k in ("pos", "vel", "accel", "image"):
                assert np.array_equal(p[k].view(np.uint32), g[k].view(np.uint32)), (method, k)
        else:
            assert np.abs(p["pos"][:, :3] - g["pos"][:, :3]).max() < pos_tol, method
            assert np.abs(p["vel"][:, :3] - g["vel"][:, :3]).max() < vel_tol, method
            assert np.array_equal(p["image"], g["image"]), method
        th = s.thermo()
        want = g["thermo"]
        for k, w in zip(("ke", "pe", "T", "P", "ndof"), want):
            assert abs(th[k] - w) <= 2e-4 * abs(w), (method, k, th[k], w)


def _kg_fields(s, g):
    s.set_bonds(g["bonds"])
    s.setup_lj(r_cut=2.0 ** (1.0 / 6.0), r_buff=0.4, shift_mode=1)
    s.setup_fene(k=30.0, r0=1.5)
    s.exclusions_from_bonds(True)


def make_kg(make, g, vel=True, image=False):
    """Kremer-Grest system of the kremer_grest fixture on implementation `make` (WCA + FENE + bond exclusions)."""
    L = float(g["L"])
    s = make(g["pos"], (L, L, L), 1, vel=g["vel"] if vel else None, image=g["image"] if image else None, n_bond_types=1)
    _kg_fields(s, g)
    return s


def _rel(a, b):
    scale = np.maximum(np.abs(b), np.abs(b).mean())
    return (np.abs(a - b) / scale).max()


def check_kremer_grest(make, rtol, bit_exact=False, traj_bit_exact=False, pos_tol=2e-4, vel_tol=2e-3):
    """filtered neighbor list (exact, incl. order), per-ForceCompute forces / virials, net force, 40-step trajectories"""
    g = load("kremer_grest")
    L = float(g["L"])
    s = make_kg(make, g)
    s.set_virial(True)
    s.compute_forces(0)
    _cmp_nlist(s, g, True)
    eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))
    fnet, vnet = s.forces()
    for which, name in ((0, "pair"), (1, "bond")):
        f, v = s.force_component(which, True)
        fr, vr = g["force_" + name], g["virial_" + name]
        if bit_exact:
            assert eq(f[:, :3], fr[:, :3]) and eq(v, vr), name
            # the energy goes through logf: allow the last bit
            assert np.abs(f[:, 3] - fr[:, 3]).max() <= 4e-6 * np.abs(fr[:, 3]).max(), name
            continue
        sc = abs_scales(lambda o: _kg_fields(o, g), g["pos"], (L, L, L), n_bond_types=1, which=which + 1)
        assert_parity(fr, vr, f, v, sc, "kremer_grest/" + name, rtol)
    sc = abs_scales(lambda o: _kg_fields(o, g), g["pos"], (L, L, L), n_bond_types=1)
    assert_parity(g["force_pair"] + g["force_bond"], g["virial_pair"] + g["virial_bond"], fnet, vnet, sc,
                  "kremer_grest/net", max(rtol, 1e-6))
    for method in ("nve", "langevin"):
        s = make_kg(make, g, image=True)
        s.set_virial(True)
        s.setup_integrator(0.005)
        s.add_nve() if method == "nve" else s.add_langevin(1.0, 4)
        s.run(40)
        p = s.particles()
        order = np.argsort(p["tag"]) if "tag" in p else slice(None)
        if traj_bit_exact:
            for k in ("pos", "vel", "accel", "image"):
                assert eq(p[k][order], g["%s_%s" % (method, k)]), (method, k)
        else:
            assert np.abs(p["pos"][order][:, :3] - g[method + "_pos"][:, :3]).max() < pos_tol, method
            assert np.abs(p["vel"][order][:, :3] - g[method + "_vel"][:, :3]).max() < vel_tol, method
            assert np.array_equal(p["image"][order], g[method + "_image"]), method
        th = s.thermo()
        for k, w in zip(("ke", "pe", "T", "P", "ndof"), g[method + "_thermo"]):
            assert abs(th[k] - w) <= 2e-4 * abs(w), (method, k, th[k], w)


PAIR_CASES = {  # as tests/golden/make_golden.py
    "gauss": (dict(epsilon=2.0, sigma=0.9), 2.5, 1.8),
    "yukawa": (dict(epsilon=1.5, kappa=1.2), 2.5, 1.8),
    "mie": (dict(epsilon=1.0, sigma=1.0, n=13.0, m=7.0), 2.5, 1.8),
    "force_shifted_lj": (dict(epsilon=1.0, sigma=1.0, alpha=1.0), 2.5, 1.8),
    "morse": (dict(D0=1.2, alpha=3.0, r0=1.1), 2.5, 1.8),
    "dpd_conservative": (dict(A=25.0), 1.6, 1.2),
}


def check_pair_evaluators(make, rtol, bit_exact=False, kinds=None, pos_tol=2e-4, vel_tol=2e-3):
    """md.pair.gauss / yukawa / mie / force_shifted_lj / morse: forces, energies, virials in the three energy-shift modes
    and a 30-step Langevin trajectory against the reference (fixture pair_evaluators)"""
    g = load("pair_evaluators")
    L = float(g["L"])
    eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))
    for kind, (co, rc, ron) in PAIR_CASES.items():
        if kinds and kind not in kinds:
            continue
        for mode, name in ((0, "none"), (1, "shift"), (2, "xplor")):
            s = make(g["pos"], (L, L, L), 1, vel=g["vel"])
            s.setup_pair(kind, r_cut=rc, r_buff=0.4, shift_mode=mode, r_on=ron if mode == 2 else None, **co)
            s.set_virial(True)
            s.compute_forces(0)
            f, v = s.forces()
            fr, vr = g["%s_%s_force" % (kind, name)], g["%s_%s_virial" % (kind, name)]
            if bit_exact:
                assert eq(f, fr) and eq(v, vr), (kind, name)
                continue
            sc = abs_scales(lambda o: o.setup_pair(kind, r_cut=rc, r_buff=0.4, shift_mode=mode,
                                                   r_on=ron if mode == 2 else None, **co), g["pos"], (L, L, L))
            assert_parity(fr, vr, f, v, sc, "%s/%s" % (kind, name), rtol)
        s = make(g["pos"], (L, L, L), 1, vel=g["vel"], image=g["image"])
        s.setup_pair(kind, r_cut=rc, r_buff=0.4, shift_mode=1, **co)
        s.setup_integrator(0.005)
        s.add_langevin(1.0, 5)
        s.run(30)
        p = s.particles()
        if bit_exact:
            assert eq(p["pos"], g[kind + "_traj_pos"]) and eq(p["vel"], g[kind + "_traj_vel"]), kind
        else:
            assert np.abs(p["pos"][:, :3] - g[kind + "_traj_pos"][:, :3]).max() < pos_tol, kind
            assert np.abs(p["vel"][:, :3] - g[kind + "_traj_vel"][:, :3]).max() < vel_tol, kind


def check_nvt(make, bit_exact, pos_tol=2e-4, vel_tol=2e-3, state_rtol=2e-5, long_run=True):
    """md.integrate.nvt against the reference's TwoStepNVTMTK + ComputeThermo (tests/golden/nvt.npz).  bit_exact: every
    array and the thermostat variables bit for bit (the C restatement sums the kinetic energy in the reference's order);
    otherwise tolerances: a device reduction adds the same doubles in another order, which can move the FP32 temperature
    by one ulp."""
    g = load("nvt")
    L = float(g["L"])
    for virial in (0, 1):
        s = make(g["pos0"], (L, L, L), 1, vel=g["vel0"])
        s.setup_lj(r_cut=2.5, r_buff=0.4)
        s.set_virial(bool(virial))
        s.setup_integrator(0.005)
        s.add_nvt(1.0, 0.5)
        for nsteps, label in ((60, "a"), (340, "b")):
            if label == "b" and not long_run:
                break
            s.run(nsteps)
            k = "v%d%s_" % (virial, label)
            p, st = s.particles(), s.nvt_state()
            if bit_exact:
                for a in ("pos", "vel", "accel", "image"):
                    assert np.array_equal(p[a].view(np.uint32), g[k + a].view(np.uint32)), (k, a)
                assert np.array_equal(st.view(np.uint32), g[k + "state"].view(np.uint32)), (k, st, g[k + "state"])
            else:
                # 400 steps are ~2 Lyapunov times: a last-bit difference has grown, the global thermostat variables less so
                rt = state_rtol if label == "a" else 100 * state_rtol
                assert np.allclose(st, g[k + "state"], rtol=rt, atol=1e-6), (k, st, g[k + "state"])
                if label == "a":  # before chaos amplifies a last-bit difference
                    assert np.abs(p["pos"][:, :3] - g[k + "pos"][:, :3]).max() < pos_tol, k
                    assert np.abs(p["vel"][:, :3] - g[k + "vel"][:, :3]).max() < vel_tol, k
                    assert np.array_equal(p["image"], g[k + "image"]), k
            th = s.thermo()
            for name, w in zip(("ke", "pe", "T", "ndof"), g[k + "thermo"][[0, 1, 2, 4]]):
                assert abs(th[name] - w) <= (1e-6 if bit_exact else 5e-4) * abs(w), (k, name, th[name], w)


BOX_RESIZE_CASES = {"compress": (True, 0.996, 40, 5), "expand_wrap": (False, 1.004, 30, 5)}


def box_resize_schedule(L, factor, segments):
    out, Lc = [], np.float64(L)
    for _ in range(segments):
        Lc = Lc * factor
        out.append((np.float32(Lc), np.float32(Lc * 1.001), np.float32(Lc * 0.999)))
    return out


def check_box_resize(make, bit_exact, pos_tol=3e-4, vel_tol=3e-3, updates=None):
    """hoomd.update.box_resize between run segments against the reference's BoxResizeUpdater + lambda-scaled distance check
    + cell-list re-dimensioning (tests/golden/box_resize.npz).  `updates(s)`: the system's neighbor-list build count."""
    g = load("box_resize")
    L = float(g["L"])
    for name, (scale, factor, segments, nsteps) in BOX_RESIZE_CASES.items():
        s = make(g["pos0"], (L, L, L), 1, vel=g["vel0"])
        s.setup_lj(r_cut=2.5, r_buff=0.4)
        s.setup_integrator(0.005)
        s.add_langevin(1.0, 3)
        for (lx, ly, lz) in box_resize_schedule(L, factor, segments):
            s.set_box(lx, ly, lz, scale)
            s.run(nsteps)
        p = s.particles()
        order = np.argsort(p["tag"]) if "tag" in p else np.arange(len(p["pos"]))
        assert np.array_equal(np.asarray(s.box_lengths(), np.float32).view(np.uint32), g[name + "_box"].view(np.uint32)), name
        if bit_exact:
            for k in ("pos", "vel", "accel", "image"):
                assert np.array_equal(p[k][order].view(np.uint32), g[name + "_" + k].view(np.uint32)), (name, k)
        else:
            Lf = g[name + "_box"]
            d = p["pos"][order][:, :3] - g[name + "_pos"][:, :3]
            d -= Lf * np.round(d / Lf)  # a particle within rounding of a face may sit in the neighboring image
            assert np.abs(d).max() < pos_tol, (name, np.abs(d).max())
            assert np.abs(p["vel"][order][:, :3] - g[name + "_vel"][:, :3]).max() < vel_tol, name
        if updates is not None:
            assert updates(s) == int(g[name + "_updates"]), (name, updates(s), int(g[name + "_updates"]))
