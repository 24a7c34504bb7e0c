"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libhoomd_ref.so, built from /root/reference by
oracle/Makefile).  Run in the build container only:  python tests/golden/make_golden.py
The fixtures pin the C restatement (oracle/oracle.c) and the CUDA path on machines where the reference does not exist.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ref  # noqa: E402
from hoomd_b200 import workloads  # noqa: E402


def csr(s):
    n, h, nl = s.nlist()
    return n, h, np.concatenate([nl[h[i]:h[i] + n[i]] for i in range(len(n))]) if len(n) else nl[:0]


def kremer_grest():
    """(8) Kremer-Grest bead-spring chains (WCA pair + FENE bonds, bond exclusions): 216 beads in chains of 8, melted with
    the reference; filtered neighbor list, per-ForceCompute forces / virials, 40-step NVE and Langevin trajectories"""
    kg = workloads.KG_PARAMS
    pos, vel, L, bonds = workloads.polymer_lattice(6, 8, rho=kg["rho"], kT=kg["kT"], seed=3)

    def make(p, v, img=None):
        q = ref.RefSystem(p, (L, L, L), vel4=v, image=img, n_bond_types=1)
        q.set_bonds(bonds)
        q.setup_lj(r_cut=kg["r_cut"], r_buff=0.4, full=True, shift_mode=1)
        q.setup_fene(k=kg["fene_k"], r0=kg["fene_r0"])
        q.exclusions_from_bonds(True)
        return q

    s = make(pos, vel)
    s.setup_integrator(0.005)
    s.add_langevin(kg["kT"], 11)
    s.run(400)
    st = s.particles()
    q = make(st["pos"], st["vel"])
    q.set_virial(True)
    q.compute_forces(0)
    n, h, flat = csr(q)
    fp, vp = q.force_component(0, True)
    fb, vb = q.force_component(1, True)
    res = dict(pos=st["pos"], vel=st["vel"], image=st["image"], L=L, bonds=bonds, n_neigh=n, head=h, nlist=flat,
               force_pair=fp, virial_pair=vp, force_bond=fb, virial_bond=vb)
    for method in ("nve", "langevin"):
        q = make(st["pos"], st["vel"], st["image"])
        q.set_virial(True)
        q.setup_integrator(0.005)
        q.add_nve() if method == "nve" else q.add_langevin(kg["kT"], 4)
        q.run(40)
        p = q.particles()
        th = q.thermo()
        for k in ("pos", "vel", "accel", "image"):
            res["%s_%s" % (method, k)] = p[k]
        res[method + "_updates"] = q.nlist_stats()["updates"]
        res[method + "_thermo"] = np.array([th[k] for k in ("ke", "pe", "T", "P", "ndof")])
    np.savez_compressed(os.path.join(HERE, "kremer_grest.npz"), **res)


PAIR_CASES = {  # evaluator: (coefficients, r_cut, r_on for xplor)
    "gauss": (dict(epsilon=2.0, sigma=0.9), 2.5, 1.8),
    "yukawa": (dict(epsilon=1.5, kappa=1.2), 2.5, 1.8),
    "mie": (dict(epsilon=1.0, sigma=1.0, n=13.0, m=7.0), 2.5, 1.8),
    "force_shifted_lj": (dict(epsilon=1.0, sigma=1.0, alpha=1.0), 2.5, 1.8),
    "morse": (dict(D0=1.2, alpha=3.0, r0=1.1), 2.5, 1.8),
    "dpd_conservative": (dict(A=25.0), 1.6, 1.2),
}


def pair_evaluators():
    """(9) the other PotentialPair<evaluator> instantiations on a melted 216-particle liquid: force / energy / virial for
    every evaluator x energy-shift mode, and a 30-step Langevin trajectory per evaluator"""
    pos, vel, L = workloads.sc_lattice(6, (1.0 / 0.5) ** (1.0 / 3.0), kT=1.0, seed=8)
    s = ref.RefSystem(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
    s.setup_integrator(0.005)
    s.add_langevin(1.0, 3)
    s.run(300)
    st = s.particles()
    res = dict(pos=st["pos"], vel=st["vel"], image=st["image"], L=L)
    for kind, (co, rc, ron) in PAIR_CASES.items():
        for mode, name in ((0, "none"), (1, "shift"), (2, "xplor")):
            q = ref.RefSystem(st["pos"], (L, L, L), vel4=st["vel"])
            q.setup_pair(kind, r_cut=rc, r_buff=0.4, full=True, shift_mode=mode, r_on=ron if mode == 2 else None, **co)
            q.set_virial(True)
            q.compute_forces(0)
            f, v = q.forces()
            res["%s_%s_force" % (kind, name)], res["%s_%s_virial" % (kind, name)] = f, v
        q = ref.RefSystem(st["pos"], (L, L, L), vel4=st["vel"], image=st["image"])
        q.setup_pair(kind, r_cut=rc, r_buff=0.4, full=True, shift_mode=1, **co)
        q.setup_integrator(0.005)
        q.add_langevin(1.0, 5)
        q.run(30)
        p = q.particles()
        res[kind + "_traj_pos"], res[kind + "_traj_vel"] = p["pos"], p["vel"]
    np.savez_compressed(os.path.join(HERE, "pair_evaluators.npz"), **res)


def nvt():
    """(10) Nose-Hoover (MTK) thermostat: an LJ liquid started at kT 1.5 and coupled to kT 1.0 with tau 0.5 -- state and
    thermostat variables after 60 and 400 steps (Nmax growth and several rebuilds on the way), virial flag off and on
    (the flag switches ComputeThermo to the pressure-tensor form of the kinetic energy)"""
    pos, vel, L = workloads.sc_lattice(8, (1.0 / 0.4) ** (1.0 / 3.0), kT=1.5, seed=21)
    res = dict(pos0=pos, vel0=vel, L=L)
    for virial in (0, 1):
        q = ref.RefSystem(pos, (L, L, L), vel4=vel)
        q.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
        q.set_virial(bool(virial))
        q.setup_integrator(0.005)
        q.add_nvt(1.0, 0.5)
        for nsteps, label in ((60, "a"), (340, "b")):
            q.run(nsteps)
            p, th = q.particles(), q.thermo()
            k = "v%d%s_" % (virial, label)
            res[k + "pos"], res[k + "vel"], res[k + "accel"], res[k + "image"] = p["pos"], p["vel"], p["accel"], p["image"]
            res[k + "state"] = q.nvt_state()
            res[k + "thermo"] = np.array([th[x] for x in ("ke", "pe", "T", "P", "ndof")])
    np.savez_compressed(os.path.join(HERE, "nvt.npz"), **res)


BOX_RESIZE_CASES = {  # name: (scale_particles, per-segment factor, segments, steps per segment)
    "compress": (True, 0.996, 40, 5),
    "expand_wrap": (False, 1.004, 30, 5),
}


def box_resize_schedule(L, factor, segments):
    """the box lengths of segment k (FP32, slightly anisotropic so that the three lambdas differ)"""
    out, Lc = [], np.float64(L)
    for _ in range(segments):
        Lc = Lc * factor
        out.append((np.float32(Lc), np.float32(Lc * 1.001), np.float32(Lc * 0.999)))
    return out


def box_resize():
    """(11) hoomd.update.box_resize: BoxResizeUpdater between run segments -- particle scaling / wrapping, the lambda-scaled
    distance check (NeighborList.cc:981-1007) and the cell list re-dimensioning (4 -> 3 cells per axis on the way down)"""
    pos, vel, L = workloads.sc_lattice(9, (1.0 / 0.3) ** (1.0 / 3.0), kT=1.2, seed=7)
    res = dict(pos0=pos, vel0=vel, L=L)
    for name, (scale, factor, segments, nsteps) in BOX_RESIZE_CASES.items():
        q = ref.RefSystem(pos, (L, L, L), vel4=vel)
        q.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
        q.setup_integrator(0.005)
        q.add_langevin(1.0, 3)
        for (lx, ly, lz) in box_resize_schedule(L, factor, segments):
            q.set_box(lx, ly, lz, scale)
            q.run(nsteps)
        p = q.particles()
        order = np.argsort(p["tag"])
        for k in ("pos", "vel", "accel", "image"):
            res[name + "_" + k] = p[k][order]
        res[name + "_box"] = q.box_lengths()
        res[name + "_updates"] = q.nlist_stats()["updates"]
    np.savez_compressed(os.path.join(HERE, "box_resize.npz"), **res)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "box":
        box_resize()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "nvt":
        nvt()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "kg":
        kremer_grest()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "pairs":
        pair_evaluators()
        return
    kremer_grest()
    pair_evaluators()
    nvt()
    box_resize()
    out = {}
    # (1) Saru stream: (tag, timestep, seed) triples
    rng = np.random.default_rng(0)
    triples = np.concatenate([rng.integers(0, 2 ** 32, (64, 3), dtype=np.uint64).astype(np.uint32),
                              np.array([[0, 0, 0], [1, 2, 3], [0xffffffff] * 3, [5, 0, 7]], np.uint32)])
    u = np.zeros((len(triples), 4), np.uint32)
    f = np.zeros((len(triples), 4), np.float32)
    for k, t in enumerate(triples):
        u[k], f[k] = ref.saru(int(t[0]), int(t[1]), int(t[2]), 4)
    np.savez_compressed(os.path.join(HERE, "saru.npz"), triples=triples, u32=u, f=f)

    # (2) neighborlist_comparison_test protocol: random particles, r_cut 3.0, r_buff 0.4, full mode
    N, L = 600, 11.0
    pos = workloads.random_system(N, L, seed=21)
    s = ref.RefSystem(pos, (L, L, L))
    s.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
    s.compute_nlist(0)
    n, h, flat = csr(s)
    cl = s.cell_list()
    np.savez_compressed(os.path.join(HERE, "nlist_random.npz"), pos=pos, L=L, n_neigh=n, head=h, nlist=flat,
                        cell_dim=cl["dim"], cell_size=cl["cell_size"])

    # (3) two types, per-pair r_cut with one disabled pair
    pos2 = workloads.random_system(500, 10.0, seed=22, ntypes=2)
    rc = np.array([[2.5, 2.0], [2.0, -1.0]])
    s = ref.RefSystem(pos2, (10.0,) * 3, ntypes=2)
    s.setup_lj(r_cut=rc, epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]], r_buff=0.3, full=True)
    s.compute_nlist(0)
    n, h, flat = csr(s)
    np.savez_compressed(os.path.join(HERE, "nlist_two_types.npz"), pos=pos2, L=10.0, rcut=rc, n_neigh=n, head=h,
                        nlist=flat)

    # (4) LJ liquid: melt a 512-particle lattice with the reference, then forces / energy / virial and trajectories
    pos, vel, L = workloads.sc_lattice(8, (1.0 / 0.2) ** (1.0 / 3.0), kT=1.2, seed=5)
    s = ref.RefSystem(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
    s.setup_integrator(0.005)
    s.add_langevin(1.2, 42)
    s.run(300)
    st = s.particles()
    for mode, name in ((0, "none"), (1, "shift"), (2, "xplor")):
        q = ref.RefSystem(st["pos"], (L, L, L), vel4=st["vel"])
        q.setup_lj(r_cut=3.0, r_buff=0.4, full=True, shift_mode=mode, r_on=2.0 if mode == 2 else None)
        q.set_virial(True)
        q.compute_forces(0)
        fo, vi = q.forces()
        out["force_" + name], out["virial_" + name] = fo, vi
    np.savez_compressed(os.path.join(HERE, "liquid_forces.npz"), pos=st["pos"], vel=st["vel"], L=L, **out)

    # (5) trajectories from that state: 40 NVE steps, 40 Langevin steps (seed 4, kT 1.2), sorter off
    for method in ("nve", "langevin"):
        q = ref.RefSystem(st["pos"], (L, L, L), vel4=st["vel"], image=st["image"])
        q.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
        q.set_virial(True)
        q.setup_integrator(0.005)
        q.add_nve() if method == "nve" else q.add_langevin(1.2, 4)
        q.run(40)
        p = q.particles()
        th = q.thermo()
        np.savez_compressed(os.path.join(HERE, "traj_%s.npz" % method), pos0=st["pos"], vel0=st["vel"],
                            image0=st["image"], L=L, pos=p["pos"], vel=p["vel"], accel=p["accel"], image=p["image"],
                            updates=q.nlist_stats()["updates"], thermo=np.array([th[k] for k in
                                                                                  ("ke", "pe", "T", "P", "ndof")]))

    # (6) isolated half steps incl. wrap / image flags and the limit option
    N, L = 512, 10.0
    rng = np.random.default_rng(11)
    pos = workloads.random_system(N, L, seed=4)
    pos[:16, 0] = np.float32(L / 2 - 1e-4)
    pos[16:32, 1] = np.float32(-L / 2 + 1e-5)
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = rng.normal(0, 1.5, (N, 3))
    vel[:, 3] = rng.uniform(0.5, 2.0, N)
    accel = rng.normal(0, 10.0, (N, 3)).astype(np.float32)
    fnet = rng.normal(0, 20.0, (N, 4)).astype(np.float32)
    res = dict(pos=pos, vel=vel, accel=accel, fnet=fnet, L=L)
    for limit, tag in ((None, "nolimit"), (0.01, "limit")):
        q = ref.RefSystem(pos, (L, L, L), vel4=vel)
        q.setup_lj(r_cut=2.0, r_buff=0.3, full=True)
        q.setup_integrator(0.005)
        q.add_nve(limit=limit)
        q.set_accel(accel)
        q.step_one(0)
        p = q.particles()
        res["s1_pos_" + tag], res["s1_vel_" + tag], res["s1_image_" + tag] = p["pos"], p["vel"], p["image"]
        q.step_two(0, fnet)
        p = q.particles()
        res["s2_vel_" + tag], res["s2_accel_" + tag] = p["vel"], p["accel"]
    q = ref.RefSystem(pos, (L, L, L), vel4=vel)
    q.setup_lj(r_cut=2.0, r_buff=0.3, full=True)
    q.setup_integrator(0.005)
    q.add_langevin(1.2, 4)
    q.step_two(123456, fnet)
    p = q.particles()
    res["lang_vel"], res["lang_accel"] = p["vel"], p["accel"]
    np.savez_compressed(os.path.join(HERE, "half_steps.npz"), **res)

    # (7) SFC sorter order
    pos, vel, L = workloads.sc_lattice(8, (1.0 / 0.2) ** (1.0 / 3.0), kT=1.2, seed=5)
    pos[:, :3] += np.random.default_rng(9).normal(0, 0.3, (len(pos), 3)).astype(np.float32)
    pos[:, :3] = np.clip(pos[:, :3], -L / 2, L / 2 - 1e-3)
    q = ref.RefSystem(pos, (L, L, L), vel4=vel)
    q.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
    q.sort_now()
    np.savez_compressed(os.path.join(HERE, "sfc_order.npz"), pos=pos, L=L, tag=q.particles()["tag"])
    print("golden fixtures written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
