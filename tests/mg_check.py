"""Multi-GPU parity check, run under torchrun with >= 2 ranks (one per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/mg_check.py [n_side] [steps]

The domain-decomposed engine (ghosts / migration over NVLink) against the UNMODIFIED reference CPU classes
(oracle/_ref, RefSystem) on the same global snapshot, every comparison keyed by particle TAG:
  * ghost sets equal to a brute-force numpy construction, ghost positions are periodic images of their owners;
  * forces and per-particle energies within 1e-5 of the sums of absolute pair terms (north_star bar,
    golden_util.assert_parity) -- neighbor sets are therefore the reference's: a missing or extra pair inside r_cut would
    show up at 1e-2;
  * neighbor-list row lengths equal to the reference's per tag (ghost rows included through their owners);
  * after `steps` NVE steps: particle count conserved through migrations, EQUAL number of list rebuilds, trajectories
    agreeing to FP32 round-off growth.
`parity_pass()` is also what `bench.py --gpus N` runs before timing (its line carries parity_checked / max_force_err).
Prints MG_CHECK_OK on rank 0.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def expected_ghost_tags(pos, L, lo, hi, per, r_ghost):
    """tags (with multiplicity) of particles whose periodic image lies in the shell [lo - r, hi + r) \\ [lo, hi) of a
    rank, restricted to the split axes; brute force over the 27 images."""
    out = []
    split = [p == 0 for p in per]
    shifts = [(-1, 0, 1) if split[a] else (0,) for a in range(3)]
    for sx in shifts[0]:
        for sy in shifts[1]:
            for sz in shifts[2]:
                q = pos[:, :3] + np.array([sx * L[0], sy * L[1], sz * L[2]], np.float32)
                inside_ext = np.ones(len(q), bool)
                inside_own = np.ones(len(q), bool)
                for a in range(3):
                    if split[a]:
                        inside_ext &= (q[:, a] >= lo[a] - r_ghost) & (q[:, a] < hi[a] + r_ghost)
                        inside_own &= (q[:, a] >= lo[a]) & (q[:, a] < hi[a])
                out.append(np.nonzero(inside_ext & ~inside_own)[0])
    return np.sort(np.concatenate(out))


def parity_pass(pos, vel, image, L, steps, local_rank, r_cut=3.0, r_buff=0.4, check_ghosts=True, verbose=True, ntypes=1,
                lj=None, method="nve", pair=None, bonds=None):
    """One domain-decomposed system against RefSystem on the same snapshot (every rank calls this; torch.distributed must be
    initialised).  Returns a dict of measured maxima; raises AssertionError on a parity failure.  ntypes / lj: a mixture with
    per-pair coefficients (type ids in pos[:, 3]); method "langevin": the noise is keyed by global tag, so the decomposed
    trajectory follows the single-domain reference as well.  pair = (kind, coefficients): another pair evaluator
    (md.pair.gauss / yukawa / mie / force_shifted_lj / morse) through setup_pair on all three sides.  bonds: (nbonds, 2)
    GLOBAL tags -- FENE bonds + bond exclusions (Kremer-Grest): every rank holds the whole tag-keyed table and reaches a
    partner on another rank through its ghost copy."""
    import torch.distributed as dist
    import golden_util as G
    from hoomd_b200 import domain
    from oracle import ref as R

    rank, world = dist.get_rank(), dist.get_world_size()
    N = len(pos)
    box = (L, L, L)
    if not R.available():
        raise RuntimeError("oracle/_ref/libhoomd_ref.so is missing: the multi-GPU parity pass needs the reference classes")

    lj = dict(lj or {})
    r_cut_max = float(np.max(lj.get("r_cut", r_cut)))
    lj.setdefault("r_cut", r_cut)
    mg = domain.DomainSystem(pos, box, r_ghost=r_cut_max + r_buff, vel4=vel, image=image, device=local_rank, ntypes=ntypes)
    def configure(sys_, **extra):
        if bonds is not None:
            sys_.set_bonds(bonds)
        if pair is not None:
            sys_.setup_pair(pair[0], r_cut=lj["r_cut"], r_buff=r_buff, **pair[1], **extra)
        else:
            sys_.setup_lj(r_buff=r_buff, **lj, **extra)
        if bonds is not None:
            sys_.setup_fene(k=30.0, r0=1.5, epsilon=1.0, sigma=1.0)
            sys_.exclusions_from_bonds(True)

    configure(mg)
    mg.compute_forces(0)
    counts = [None] * world
    dist.all_gather_object(counts, mg._sync_n())
    assert sum(counts) == N, counts

    if check_ghosts:
        lo, hi, per = domain.local_bounds(box, mg.grid, rank)
        gpos, gtag = mg.ghosts()
        exp = expected_ghost_tags(pos, box, lo, hi, per, np.float32(r_cut_max + r_buff))
        assert np.array_equal(np.sort(gtag), exp), (len(gtag), len(exp))
        d = gpos[:, :3] - pos[gtag, :3]
        d -= np.round(d / np.float32(L)) * np.float32(L)
        assert np.abs(d).max() < 1e-5

    # ---- the reference's answer (CPU classes, full list) and the scales of the 1e-5 bar, on every rank's host cores ----
    r = R.RefSystem(pos, box, ntypes=ntypes, vel4=vel, image=image, **(dict(n_bond_types=1) if bonds is not None else {}))
    configure(r, full=True)
    r.compute_forces(0)
    f_ref = r.forces(False)[0]
    if bonds is not None:  # Integrator::computeNetForce: net = pair + bond, added in FP32
        f_ref = f_ref + r.force_component(1)[0]
    rn = r.nlist()[0]
    sc = G.abs_scales(configure, pos, box, ntypes=ntypes, n_bond_types=1 if bonds is not None else 0)

    p = mg.particles()
    tag = p["tag"]
    f_mg = mg.forces(False)[0]
    out = G.assert_parity(f_ref[tag], None, f_mg, None, (sc[0][tag], sc[1][tag], sc[2][:, tag]),
                          "rank %d of %d vs reference" % (rank, world), verbose=False)
    gn = mg.nlist(counts_only=True)[0][:mg.N]
    assert np.array_equal(gn, rn[tag]), "neighbor-list row lengths differ from the reference's"

    # ---- trajectories ----
    res = dict(max_force_err=out["force"], max_energy_err=out["energy"])
    if steps:
        for s in (r, mg):
            s.setup_integrator(0.005)
            if method == "langevin":
                s.add_langevin(1.0, 11, gamma=[1.0, 0.5][:ntypes] if ntypes <= 2 else [1.0] * ntypes)
            elif method == "nvt":
                s.add_nvt(1.0, 0.4)
            else:
                s.add_nve()
            s.set_sorter(100)
        r.run(steps)
        mg.run(steps)
        dist.all_gather_object(counts, mg._sync_n())
        assert sum(counts) == N, counts
        g = mg.gather()
        assert np.array_equal(g["tag"], np.arange(N, dtype=np.uint32)), "particles lost or duplicated"
        a = r.particles()
        order = np.argsort(a["tag"])
        rp, rv, ri = a["pos"][order], a["vel"][order], a["image"][order]
        dpos = g["pos"][:, :3] - rp[:, :3]
        dpos -= np.round(dpos / np.float32(L)) * np.float32(L)     # migration re-wraps: compare modulo the box
        unw_mg = g["pos"][:, :3].astype(np.float64) + g["image"] * float(L)
        unw_rf = rp[:, :3].astype(np.float64) + ri * float(L)
        res.update(pos_err=float(np.abs(dpos).max()), vel_err=float(np.abs(g["vel"][:, :3] - rv[:, :3]).max()),
                   unwrapped_err=float(np.abs(unw_mg - unw_rf).max()), rebuilds_ref=int(r.nlist_stats()["updates"]),
                   # NeighborList::getNumUpdates = normal + forced updates (hoomd/md/NeighborList.h:423-426)
                   rebuilds_mg=int(mg.stats()["updates"] + mg.stats()["forced"]), mg_stats=mg.mg_stats(),
                   mg_timing=mg.mg_timing())
        if method == "nvt":
            # the thermostat variables of every rank follow the single-domain reference (kinetic energy summed over the ranks)
            res.update(nvt_ref=[float(x) for x in r.nvt_state()], nvt_mg=[float(x) for x in mg.nvt_state()])
            assert np.allclose(res["nvt_mg"], res["nvt_ref"], rtol=1e-3, atol=1e-5), res
            assert abs(res["nvt_ref"][0]) > 1e-3, "the thermostat did not act"
        tol = 5e-3 * max(1.0, steps / 300.0)
        assert res["pos_err"] < tol and res["unwrapped_err"] < tol, res
        assert res["rebuilds_ref"] == res["rebuilds_mg"], res
    # the worst rank's figures
    allres = [None] * world
    dist.all_gather_object(allres, {k: v for k, v in res.items() if k not in ("mg_stats", "mg_timing")})
    for k in ("max_force_err", "max_energy_err", "pos_err", "vel_err", "unwrapped_err"):
        if k in res:
            res[k] = max(q[k] for q in allres)
    res.update(N=N, grid=list(mg.grid), counts=counts, steps=steps)
    if verbose and rank == 0:
        print("mg parity vs reference: %s" % res, flush=True)
    mg.close()
    return res


def main():
    import torch
    import torch.distributed as dist
    from hoomd_b200 import workloads

    n_side = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
    ntypes = int(sys.argv[3]) if len(sys.argv) > 3 else 1     # 2: Kob-Andersen-like mixture under Langevin
    pair_kind = sys.argv[4] if len(sys.argv) > 4 else "lj"    # morse | yukawa | gauss: another evaluator (one type, NVE)
    method = sys.argv[5] if len(sys.argv) > 5 else ""         # nvt: Nose-Hoover thermostat (one type)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    rho, kT = 0.2, 1.2
    kw = {}
    if ntypes == 2:
        rho = 0.6
        pos, vel, L = workloads.sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=kT, seed=11, ntypes=2,
                                           type_fractions=[0.8, 0.2])
        kw = dict(ntypes=2, method="langevin", r_cut=2.5,
                  lj=dict(r_cut=[[2.5, 2.0], [2.0, 2.2]], epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]],
                          shift_mode=1))
    else:
        pos, vel, L = workloads.sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=kT, seed=11)
        coeffs = {"morse": dict(D0=1.0, alpha=3.0, r0=1.2), "yukawa": dict(epsilon=2.0, kappa=1.5),
                  "gauss": dict(epsilon=1.5, sigma=0.9)}
        if pair_kind in coeffs:
            kw = dict(pair=(pair_kind, coeffs[pair_kind]), r_cut=2.5)
    rng = np.random.default_rng(5)
    pos[:, :3] += rng.normal(0, 0.12 if ntypes == 1 else 0.05, (len(pos), 3)).astype(np.float32)
    pos[:, :3] = np.clip(pos[:, :3], -L / 2, np.nextafter(np.float32(L / 2), np.float32(0)))
    if pair_kind == "kg":
        # Kremer-Grest: chains of 8 beads threaded through the lattice, WCA + FENE, bond exclusions
        pos, vel, L, bonds = workloads.polymer_lattice(n_side, 8, rho=0.85, kT=1.0, seed=11)
        rng = np.random.default_rng(5)
        pos[:, :3] += rng.normal(0, 0.03, (len(pos), 3)).astype(np.float32)
        pos[:, :3] = np.clip(pos[:, :3], -L / 2, np.nextafter(np.float32(L / 2), np.float32(0)))
        kw = dict(bonds=bonds, r_cut=2.0 ** (1.0 / 6.0), lj=dict(r_cut=2.0 ** (1.0 / 6.0), shift_mode=1))
    if method:
        kw["method"] = method
    res = parity_pass(pos, vel, None, L, steps, local_rank, **kw)
    assert res["mg_stats"]["max_migrants"] > 0 or steps < 100, "no particle ever migrated: migration was not exercised"
    dist.barrier()
    if rank == 0:
        print("MG_CHECK_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
