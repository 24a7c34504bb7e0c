"""Multi-GPU parity check, run under torchrun with >= 2 ranks (one per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/mg_check.py [n_side] [steps]

The domain-decomposed engine (ghosts/migration over NVLink) against the single-GPU engine on the same global snapshot:
ghost sets equal to a brute-force numpy construction, forces within 1e-5 relative, particle count conserved through
migrations, trajectories agreeing to FP32 round-off growth, same rebuild count.  Prints MG_CHECK_OK on rank 0.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


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


def main():
    import torch
    import torch.distributed as dist
    import hoomd_b200
    from hoomd_b200 import workloads, domain

    n_side = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    rho, kT, r_cut, r_buff = 0.2, 1.2, 3.0, 0.4
    pos, vel, L = workloads.sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=kT, seed=11)
    rng = np.random.default_rng(5)
    pos[:, :3] += rng.normal(0, 0.12, (len(pos), 3)).astype(np.float32)
    pos[:, :3] = np.clip(pos[:, :3], -L / 2, np.nextafter(np.float32(L / 2), np.float32(0)))
    N = len(pos)
    box = (L, L, L)

    # ---- single-GPU reference on every rank's own GPU (cheap at this size) ----
    ref = hoomd_b200.System(pos, box, vel4=vel, device=local_rank)
    ref.set_pair_mode(0)
    ref.setup_lj(r_cut=r_cut, r_buff=r_buff)
    ref.compute_forces(0)
    f_ref = ref.forces(False)[0]

    mg = domain.DomainSystem(pos, box, r_ghost=r_cut + r_buff, vel4=vel, device=local_rank)
    mg.setup_lj(r_cut=r_cut, r_buff=r_buff)
    mg.compute_forces(0)
    counts = [None] * world
    dist.all_gather_object(counts, mg._sync_n())
    assert sum(counts) == N, counts

    # ghosts vs brute force
    lo, hi, per = domain.local_bounds(box, mg.grid, rank)
    gpos, gtag = mg.ghosts()
    exp = expected_ghost_tags(pos, box, lo, hi, per, np.float32(r_cut + r_buff))
    assert np.array_equal(np.sort(gtag), exp), (len(gtag), len(exp))
    # ghost positions are periodic images of their owners
    d = gpos[:, :3] - pos[gtag, :3]
    d -= np.round(d / np.float32(L)) * np.float32(L)
    assert np.abs(d).max() < 1e-5

    # forces by tag
    p = mg.particles()
    f_mg = mg.forces(False)[0]
    fr = f_ref[p["tag"]]
    scale = np.maximum(np.linalg.norm(fr[:, :3], axis=1, keepdims=True), np.abs(fr[:, :3]).mean())
    err = (np.abs(fr[:, :3] - f_mg[:, :3]) / scale).max()
    assert err < 1e-4, err
    eerr = np.abs(fr[:, 3] - f_mg[:, 3]).max() / np.abs(fr[:, 3]).mean()
    assert eerr < 1e-4, eerr

    # ---- trajectories ----
    for s in (ref, mg):
        s.setup_integrator(0.005)
        s.add_nve()
        s.set_sorter(100)
    ref.run(steps)
    mg.run(steps)
    counts = [None] * world
    dist.all_gather_object(counts, mg._sync_n())
    assert sum(counts) == N, counts
    g = mg.gather()
    assert np.array_equal(g["tag"], np.arange(N, dtype=np.uint32)), "particles lost or duplicated"
    a = ref.particles()
    order = np.argsort(a["tag"])
    rp, rv, ri = a["pos"][order], a["vel"][order], a["image"][order]
    # unwrapped positions (migration re-wraps; compare modulo the box)
    dpos = g["pos"][:, :3] - rp[:, :3]
    dpos -= np.round(dpos / np.float32(L)) * np.float32(L)
    perr, verr = np.abs(dpos).max(), np.abs(g["vel"][:, :3] - rv[:, :3]).max()
    unw_mg = g["pos"][:, :3].astype(np.float64) + g["image"] * float(L)
    unw_rf = rp[:, :3].astype(np.float64) + ri * float(L)
    uerr = np.abs(unw_mg - unw_rf).max()
    st_ref, st_mg, ms = ref.stats(), mg.stats(), mg.mg_stats()
    if rank == 0:
        print("N=%d grid=%s counts=%s force_err=%.2e pe_err=%.2e pos_err=%.2e vel_err=%.2e unwrapped_err=%.2e "
              "rebuilds ref=%d mg=%d mg_stats=%s" % (N, mg.grid, counts, err, eerr, perr, verr, uerr,
                                                     st_ref["updates"], st_mg["updates"], ms), flush=True)
    tol = 5e-3 * max(1.0, steps / 300.0)
    assert perr < tol and uerr < tol, (perr, uerr)
    assert abs(st_ref["updates"] - st_mg["updates"]) <= max(1, st_ref["updates"] // 20)
    assert ms["max_migrants"] > 0 or steps < 100, "no particle ever migrated: the test did not exercise migration"
    dist.barrier()
    if rank == 0:
        print("MG_CHECK_OK", flush=True)
    mg.close()
    ref.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
