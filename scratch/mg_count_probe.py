"""debug probe (2 GPUs): rebuild counters of the domain-decomposed engine vs the reference for growing step counts"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.distributed as dist
from hoomd_b200 import workloads, domain
import hoomd_b200
from oracle import ref as R
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
pos, vel, L = workloads.sc_lattice(24, (1 / 0.2) ** (1 / 3), kT=1.2, seed=11)
pos[:, :3] += np.random.default_rng(5).normal(0, 0.12, (len(pos), 3)).astype(np.float32)
pos[:, :3] = np.clip(pos[:, :3], -L / 2, np.nextafter(np.float32(L / 2), np.float32(0)))
for pre in (True, False):
  for steps in (1, 2, 3, 5, 8, 12, 20, 40):
    out = []
    for kind in ("ref", "one", "mg"):
        if kind == "ref":
            s = R.RefSystem(pos, (L, L, L), vel4=vel); s.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
        elif kind == "one":
            s = hoomd_b200.System(pos, (L, L, L), vel4=vel, device=lr); s.setup_lj(r_cut=3.0, r_buff=0.4)
        else:
            s = domain.DomainSystem(pos, (L, L, L), r_ghost=3.4, vel4=vel, device=lr); s.setup_lj(r_cut=3.0, r_buff=0.4)
        if pre: s.compute_forces(0)
        s.setup_integrator(0.005); s.add_nve(); s.set_sorter(100)
        s.run(steps)
        if kind == "ref": out.append(("ref", s.nlist_stats()["updates"]))
        else:
            st = s.stats(); out.append((kind, st["updates"], st["forced"], st["overflows"]))
        if kind != "ref": s.close()
    if rank == 0: print("pre" if pre else "nopre", steps, out, flush=True)
dist.barrier(); dist.destroy_process_group()
