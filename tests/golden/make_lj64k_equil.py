"""Generates tests/golden/lj64k_equil.npz: the equilibrated start state of BASELINE configs[1] (LJ liquid 64,000 particles,
rho 0.2) that BOTH arms of bench.py load -- made by the UNMODIFIED reference (oracle/_ref/libhoomd_ref.so): 40^3 simple-cubic
lattice, Gaussian velocities kT 1.2 (numpy default_rng(42)), 10,000 Langevin steps kT 1.2 seed 42 dt 0.005 with md.pair.lj
r_cut 3.0 / nlist.cell r_buff 0.4.  Also stores the reference's neighbor-list row lengths and
thermodynamic quantities of that state (the full-size parity test compares forces live against oracle/_ref).
Run in the build container only (about five minutes on one core):  python tests/golden/make_lj64k_equil.py
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ref  # noqa: E402
from hoomd_b200 import workloads  # noqa: E402


def main():
    pos, vel, L = workloads.lj_liquid(64000, rho=0.2, kT=1.2, seed=42)
    s = ref.RefSystem(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=3.0, r_buff=0.4, full=False)  # the CPU md.pair.lj default: half list
    s.setup_integrator(0.005)
    s.add_langevin(1.2, 42)
    s.set_sorter(100)                                # hoomd.update.sort default period on the CPU
    t0 = time.time()
    s.run(10000)
    print("equilibrated in %.0f s" % (time.time() - t0))
    p = s.particles()
    order = np.argsort(p["tag"], kind="stable")     # back to tag order (= index order of a fresh system)
    st = {k: np.ascontiguousarray(p[k][order]) for k in ("pos", "vel", "image")}
    # the reference's own record of that state (full list): row lengths and thermodynamics, checked by
    # tests/test_full_size_gpu.py next to the live comparison
    q = ref.RefSystem(st["pos"], (L, L, L), vel4=st["vel"], image=st["image"])
    q.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
    q.set_virial(True)
    q.setup_integrator(0.005)
    q.add_nve()
    q.compute_forces(0)
    n = q.nlist()[0]
    th = q.thermo()
    np.savez_compressed(os.path.join(HERE, "lj64k_equil.npz"), L=np.float32(L), pos=st["pos"], vel=st["vel"],
                        image=st["image"].astype(np.int8), n_neigh=n.astype(np.uint16),
                        thermo=np.array([th[k] for k in ("ke", "pe", "T", "P")]))
    print("T = %.4f, mean neighbors %.2f" % (th["T"], n.mean()))


if __name__ == "__main__":
    main()
