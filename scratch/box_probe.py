import time, numpy as np, torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hoomd_b200 as hb
from hoomd_b200 import workloads
pos, vel, L = workloads.sc_lattice(40, (1.0/0.2)**(1/3), kT=1.2, seed=1)
s = hb.System(pos, (L, L, L), vel4=vel)
s.setup_lj(r_cut=3.0, r_buff=0.4); s.setup_integrator(0.005); s.add_langevin(1.2, 3); s.set_sorter(300)
s.run(3000)
Lc = float(L); tb = tr = tl = tg = 0.0; n = 1000
for k in range(n):
    Lc *= (1 - 2e-4)
    t0 = time.perf_counter(); s.set_box(np.float32(Lc), np.float32(Lc), np.float32(Lc), True); t1 = time.perf_counter()
    el, ms = s.run(1); t2 = time.perf_counter()
    tb += t1 - t0; tr += t2 - t1; tl += el; tg += ms * 1e-3
print("per iteration us: set_box %.1f  run(1) %.1f  (step loop wall %.1f, gpu events %.1f)" % (tb/n*1e6, tr/n*1e6, tl/n*1e6, tg/n*1e6))
t0 = time.perf_counter()
for k in range(n): s.run(1)
print("run(1) without box change us: %.1f" % ((time.perf_counter()-t0)/n*1e6))
print(s.stats())
