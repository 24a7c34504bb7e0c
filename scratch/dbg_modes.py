import sys, numpy as np
sys.path.insert(0, '.')
import hoomd_b200 as hb
def mk(ug, method, sorter):
    pos, vel, L = hb.workloads.sc_lattice(12, (1/0.2)**(1/3), kT=1.2, seed=5)
    g = hb.System(pos, (L,L,L), vel4=vel)
    g.setup_lj(r_cut=3.0, r_buff=0.4); g.setup_integrator(0.005)
    g.add_nve() if method=='nve' else g.add_langevin(1.2, 7)
    if sorter: g.set_sorter(sorter)
    g.set_tuning(pair_tpp=4, use_graph=ug)
    return g
import itertools
for method, sorter, inc in itertools.product(('nve','langevin'), (0,100), (450, 150)):
  if True:
    gs = [mk(ug, method, sorter) for ug in (0,1,2)]
    for it in range(450//inc):
        for g in gs: g.run(inc)
        ps = [g.particles() for g in gs]
        sts = [g.stats() for g in gs]
        eq = [all(np.array_equal(ps[0][k].view(np.uint32), p[k].view(np.uint32)) for k in ('pos','vel','accel','image','tag')) for p in ps[1:]]
        print(method, sorter, inc, 'ts', gs[0].timestep, eq, [(s['updates'], s['forced'], s['overflows'], s['Nmax0']) for s in sts], flush=True)
        if not all(eq): break
