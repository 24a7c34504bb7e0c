"""Step-loop overhead probe: TPS with rebuilds disabled (check_period huge) per stepping mode; timing only."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, hoomd_b200
bench.EQUIL_STEPS = 2000
state, L = bench.make_state("lj64k")
for cp in (1, 10**9):
    for ug in (2, 1, 3, 0):
        s = hoomd_b200.System(state["pos"], (L, L, L), vel4=state["vel"], image=state["image"])
        s.setup_lj(r_cut=3.0, r_buff=0.4, check_period=cp)
        s.setup_integrator(0.005)
        s.add_nve()
        s.set_tuning(use_graph=ug)
        s.run(200)
        ms = min(s.run(400)[1] for _ in range(3))
        print(json.dumps(dict(check_period=cp, use_graph=ug, us_per_step=1e3 * ms / 400, pair_us=1e3 * s.time_kernel(0, 200),
                              rebuild_us=1e3 * s.time_kernel(1, 50), stats=s.stats())), flush=True)
        s.close()
