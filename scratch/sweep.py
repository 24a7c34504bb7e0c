"""TPS sweep over pair threads-per-particle / stepping mode on one workload (run on the GPU box)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import numpy as np

TPPS = [int(x) for x in os.environ.get("TPPS", "1,2,4,8").split(",")]
wl = sys.argv[1] if len(sys.argv) > 1 else "lj64k"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
bench.EQUIL_STEPS = int(sys.argv[3]) if len(sys.argv) > 3 else None
state, L = bench.make_state(wl)
for ug in (2,):
    for tpp in TPPS:
        bench.USE_GRAPH = ug
        s = bench.new_system(state, L)
        s.set_tuning(pair_tpp=tpp, use_graph=ug)
        pm = int(os.environ.get("HOOMD_B200_PAIR_MODE", "-1"))
        if pm != 0:
            tile = tuple(int(x) for x in os.environ.get("TILE", "0,0,0").split(","))
            s.set_pair_mode(pm, tile=tile, tpp=tpp)
        s.run(steps)
        ms = min(s.run(steps)[1] for _ in range(3))
        print(json.dumps(dict(workload=wl, use_graph=ug, tpp=tpp, tps=steps / (ms * 1e-3), us_per_step=1e3 * ms / steps,
                              pair_us=1e3 * s.time_kernel(0, 200), rebuild_us=1e3 * s.time_kernel(1, 50),
                              stats=s.stats(), pair_mode=s.pair_mode())), flush=True)
        s.close()
        if ug == 0 and tpp >= 8:
            break
