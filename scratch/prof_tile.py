"""profile target: a few stream-launched cell-tile MD steps and one rebuild on workload argv[1]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "lj1m"
bench.EQUIL_STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 300
state, L = bench.make_state(wl)
s = bench.new_system(state, L)
s.run(200)
print(json.dumps(dict(step_ms=s.time_kernel(2, 6), pair_ms=s.time_kernel(0, 3), rebuild_ms=s.time_kernel(1, 2),
                      pair_mode=s.pair_mode())))
