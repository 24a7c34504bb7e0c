"""profile target: warm a workload up, then run the micro-timers inside a cudaProfilerStart/Stop range
(ncu --profile-from-start off): which = 2 the single-kernel MD step, 1 the list rebuild"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "lj1m"
which = int(sys.argv[2]) if len(sys.argv) > 2 else 2
bench.EQUIL_STEPS = int(sys.argv[3]) if len(sys.argv) > 3 else 600
state, L = bench.make_state(wl)
s = bench.new_system(state, L, workload=wl)
s.run(300)
rt = torch.cuda.cudart()
rt.cudaProfilerStart()
ms = s.time_kernel(which, 40 if which == 5 else 4)
rt.cudaProfilerStop()
print(json.dumps(dict(workload=wl, which=which, ms=ms)))
