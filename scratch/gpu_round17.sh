timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -8
B="--no-cpu-baseline --no-scale-ref"
for w in lj64k lj1m kg1m; do
timeout 300 python bench.py $B --workload $w > gpurun_out/r1s_${w}.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python - <<PY
import json; d=json.load(open("gpurun_out/r1s_${w}.json")); print("$w", round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["frac"])
PY
done
python - <<'PY'
# box_resize throughput: 64k LJ liquid, compression by 0.02 % every step (period 1) and every 10 steps
import time, numpy as np, torch
import hoomd_b200 as hb
from hoomd_b200 import workloads
pos, vel, L = workloads.sc_lattice(40, (1.0/0.2)**(1/3), kT=1.2, seed=1)
for period in (1, 10):
    s = hb.System(pos, (L, L, L), vel4=vel)
    s.setup_lj(r_cut=3.0, r_buff=0.4); s.setup_integrator(0.005); s.add_langevin(1.2, 3); s.set_sorter(300)
    s.run(2000)
    Lc = float(L); torch.cuda.synchronize(); t0 = time.time(); n = 0
    for k in range(2000 // period):
        Lc *= (1 - 2e-4) ** period
        s.set_box(np.float32(Lc), np.float32(Lc), np.float32(Lc), True)
        s.run(period); n += period
    torch.cuda.synchronize(); dt = time.time() - t0
    print("box_resize period", period, "steps/s", round(n / dt), "builds", s.stats()["updates"], "overflows", s.stats()["overflows"])
PY
