# cell-tile pair path (pair mode 1) against the CSR walk (mode 0) at 1M and 64k: bench lines + launch list of the tile rebuild
mkdir -p gpurun_out
for w in lj1m lj64k; do
for m in 0 1; do
HOOMD_B200_PAIR_MODE=$m timeout 600 python bench.py --steps 8 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/tile_bench_${w}_mode$m.json 2> gpurun_out/tile_bench_${w}_mode$m.err; echo "$w mode $m rc=$?"
done; done
python - <<PY
import json
for w in ("lj1m","lj64k"):
  for m in (0,1):
    try:
        d=json.loads(open("gpurun_out/tile_bench_%s_mode%d.json"%(w,m)).read().strip().splitlines()[-1]); r=d.get("roofline") or {}
        print(w, m, "TPS %.1f e2e %.1f kernel %s rebuild %s cell %s build %s frac %s" % (d["value"], d["e2e"]["value"], r.get("kernel_ms"), r.get("nlist_rebuild_ms"), r.get("cell_list_ms"), r.get("nlist_kernel_ms"), r.get("frac")))
    except Exception as e: print(w, m, "ERR", e); print(open("gpurun_out/tile_bench_%s_mode%d.err"%(w,m)).read()[-600:])
PY
HOOMD_B200_PAIR_MODE=1 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tile_launches_rebuild_1m.csv python tools/prof_step.py lj1m 1 300 > gpurun_out/tile_ncu1.log 2>&1; echo "ncu rebuild rc=$?"
HOOMD_B200_PAIR_MODE=1 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tile_launches_step_1m.csv python tools/prof_step.py lj1m 2 300 > gpurun_out/tile_ncu2.log 2>&1; echo "ncu step rc=$?"
python tools/ncu_summary.py launches gpurun_out/tile_launches_rebuild_1m.csv gpurun_out/tile_launches_rebuild_1m.md; head -40 gpurun_out/tile_launches_rebuild_1m.md
python tools/ncu_summary.py launches gpurun_out/tile_launches_step_1m.csv gpurun_out/tile_launches_step_1m.md; head -20 gpurun_out/tile_launches_step_1m.md
