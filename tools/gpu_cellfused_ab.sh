# single GPU: one-launch cell list (custom grid barrier) on / off at 64k and 27k, then the full suite
mkdir -p gpurun_out
for c in 1 0; do
HOOMD_B200_CELL_FUSED=$c timeout 300 python bench.py --steps 8 --warmup 3 --workload lj64k --no-scale-ref --no-cpu-baseline > gpurun_out/cf_lj64k_c$c.json 2> gpurun_out/cf_lj64k_c$c.err; echo "bench fused=$c rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/cf_lj64k_c$c.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("lj64k cell_fused=$c TPS %.1f e2e %.1f kernel %.4f rebuild %.4f cell %.4f nlist %.4f launches %s" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["nlist_rebuild_ms"], r["cell_list_ms"], r["nlist_kernel_ms"], d["gpu_launches"]))
except Exception as e: print("ERR", e)
PY
done
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/cf_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/cf_tests.log | cut -c1-400
