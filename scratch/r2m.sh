# single GPU: multi-step kernel (k_md_steps) A/B at 64k + the suites that exercise it
mkdir -p gpurun_out
for p in 0 1; do
HOOMD_B200_PERSIST=$p timeout 300 python bench.py --steps 8 --warmup 3 --workload lj64k --no-scale-ref --no-cpu-baseline > gpurun_out/r2m_lj64k_p$p.json 2> gpurun_out/r2m_lj64k_p$p.err; echo "bench persist=$p rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2m_lj64k_p$p.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("lj64k persist=$p TPS %.1f e2e %.1f kernel %.4f rebuild %.4f launches %s ms_runs %s" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["nlist_rebuild_ms"], d["gpu_launches"], [round(x,1) for x in d["ms_runs"]]))
except Exception as e: print("ERR", e)
PY
done
tail -3 gpurun_out/r2m_lj64k_p1.err | cut -c1-300
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/r2m_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2m_tests.log | cut -c1-400
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2m_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2m_smoke.log | cut -c1-300
