# single GPU: multi-step kernel with release-atomic barrier: A/B, new test, per-step timing, ncu capture
mkdir -p gpurun_out
for p in 1 0; do
HOOMD_B200_PERSIST=$p timeout 300 python bench.py --steps 8 --warmup 3 --workload lj64k --no-scale-ref --no-cpu-baseline > gpurun_out/ms_lj64k_p$p.json 2> gpurun_out/ms_lj64k_p$p.err; echo "bench persist=$p rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ms_lj64k_p$p.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("lj64k persist=$p TPS %.1f e2e %.1f kernel %.4f one-step %.4f rebuild %.4f launches %s frac %.3f" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["one_step_kernel_ms"], r["nlist_rebuild_ms"], d["gpu_launches"], r["frac"]))
except Exception as e: print("ERR", e)
PY
done
timeout 600 python -m pytest tests/test_parity_gpu.py -q -x > gpurun_out/ms_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/ms_tests.log | cut -c1-400
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
timeout 600 $NCU -k regex:k_md_steps -c 1 -o gpurun_out/prof_mdsteps_ms_64k python tools/prof_step.py lj64k 5 > gpurun_out/ms_ncu1.log 2>&1; tail -2 gpurun_out/ms_ncu1.log | cut -c1-300
