mkdir -p gpurun_out
HOOMD_B200_LIB=$PWD/hoomd_b200/libhoomd_b200_trace.so timeout 600 python tools/trace_persist.py lj64k > gpurun_out/trace_persist_64k.json 2> gpurun_out/trace_persist_64k.err; echo "rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/trace_persist_64k.json"))
for k,v in d.items():
    print(k, (v if not isinstance(v,dict) or "mean_us" not in v else "mean %.2f p10 %.2f p50 %.2f p90 %.2f max %.2f"%(v["mean_us"],v["p10"],v["p50"],v["p90"],v["max"])))
PY
tail -5 gpurun_out/trace_persist_64k.err
timeout 600 python -m pytest tests/test_parity_gpu.py -q -x > gpurun_out/tr_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tr_tests.log | cut -c1-300
timeout 600 python bench.py --no-scale-ref --no-cpu-baseline > gpurun_out/tr_bench_lj64k.json 2> gpurun_out/tr_bench_lj64k.err; echo "64k rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/tr_bench_lj64k.json").read().strip().splitlines()[-1]); r=d["roofline"]
print("lj64k TPS %.1f e2e %.1f kernel %.5f rebuild %.5f" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["nlist_rebuild_ms"]), d["e2e"]["ms_runs"])
PY
HOOMD_B200_PERSIST_EARLY_FENCE=1 timeout 600 python bench.py --no-scale-ref --no-cpu-baseline > gpurun_out/tr_bench_lj64k_earlyfence.json 2> gpurun_out/tr_bench_lj64k_earlyfence.err; echo "64k early-fence rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/tr_bench_lj64k_earlyfence.json").read().strip().splitlines()[-1]); r=d["roofline"]
print("lj64k EARLY-FENCE TPS %.1f e2e %.1f kernel %.5f rebuild %.5f" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["nlist_rebuild_ms"]))
PY
