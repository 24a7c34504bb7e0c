# last check of a tree: full GPU suite, smoke, the default bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -x -m gpu > gpurun_out/fin_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/fin_tests.log | cut -c1-300
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/fin_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/fin_smoke.log | cut -c1-300
timeout 300 python bench.py > gpurun_out/fin_bench_lj64k.json 2> gpurun_out/fin_bench_lj64k.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/fin_bench_lj64k.json").read().strip().splitlines()[-1]); r=d["roofline"]
print("lj64k TPS %.1f e2e %.1f kernel %.5f rebuild %.5f frac %.3f launches %s" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["nlist_rebuild_ms"], r["frac"], d["gpu_launches"]))
PY
