# single GPU: full suite + smoke + default bench lines (64k with cpu_baseline, reference arm, 1M, KA, KG)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/val_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/val_tests.log | cut -c1-400
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/val_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/val_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/val_bench_lj64k.json 2> gpurun_out/val_bench_lj64k.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/val_bench_ref.json 2> gpurun_out/val_bench_ref.err; echo "ref rc=$?"
for w in lj1m ka256k kg1m; do
timeout 600 python bench.py --steps 8 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/val_bench_$w.json 2> gpurun_out/val_bench_$w.err; echo "$w rc=$?"
done
python - <<PY
import json
for w in ("lj64k","ref","lj1m","ka256k","kg1m"):
    try:
        d=json.loads(open("gpurun_out/val_bench_%s.json"%w).read().strip().splitlines()[-1]); r=d.get("roofline") or {}
        print(w, "TPS %.1f e2e %.1f kernel %s rebuild %s frac %s launches %s cpu %s" % (d["value"], d["e2e"]["value"], r.get("kernel_ms"), r.get("nlist_rebuild_ms"), r.get("frac"), d.get("gpu_launches"), json.dumps(d.get("cpu_baseline"))[:200]))
    except Exception as e: print(w, "ERR", e)
PY
