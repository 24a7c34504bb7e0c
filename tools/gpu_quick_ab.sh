# quick regression pass after a kernel change: list / parity tests, then the 64k and 1M bench lines
mkdir -p gpurun_out
T=${TAG:-q}
timeout 900 python -m pytest tests/test_drivers_gpu.py tests/test_golden_gpu.py tests/test_parity_gpu.py tests/test_bond_gpu.py tests/test_tile_gpu.py -q -x > gpurun_out/${T}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${T}_tests.log | cut -c1-400
timeout 600 python bench.py --no-scale-ref --no-cpu-baseline > gpurun_out/${T}_bench_lj64k.json 2> gpurun_out/${T}_bench_lj64k.err; echo "64k rc=$?"
for w in lj1m kg1m; do
timeout 600 python bench.py --steps 8 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err; echo "$w rc=$?"
done
python - <<PY
import json
for w in ("lj64k","lj1m","kg1m"):
    try:
        d=json.loads(open("gpurun_out/${T}_bench_%s.json"%w).read().strip().splitlines()[-1]); r=d.get("roofline") or {}
        print(w, "TPS %.1f e2e %.1f kernel %.5f rebuild %.5f cell %.5f build %.5f launches %s" % (d["value"], d["e2e"]["value"], r.get("kernel_ms"), r.get("nlist_rebuild_ms"), r.get("cell_list_ms"), r.get("nlist_kernel_ms"), d.get("gpu_launches")))
    except Exception as e: print(w, "ERR", e); print(open("gpurun_out/${T}_bench_%s.err"%w).read()[-800:])
PY
