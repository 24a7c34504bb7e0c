# one-cluster cell list (HOOMD_B200_CELL_CLUSTER = 0 off | 8 | 16) at 64k: tests first, then bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_drivers_gpu.py tests/test_golden_gpu.py tests/test_parity_gpu.py tests/test_tile_gpu.py tests/test_full_size_gpu.py -q -x > gpurun_out/cc_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/cc_tests.log | cut -c1-400
for c in 0 16 8; do
HOOMD_B200_CELL_CLUSTER=$c timeout 600 python bench.py --no-scale-ref --no-cpu-baseline > gpurun_out/cc_bench_lj64k_c$c.json 2> gpurun_out/cc_bench_lj64k_c$c.err; echo "c=$c rc=$?"
done
timeout 600 python bench.py --steps 8 --warmup 3 --workload lj1m --no-scale-ref --no-cpu-baseline > gpurun_out/cc_bench_lj1m.json 2> gpurun_out/cc_bench_lj1m.err; echo "lj1m rc=$?"
python - <<PY
import json
for w in ("lj64k_c0","lj64k_c16","lj64k_c8","lj1m"):
    try:
        d=json.loads(open("gpurun_out/cc_bench_%s.json"%w).read().strip().splitlines()[-1]); r=d.get("roofline") or {}
        print(w, "TPS %.1f e2e %.1f kernel %.5f rebuild %.5f cell %.5f build %.5f launches %s" % (d["value"], d["e2e"]["value"], r.get("kernel_ms"), r.get("nlist_rebuild_ms"), r.get("cell_list_ms"), r.get("nlist_kernel_ms"), d.get("gpu_launches")))
    except Exception as e: print(w, "ERR", e); print(open("gpurun_out/cc_bench_%s.err"%w).read()[-800:])
PY
