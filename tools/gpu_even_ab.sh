# single GPU: k_md_steps with the particles spread evenly over a grid that is a multiple of the SM count, on / off
mkdir -p gpurun_out
for e in 1 0 1 0; do
HOOMD_B200_PERSIST_EVEN=$e timeout 300 python bench.py --steps 8 --warmup 3 --workload lj64k --no-scale-ref --no-cpu-baseline > gpurun_out/even_lj64k_e$e.json 2> gpurun_out/even_lj64k_e$e.err; echo "bench even=$e rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/even_lj64k_e$e.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("lj64k even=$e TPS %.1f e2e %.1f kernel %.4f one-step %.4f rebuild %.4f" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["one_step_kernel_ms"], r["nlist_rebuild_ms"]))
except Exception as e: print("ERR", e)
PY
done
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_group_gpu.py -q -x > gpurun_out/even_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/even_tests.log | cut -c1-300
