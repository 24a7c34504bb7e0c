# back-off between two polls of the k_md_steps release word: 0 / 100 (default) / 300 / 1000 ns at 64k, then the bit-identity tests
mkdir -p gpurun_out
for ns in 100 0 300 1000; do
HOOMD_B200_PERSIST_SLEEP=$ns timeout 300 python bench.py --no-scale-ref --no-cpu-baseline > gpurun_out/sl_bench_lj64k_$ns.json 2> gpurun_out/sl_bench_lj64k_$ns.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/sl_bench_lj64k_$ns.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("sleep $ns ns: TPS %.1f e2e %.1f kernel %.5f" % (d["value"], d["e2e"]["value"], r["kernel_ms"]))
except Exception as e: print("$ns ERR", e)
PY
done
timeout 300 python -m pytest tests/test_parity_gpu.py -q -x -k "multistep or bit" 2>&1 | tail -2
