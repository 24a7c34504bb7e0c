mkdir -p gpurun_out
python -m pytest tests -q -m gpu -s > gpurun_out/r2c_gpu_tests.log 2>&1; echo "tests rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r2c_gpu_tests.log | cut -c1-300 | head -40
grep "parity " gpurun_out/r2c_gpu_tests.log | cut -c1-250 | sort | uniq | head -80 > gpurun_out/r2c_parity_lines.txt
for w in lj64k lj1m kg1m; do
  python bench.py --steps 6 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/r2c_${w}_exact.json 2> gpurun_out/r2c_${w}_exact.err; echo "$w exact rc=$?"
  HOOMD_B200_LIB=$PWD/hoomd_b200/libhoomd_b200_fast.so python bench.py --steps 6 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/r2c_${w}_fast.json 2> gpurun_out/r2c_${w}_fast.err; echo "$w fast rc=$?"
  python - <<PY
import json
for v in ("exact","fast"):
    try:
        d=json.loads(open("gpurun_out/r2c_${w}_%s.json"%v).read().strip().splitlines()[-1])
        print("${w}", v, "TPS %.1f e2e %.1f kernel_ms %.4f rebuild_ms %.4f frac %.3f" % (d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["roofline"]["nlist_rebuild_ms"], d["roofline"]["frac"]))
    except Exception as e: print("${w}", v, "ERR", e)
PY
done
