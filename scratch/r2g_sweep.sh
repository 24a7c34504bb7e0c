# single GPU: full test suite, then the k_md_step latency-trim A/B (arithmetic head list, L1 prefetch, lanes per particle)
mkdir -p gpurun_out
python -m pytest tests -q -m gpu > gpurun_out/r2g_gpu_tests.log 2>&1; echo "tests rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r2g_gpu_tests.log | cut -c1-300 | head -20
for w in lj64k lj1m; do
  for cfg in "1 1 0" "0 1 0" "1 0 0" "0 0 0" "1 1 4"; do
    set -- $cfg
    HOOMD_B200_HEAD_ARITH=$1 HOOMD_B200_PREFETCH=$2 HOOMD_B200_PAIR_TPP=$3 python bench.py --steps 8 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/r2g_${w}_h$1_p$2_t$3.json 2> gpurun_out/r2g_${w}_h$1_p$2_t$3.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2g_${w}_h$1_p$2_t$3.json").read().strip().splitlines()[-1])
    print("${w} head_arith=$1 prefetch=$2 tpp=$3: TPS %.1f e2e %.1f kernel_ms %.4f rebuild_ms %.4f" % (d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["roofline"]["nlist_rebuild_ms"]))
except Exception as e: print("${w} $cfg ERR", e)
PY
  done
done
