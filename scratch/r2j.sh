# single GPU: conditional-graph unroll A/B at 64k (and KA 256k), engine tests under unroll 8
mkdir -p gpurun_out
HOOMD_B200_GRAPH_UNROLL=8 python -m pytest tests/test_group_gpu.py tests/test_parity_gpu.py tests/test_nvt_gpu.py tests/test_bond_gpu.py tests/test_box_resize_gpu.py -q -x > gpurun_out/r2j_tests_u8.log 2>&1; echo "tests(u8) rc=$?"; tail -3 gpurun_out/r2j_tests_u8.log | cut -c1-300
for w in lj64k ka256k; do
for u in 2 4 8 16; do
HOOMD_B200_GRAPH_UNROLL=$u python bench.py --steps 8 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/r2j_${w}_u$u.json 2> gpurun_out/r2j_${w}_u$u.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2j_${w}_u$u.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("$w unroll $u TPS %.1f e2e %.1f kernel %.4f rebuild %.4f launches %s" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["nlist_rebuild_ms"], d["gpu_launches"]))
except Exception as e: print("$w $u ERR", e)
PY
done
done
