# threads per cell of k_nlist_tiled: 128 (default library) against 256 (libhoomd_b200_nl256.so) at 64k, 1M, Kob-Andersen
mkdir -p gpurun_out
for v in 128 256; do
L=$PWD/hoomd_b200/libhoomd_b200.so; [ $v = 256 ] && L=$PWD/hoomd_b200/libhoomd_b200_nl256.so
for w in lj64k lj1m ka256k; do
HOOMD_B200_LIB=$L timeout 600 python bench.py --steps 6 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/nlb_${w}_$v.json 2> gpurun_out/nlb_${w}_$v.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/nlb_${w}_$v.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("$w block $v TPS %.1f kernel %.5f rebuild %.5f cell %.5f build %.5f" % (d["value"], r["kernel_ms"], r["nlist_rebuild_ms"], r["cell_list_ms"], r["nlist_kernel_ms"]))
except Exception as e: print("$w $v ERR", e)
PY
done; done
HOOMD_B200_LIB=$PWD/hoomd_b200/libhoomd_b200_nl256.so timeout 600 python -m pytest tests/test_drivers_gpu.py tests/test_golden_gpu.py -q -x 2>&1 | tail -2
