# single GPU: NPT tests, then bench lines with the cell-list / nlist split
mkdir -p gpurun_out
python -m pytest tests/test_npt_gpu.py -q -x > gpurun_out/r2i_npt.log 2>&1; echo "npt rc=$?"; tail -30 gpurun_out/r2i_npt.log | cut -c1-400
for w in lj64k lj1m; do
python bench.py --steps 8 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/r2i_$w.json 2> gpurun_out/r2i_$w.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2i_$w.json").read().strip().splitlines()[-1]); r=d["roofline"]
print("$w TPS %.1f kernel %.4f rebuild %.4f cell %.4f nlist %.4f pair %.4f" % (d["value"], r["kernel_ms"], r["nlist_rebuild_ms"], r["cell_list_ms"], r["nlist_kernel_ms"], r["pair_kernel_ms"]))
PY
done
