# single GPU: NPT script test, 64k / 1M bench with the EXTRAS-free step kernel + 16-step graph, ncu captures
mkdir -p gpurun_out
python -m pytest tests/test_npt_gpu.py tests/test_group_gpu.py tests/test_parity_gpu.py -q -x > gpurun_out/r2k_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2k_tests.log | cut -c1-300
for w in lj64k lj1m; do
python bench.py --steps 8 --warmup 3 --workload $w --no-scale-ref --no-cpu-baseline > gpurun_out/r2k_$w.json 2> gpurun_out/r2k_$w.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2k_$w.json").read().strip().splitlines()[-1]); r=d["roofline"]
print("$w TPS %.1f e2e %.1f kernel %.4f rebuild %.4f cell %.4f nlist %.4f pair %.4f" % (d["value"], d["e2e"]["value"], r["kernel_ms"], r["nlist_rebuild_ms"], r["cell_list_ms"], r["nlist_kernel_ms"], r["pair_kernel_ms"]))
PY
done
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
$NCU -k regex:k_md_step -c 1 -o gpurun_out/prof_mdstep_r2k_64k python tools/prof_step.py lj64k 2 > gpurun_out/r2k_ncu1.log 2>&1; tail -1 gpurun_out/r2k_ncu1.log
$NCU -k regex:k_nlist_tiled -c 1 -o gpurun_out/prof_nlist_r2k_64k python tools/prof_step.py lj64k 1 > gpurun_out/r2k_ncu2.log 2>&1; tail -1 gpurun_out/r2k_ncu2.log
du -sh gpurun_out
