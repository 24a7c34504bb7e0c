python -m pytest tests/test_bond_gpu.py tests/test_golden_gpu.py -q -x 2>&1 | tail -4
for V in 1 0; do
HOOMD_B200_NLIST_WARPCELL=$V python bench.py --workload kg1m --no-scale-ref --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ab_wc$V.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python -c "
import json; d=json.load(open('gpurun_out/ab_wc$V.json')); print('warpcell=$V', round(d['value']), d['roofline']['nlist_rebuild_ms'], round(d['e2e']['value']))"
done
