python -m pytest tests/test_bond_gpu.py tests/test_golden_gpu.py tests/test_parity_gpu.py -q -x 2>&1 | tail -4
HOOMD_B200_NLIST_WARPCELL=1 python -m pytest tests/test_golden_gpu.py tests/test_parity_gpu.py -q -x 2>&1 | tail -3
python bench.py --workload kg1m --no-scale-ref --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r1o_bench_kg1m.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python -c "
import json; d=json.load(open('gpurun_out/r1o_bench_kg1m.json')); print('kg1m', round(d['value']), d['roofline']['nlist_rebuild_ms'], round(d['e2e']['value']))"
