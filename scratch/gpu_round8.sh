set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -6
python bench.py --workload ka256k --no-scale-ref --steps 3 --no-cpu-baseline > gpurun_out/r1l_bench_ka256k.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cat gpurun_out/r1l_bench_ka256k.json
