set -x
python bench.py --workload ka256k --no-scale-ref --steps 3 > gpurun_out/r1k_bench_ka256k.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cat gpurun_out/r1k_bench_ka256k.json
