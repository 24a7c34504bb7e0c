set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -6
python bench.py --workload kg1m --no-scale-ref --steps 3 --no-cpu-baseline > gpurun_out/r1i_bench_kg1m.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cat gpurun_out/r1i_bench_kg1m.json
HOOMD_B200_TAG_TABLES=1 python bench.py --workload kg1m --no-scale-ref --steps 3 --no-cpu-baseline > gpurun_out/r1i_bench_kg1m_tagtables.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cat gpurun_out/r1i_bench_kg1m_tagtables.json
python bench.py --workload lj1m --no-scale-ref --steps 3 --no-cpu-baseline > gpurun_out/r1i_bench_lj1m.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cat gpurun_out/r1i_bench_lj1m.json
