set -x
python -m pytest tests -m gpu -q 2>&1 | tail -15
python bench.py --no-cpu-baseline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; tail -3 gpurun_out/bench_a.err; cat gpurun_out/bench_a.json
python bench.py --no-cpu-baseline --no-flush > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; tail -3 gpurun_out/bench_b.err; cat gpurun_out/bench_b.json
