set -x
python -m pytest tests -m gpu -q 2>&1 | tail -5
python bench.py > gpurun_out/bench_lj64k.json 2> gpurun_out/bench_lj64k.err; tail -3 gpurun_out/bench_lj64k.err; cat gpurun_out/bench_lj64k.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
HOOMD_B200_USE_GRAPH=0 python bench.py --steps 1 --warmup 1 --md-steps 50 --equil-steps 200 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
HOOMD_B200_USE_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -s 1200 -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 1 --warmup 1 --md-steps 50 --equil-steps 200 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/ncu1.log
HOOMD_B200_USE_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:k_pair_forces -s 300 -c 3 -o gpurun_out/prof_pair_r1 python bench.py --steps 1 --warmup 1 --md-steps 50 --equil-steps 200 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
