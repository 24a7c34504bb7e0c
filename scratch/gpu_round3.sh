set -x
python -m pytest tests -m gpu -q 2>&1 | tail -8
python bench.py > gpurun_out/r1f_bench_lj64k.json 2> gpurun_out/r1f_bench_lj64k.err; tail -3 gpurun_out/r1f_bench_lj64k.err; cat gpurun_out/r1f_bench_lj64k.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1f_bench_ref.json 2>gpurun_out/r1f_bench_ref.err; cat gpurun_out/r1f_bench_ref.json
python bench.py --workload lj1m --no-cpu-baseline --no-scale-ref > gpurun_out/r1f_bench_lj1m.json 2> gpurun_out/r1f_bench_lj1m.err; cat gpurun_out/r1f_bench_lj1m.json
# launch list of the bench command in stream-launch mode (graph kernels launched on the stream so ncu sees them)
HOOMD_B200_USE_GRAPH=3 python bench.py --steps 1 --warmup 3 --md-steps 50 --equil-steps 200 --no-cpu-baseline --no-scale-ref > gpurun_out/plain.log 2>&1 && \
HOOMD_B200_USE_GRAPH=3 ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 400 --csv --log-file gpurun_out/launches_r1f.csv python bench.py --steps 1 --warmup 3 --md-steps 50 --equil-steps 200 --no-cpu-baseline --no-scale-ref > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/ncu1.log
HOOMD_B200_USE_GRAPH=3 ncu --set full --clock-control none --import-source on -k regex:k_md_step -s 100 -c 2 -o gpurun_out/prof_mdstep_r1f_64k python scratch/prof_tile.py lj64k > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
HOOMD_B200_USE_GRAPH=3 ncu --set full --clock-control none --import-source on -k regex:k_md_step -s 100 -c 2 -o gpurun_out/prof_mdstep_r1f_1m python scratch/prof_tile.py lj1m > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/ncu3.log
