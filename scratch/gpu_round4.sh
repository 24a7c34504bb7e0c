set -x
python -m pytest tests/test_bond_gpu.py tests/test_script.py -m gpu -q 2>&1 | tail -8
python bench.py --workload kg1m --no-scale-ref --steps 3 > gpurun_out/r1g_bench_kg1m.json 2> gpurun_out/r1g_bench_kg1m.err; tail -5 gpurun_out/r1g_bench_kg1m.err; cat gpurun_out/r1g_bench_kg1m.json
python bench.py --impl reference --workload kg1m --steps 2 --warmup 1 --ref-md-steps 5 > gpurun_out/r1g_bench_kg1m_ref.json 2>gpurun_out/r1g_ref.err; tail -3 gpurun_out/r1g_ref.err; cat gpurun_out/r1g_bench_kg1m_ref.json
# launch list of the timed region, stream-launch mode (64k LJ): skip equilibration + warm-up launches
HOOMD_B200_USE_GRAPH=3 ncu --metrics gpu__time_duration.sum --clock-control none -s 4000 -c 400 --csv --log-file gpurun_out/launches_r1g.csv python bench.py --steps 2 --warmup 3 --md-steps 50 --equil-steps 200 --no-cpu-baseline --no-scale-ref > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log
HOOMD_B200_USE_GRAPH=3 ncu --metrics gpu__time_duration.sum --clock-control none -s 2500 -c 300 --csv --log-file gpurun_out/launches_r1g_kg1m.csv python bench.py --workload kg1m --steps 1 --warmup 3 --md-steps 30 --equil-steps 100 --no-cpu-baseline --no-scale-ref > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
