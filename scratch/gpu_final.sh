set -x
python -m pytest tests -m gpu -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r1n_bench_lj64k.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-400 gpurun_out/r1n_bench_lj64k.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1n_bench_ref.json 2>gpurun_out/err.log; cut -c1-200 gpurun_out/r1n_bench_ref.json
HOOMD_B200_USE_GRAPH=3 ncu --metrics gpu__time_duration.sum --clock-control none -s 5000 -c 400 --csv --log-file gpurun_out/launches_r1n.csv python bench.py --steps 2 --warmup 3 --md-steps 100 --equil-steps 400 --no-cpu-baseline --no-scale-ref > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log | cut -c1-200
HOOMD_B200_USE_GRAPH=3 ncu --set full --clock-control none --import-source on -k regex:k_md_step -s 100 -c 2 -o gpurun_out/prof_mdstep_r1n_64k python scratch/prof_tile.py lj64k > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
HOOMD_B200_USE_GRAPH=3 ncu --set full --clock-control none --import-source on -k regex:k_md_step -s 100 -c 2 -o gpurun_out/prof_mdstep_r1n_1m python scratch/prof_tile.py lj1m > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu3.log
