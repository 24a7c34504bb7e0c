set -x
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r1s_bench_lj64k.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-300 gpurun_out/r1s_bench_lj64k.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1s_bench_ref.json 2>gpurun_out/err.log; cut -c1-200 gpurun_out/r1s_bench_ref.json
HOOMD_B200_USE_GRAPH=3 ncu --metrics gpu__time_duration.sum --clock-control none -s 5000 -c 400 --csv --log-file gpurun_out/launches_r1s.csv python bench.py --steps 2 --warmup 3 --md-steps 100 --equil-steps 400 --no-cpu-baseline --no-scale-ref > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log | cut -c1-200
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
HOOMD_B200_USE_GRAPH=3 $NCU -k regex:k_md_step -c 1 -o gpurun_out/prof_mdstep_r1s_64k python scratch/prof_step.py lj64k 2 > gpurun_out/ncu2.log 2>&1; tail -1 gpurun_out/ncu2.log
HOOMD_B200_USE_GRAPH=3 $NCU -k regex:k_md_step -c 1 -o gpurun_out/prof_mdstep_r1s_1m python scratch/prof_step.py lj1m 2 > gpurun_out/ncu3.log 2>&1; tail -1 gpurun_out/ncu3.log
ls -la gpurun_out/*.ncu-rep
