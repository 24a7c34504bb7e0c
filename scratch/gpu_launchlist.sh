HOOMD_B200_USE_GRAPH=3 ncu --metrics gpu__time_duration.sum --clock-control none -s 26000 -c 400 --csv --log-file gpurun_out/launches_r1s.csv python bench.py --steps 2 --warmup 3 --md-steps 100 --equil-steps 4000 --no-cpu-baseline --no-scale-ref > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log | cut -c1-200
wc -l gpurun_out/launches_r1s.csv
