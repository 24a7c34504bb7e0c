set -x
for WL in lj1m kg1m; do
HOOMD_B200_USE_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 300 --csv --log-file gpurun_out/launches_r1h_host_$WL.csv python bench.py --workload $WL --steps 1 --warmup 3 --md-steps 30 --equil-steps 100 --no-cpu-baseline --no-scale-ref > gpurun_out/ncu_$WL.log 2>&1
tail -1 gpurun_out/ncu_$WL.log | cut -c1-300
done
