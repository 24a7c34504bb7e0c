CMD="python bench.py --steps 1 --warmup 1 --md-steps 60 --equil-steps 300 --no-cpu-baseline --no-scale-ref"
HOOMD_B200_USE_GRAPH=0 $CMD > gpurun_out/plain.log 2>&1 && \
HOOMD_B200_USE_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:k_nlist_tiled -s 20 -c 1 -o gpurun_out/prof_nlist_r1g $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
