# full captures of the cell-tile step kernel and the tile list build at 1M
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
HOOMD_B200_PAIR_MODE=1 timeout 600 $NCU -k regex:k_md_step_tile -c 1 -o gpurun_out/prof_tile_step_1m python tools/prof_step.py lj1m 2 300 > gpurun_out/tile_ncu3.log 2>&1; tail -2 gpurun_out/tile_ncu3.log | cut -c1-300
HOOMD_B200_PAIR_MODE=1 timeout 600 $NCU -k regex:k_tile_build -c 1 -o gpurun_out/prof_tile_build_1m python tools/prof_step.py lj1m 1 300 > gpurun_out/tile_ncu4.log 2>&1; tail -2 gpurun_out/tile_ncu4.log | cut -c1-300
