CMD="python scratch/prof_tile.py ${WL:-lj64k}"
HOOMD_B200_PAIR_MODE=1 HOOMD_B200_TILE=4,2,2,2 ncu --set full --clock-control none --import-source on -k regex:"k_tile_build" -s 1 -c 1 -o gpurun_out/prof_tilebuild_${WL:-lj64k} $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
