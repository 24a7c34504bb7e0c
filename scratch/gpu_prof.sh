CMD="python scratch/prof_tile.py ${WL:-lj64k}"
HOOMD_B200_PAIR_MODE=0 ncu --set full --clock-control none --import-source on -k regex:"k_nlist_cellscan" -s 1 -c 1 -o gpurun_out/prof_cellscan_${WL:-lj64k} $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
