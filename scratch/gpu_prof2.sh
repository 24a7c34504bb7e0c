set -x
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
$NCU -k regex:k_md_step -c 1 -o gpurun_out/prof_mdstep_r1n_64k python scratch/prof_step.py lj64k 2 > gpurun_out/ncu2.log 2>&1; tail -1 gpurun_out/ncu2.log
$NCU -k regex:k_md_step -c 1 -o gpurun_out/prof_mdstep_r1n_1m python scratch/prof_step.py lj1m 2 > gpurun_out/ncu3.log 2>&1; tail -1 gpurun_out/ncu3.log
$NCU -k regex:"k_nlist_warpcell" -c 1 -o gpurun_out/prof_warpcell_r1n_kg1m python scratch/prof_step.py kg1m 1 > gpurun_out/ncu4.log 2>&1; tail -1 gpurun_out/ncu4.log
$NCU -k regex:"k_bond_forces|k_md_step" -c 2 -o gpurun_out/prof_mdstep_r1n_kg1m python scratch/prof_step.py kg1m 2 > gpurun_out/ncu5.log 2>&1; tail -1 gpurun_out/ncu5.log
du -sh gpurun_out
