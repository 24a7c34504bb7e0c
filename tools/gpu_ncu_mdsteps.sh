# full capture of one 40-step k_md_steps launch at 64k (final tree)
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
timeout 600 $NCU -k regex:k_md_steps -c 1 -o gpurun_out/prof_mdsteps_r2z_64k python tools/prof_step.py lj64k 5 > gpurun_out/ncu_mdsteps.log 2>&1; tail -2 gpurun_out/ncu_mdsteps.log | cut -c1-300
