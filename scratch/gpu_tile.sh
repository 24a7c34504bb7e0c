timeout 600 python -m pytest tests/test_tile_gpu.py -x -q -m gpu > gpurun_out/tile_tests.log 2>&1; echo "tile tests rc=$?"; tail -5 gpurun_out/tile_tests.log
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_tile_gpu.py > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/gpu_tests.log
for wl in lj64k lj1m; do
  HOOMD_B200_PAIR_MODE=0 TPPS=2 timeout 300 python scratch/sweep.py $wl 2000 > gpurun_out/sweep_${wl}_pm0.jsonl 2> gpurun_out/sweep.err; echo "sweep $wl pm=0 rc=$?"; cat gpurun_out/sweep_${wl}_pm0.jsonl | cut -c1-200
  for tile in 4,2,2 2,2,2 3,2,2; do
    HOOMD_B200_PAIR_MODE=1 TILE=$tile TPPS=1,2,4 timeout 300 python scratch/sweep.py $wl 2000 > gpurun_out/sweep_${wl}_pm1_$tile.jsonl 2> gpurun_out/sweep.err; echo "sweep $wl pm=1 tile=$tile rc=$?"; cat gpurun_out/sweep_${wl}_pm1_$tile.jsonl | cut -c1-200
  done
done
HOOMD_B200_PAIR_MODE=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_tile64k.csv python scratch/prof_tile.py lj64k > gpurun_out/ncu3.log 2>&1; tail -2 gpurun_out/ncu3.log
