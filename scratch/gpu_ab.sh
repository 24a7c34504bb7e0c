timeout 900 python -m pytest tests/test_golden_gpu.py tests/test_parity_gpu.py -x -q -m gpu -k "not drift" > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/gpu_tests.log
for wl in lj64k lj1m; do
  HOOMD_B200_PAIR_MODE=0 TPPS=2 timeout 300 python scratch/sweep.py $wl 1000 > gpurun_out/ab.jsonl 2> gpurun_out/ab.err; echo "cellscan-bulk $wl rc=$?"; cut -c100-200 gpurun_out/ab.jsonl
  HOOMD_B200_NC_MANUAL=1 HOOMD_B200_PAIR_MODE=0 TPPS=2 timeout 300 python scratch/sweep.py $wl 1000 > gpurun_out/ab.jsonl 2> gpurun_out/ab.err; echo "cellscan-manual $wl rc=$?"; cut -c100-200 gpurun_out/ab.jsonl
  HOOMD_B200_NLIST_TILED=1 HOOMD_B200_PAIR_MODE=0 TPPS=2 timeout 300 python scratch/sweep.py $wl 1000 > gpurun_out/ab.jsonl 2> gpurun_out/ab.err; echo "tiled $wl rc=$?"; cut -c100-200 gpurun_out/ab.jsonl
done
