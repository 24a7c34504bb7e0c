# A/B of two builds of the library on the same box: vector vs scalar index loads in the CSR pair walk
for lib in hoomd_b200/libhoomd_b200.so hoomd_b200/libhoomd_b200_novec.so; do
  for wl in lj64k lj1m; do
    HOOMD_B200_LIB=$PWD/$lib HOOMD_B200_PAIR_MODE=0 TPPS=${TPPS:-2,4} timeout 300 python scratch/sweep.py $wl 2000 > gpurun_out/ab.jsonl 2> gpurun_out/ab.err; echo "$lib $wl rc=$?"; cut -c1-200 gpurun_out/ab.jsonl
  done
done
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/gpu_tests.log
