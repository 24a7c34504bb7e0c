for v in libhoomd_b200 libv_u1s libv_u2s libv_u4n; do
  echo "== $v"
  HOOMD_B200_LIB=$PWD/hoomd_b200/$v.so TPPS=1,2,4 python scratch/sweep.py lj1m 400 1500 2>&1 | grep '^{' | cut -c1-190
done
