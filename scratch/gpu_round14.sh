set -x
timeout 600 python -m pytest tests/test_nvt_gpu.py tests/test_script.py tests/test_golden_gpu.py -m gpu -q -x 2>&1 | tail -12
B="--no-cpu-baseline --no-scale-ref"
run() { n=$1; shift; env "$@" timeout 300 python bench.py $B $W > gpurun_out/r1q_$n.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log; cut -c1-90 gpurun_out/r1q_$n.json; }
W=""
run lj64k_nve A=1
run lj64k_nvt HOOMD_B200_BENCH_METHOD=nvt
W="--workload lj1m"
run lj1m_nve A=1
run lj1m_nvt HOOMD_B200_BENCH_METHOD=nvt
