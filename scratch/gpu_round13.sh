set -x
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -6
B="--no-cpu-baseline --no-scale-ref"
run() { # name, env...
  n=$1; shift
  env "$@" timeout 300 python bench.py $B $W > gpurun_out/r1p_$n.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log; cut -c1-90 gpurun_out/r1p_$n.json
}
W=""
run lj64k_default A=1
run lj64k_cellfused HOOMD_B200_CELL_FUSED=1
run lj64k_pdl HOOMD_B200_PDL=1
run lj64k_default2 A=1
run lj64k_nvt HOOMD_B200_BENCH_METHOD=nvt
W="--workload ka256k"
run ka256k_default A=1
run ka256k_cellfused HOOMD_B200_CELL_FUSED=1
run ka256k_pdl HOOMD_B200_PDL=1
W="--workload lj1m"
run lj1m_default A=1
run lj1m_nvt HOOMD_B200_BENCH_METHOD=nvt
