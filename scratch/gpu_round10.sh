set -x
timeout 400 python -m pytest tests -m gpu -q -x -k "cell or nlist or engine or overflow or step" 2>&1 | tail -5
B="--no-cpu-baseline --no-scale-ref"
timeout 300 python bench.py $B > gpurun_out/r1p_lj64k_fused.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-300 gpurun_out/r1p_lj64k_fused.json
HOOMD_B200_CELL_FUSED=0 timeout 300 python bench.py $B > gpurun_out/r1p_lj64k_4k.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-300 gpurun_out/r1p_lj64k_4k.json
timeout 300 python bench.py $B --workload ka256k > gpurun_out/r1p_ka_fused.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-300 gpurun_out/r1p_ka_fused.json
HOOMD_B200_CELL_FUSED=0 timeout 300 python bench.py $B --workload ka256k > gpurun_out/r1p_ka_4k.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-300 gpurun_out/r1p_ka_4k.json
