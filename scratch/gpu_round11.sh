set -x
HOOMD_B200_PDL=1 timeout 400 python -m pytest tests -m gpu -q -x -k "cell or nlist or engine or overflow or step" 2>&1 | tail -5
B="--no-cpu-baseline --no-scale-ref"
for w in lj64k ka256k lj1m; do
HOOMD_B200_PDL=1 timeout 300 python bench.py $B --workload $w > gpurun_out/r1p_${w}_pdl.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-120 gpurun_out/r1p_${w}_pdl.json
timeout 300 python bench.py $B --workload $w > gpurun_out/r1p_${w}_nopdl.json 2> gpurun_out/err.log; tail -3 gpurun_out/err.log; cut -c1-120 gpurun_out/r1p_${w}_nopdl.json
done
