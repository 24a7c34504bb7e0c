for V in NONE HOOMD_B200_NLIST_WARP HOOMD_B200_NLIST_CELLSCAN; do
env $V=1 python bench.py --workload kg1m --no-scale-ref --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$V.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python -c "
import json; d=json.load(open('gpurun_out/ab_$V.json')); print('$V', round(d['value']), d['roofline']['nlist_rebuild_ms'])"
done
