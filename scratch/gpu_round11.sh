for WL in lj64k lj1m; do for V in 1 0; do
HOOMD_B200_NLIST_WARPCELL=$V python bench.py --workload $WL --no-scale-ref --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$WL$V.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python -c "
import json; d=json.load(open('gpurun_out/ab_$WL$V.json')); print('$WL warpcell=$V', round(d['value']), d['roofline']['nlist_rebuild_ms'], round(d['e2e']['value']))"
done; done
