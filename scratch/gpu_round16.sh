B="--no-cpu-baseline --no-scale-ref"
for w in lj64k lj1m; do
for v in "" _v12 _v8 ""; do
HOOMD_B200_LIB=$PWD/hoomd_b200/libhoomd_b200$v.so timeout 300 python bench.py $B --workload $w > gpurun_out/r1r_${w}${v}.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python - <<PY
import json; d=json.load(open("gpurun_out/r1r_${w}${v}.json")); print("$w", "$v", round(d["value"]), d["roofline"]["frac"])
PY
done; done
