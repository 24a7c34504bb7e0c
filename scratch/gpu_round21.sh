timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
B="--no-cpu-baseline --no-scale-ref"
for w in lj64k lj1m ka256k; do
for c in 1 0; do
HOOMD_B200_NLIST_CHUNKED=$c timeout 300 python bench.py $B --workload $w > gpurun_out/r1u_${w}_chunk$c.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python - <<PY
import json; d=json.load(open("gpurun_out/r1u_${w}_chunk$c.json")); r=d["roofline"]; print("$w chunked=$c", round(d["value"]), round(d["e2e"]["value"]), r["nlist_rebuild_ms"])
PY
done; done
