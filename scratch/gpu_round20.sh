timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6
B="--no-cpu-baseline --no-scale-ref"
for w in lj64k lj1m ka256k; do
timeout 300 python bench.py $B --workload $w > gpurun_out/r1t_${w}.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log
python - <<PY
import json; d=json.load(open("gpurun_out/r1t_${w}.json")); r=d["roofline"]; print("$w", round(d["value"]), round(d["e2e"]["value"]), r["frac"], r["nlist_rebuild_ms"], r.get("traffic"))
PY
done
