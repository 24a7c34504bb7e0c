mkdir -p gpurun_out
for r in 1 2 3; do
timeout 600 python bench.py --steps 8 --warmup 3 --workload ka256k --no-scale-ref --no-cpu-baseline > gpurun_out/ka_rep$r.json 2> gpurun_out/ka_rep$r.err; echo "rep $r rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/ka_rep$r.json").read().strip().splitlines()[-1])
print(round(d["value"],1), [round(x,1) for x in d["ms_runs"]], [round(x,1) for x in d["e2e"]["ms_runs"]], d["engine"]["Nmax"], d["engine"]["rebuilds_total"])
PY
done
