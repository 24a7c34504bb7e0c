# N-GPU (default 8): mg_check in every variant against the reference classes, then the default bench line
mkdir -p gpurun_out
N=${NGPU:-8}
tr() { port=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
run() { name=$1; shift; tr $1 tests/mg_check.py "${@:2}" > gpurun_out/mgf_${name}_${N}gpu.log 2>&1; echo "$name rc=$?"; grep -E "mg parity|MG_CHECK_OK|Error" gpurun_out/mgf_${name}_${N}gpu.log | cut -c1-420 | head -3; }
run one 29601 40 200
run two 29602 32 200 2
run morse 29603 32 150 1 morse
run nvt 29604 32 200 1 lj nvt
run kg 29605 32 200 1 kg
tr 29606 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/mgf_lj8m_${N}gpu.json 2> gpurun_out/mgf_lj8m_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/mgf_lj8m_${N}gpu.json").read().strip().splitlines()[-1])
    bw=d["domain"]["barrier_wait"]
    print("N=$N TPS %.1f e2e %.1f kernel %.4f rebuild %.4f step wait %.1f parity %s %s strong %s" % (d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["roofline"]["nlist_rebuild_ms"], bw["step"]["mean_us"], d.get("parity_checked"), d.get("max_force_err"), json.dumps(d.get("strong_scaling"))[:260]))
except Exception as e: print("ERR", e)
PY
