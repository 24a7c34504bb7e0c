# N-GPU: mg_check (one type NVE; two types Langevin; Morse), then the default bench line (parity pass + strong_scaling)
mkdir -p gpurun_out
N=${NGPU:-2}
tr() { port=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
tr 29561 tests/mg_check.py 28 200 > gpurun_out/mgq_one_${N}gpu.log 2>&1; echo "one type rc=$?"; grep -E "MG_CHECK_OK|Error" gpurun_out/mgq_one_${N}gpu.log | cut -c1-300 | head -3
tr 29562 tests/mg_check.py 28 200 2 > gpurun_out/mgq_two_${N}gpu.log 2>&1; echo "two types rc=$?"; grep -E "MG_CHECK_OK|Error" gpurun_out/mgq_two_${N}gpu.log | cut -c1-300 | head -3
tr 29563 tests/mg_check.py 28 150 1 morse > gpurun_out/mgq_morse_${N}gpu.log 2>&1; echo "morse rc=$?"; grep -E "MG_CHECK_OK|mg parity|Error" gpurun_out/mgq_morse_${N}gpu.log | cut -c1-400 | head -4
tr 29564 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/mgq_lj8m_${N}gpu.json 2> gpurun_out/mgq_lj8m_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/mgq_lj8m_${N}gpu.json").read().strip().splitlines()[-1])
    print("N=$N TPS %.1f e2e %.1f parity %s %s strong %s" % (d["value"], d["e2e"]["value"], d.get("parity_checked"), d.get("max_force_err"), json.dumps(d.get("strong_scaling"))[:300]))
except Exception as e: print("ERR", e)
PY
tail -3 gpurun_out/mgq_lj8m_${N}gpu.err | cut -c1-300
