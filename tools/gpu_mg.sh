# N-GPU (default 8): parity pass inside bench (64k vs the reference classes), lj8m in the three ghost/barrier modes
mkdir -p gpurun_out
N=${NGPU:-8}
tr() { port=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
tr 29541 tests/mg_check.py 40 200 > gpurun_out/mg_mgcheck_${N}gpu.log 2>&1; echo "mg_check rc=$?"; grep -E "MG_CHECK_OK|mg parity|RuntimeError|AssertionError" gpurun_out/mg_mgcheck_${N}gpu.log | cut -c1-900 | head -3
tr 29547 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/mg_lj8m_${N}gpu_default.json 2> gpurun_out/mg_lj8m_${N}gpu_default.err; echo "default rc=$?"
HOOMD_B200_MG_MODE=7 tr 29548 bench.py --gpus $N --steps 10 --warmup 3 --no-scale-ref --no-parity > gpurun_out/mg_lj8m_${N}gpu_mode7.json 2> gpurun_out/mg_lj8m_${N}gpu_mode7.err; echo "mode 1 rc=$?"
HOOMD_B200_MG_MODE=0 tr 29549 bench.py --gpus $N --steps 10 --warmup 3 --no-scale-ref --no-parity > gpurun_out/mg_lj8m_${N}gpu_mode0.json 2> gpurun_out/mg_lj8m_${N}gpu_mode0.err; echo "mode 0 rc=$?"
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/mg_*gpu_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        bw=d["domain"]["barrier_wait"]
        print(f.split("/")[-1], "TPS %.1f e2e %.1f kernel_ms %.4f rebuild_ms %.4f step wait %.1f us, rebuild waits %.1f/%.1f/%.1f us, boundary chunks %d/%d" % (d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["roofline"]["nlist_rebuild_ms"], bw["step"]["mean_us"], bw["migrate"]["mean_us"], bw["ghost_plan"]["mean_us"], bw["overflow_vote"]["mean_us"], bw["boundary_chunks"], bw["chunks"]), d.get("parity_checked"), d.get("max_force_err"), d.get("strong_scaling"))
    except Exception as e: print(f, "ERR", e)
PY
grep -h "Error" gpurun_out/mg_*err | cut -c1-400 | sort | uniq -c | head
