# 2-GPU validation of the domain-decomposed path: parity vs the reference classes in both ghost modes, then A/B timing
mkdir -p gpurun_out
N=${NGPU:-2}
tr() { port=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
tr 29511 tests/mg_check.py 24 300 > gpurun_out/r2e_mgcheck_fused.log 2>&1; echo "mg_check fused rc=$?"; grep -E "MG_CHECK_OK|mg parity|Error|error" gpurun_out/r2e_mgcheck_fused.log | cut -c1-600 | head -5
HOOMD_B200_FUSED_GHOSTS=0 tr 29512 tests/mg_check.py 24 300 > gpurun_out/r2e_mgcheck_r1.log 2>&1; echo "mg_check round-1 mode rc=$?"; grep -E "MG_CHECK_OK|mg parity|Error|error" gpurun_out/r2e_mgcheck_r1.log | cut -c1-600 | head -5
tr 29513 tests/mg_check.py 40 600 > gpurun_out/r2e_mgcheck_fused_64k.log 2>&1; echo "mg_check fused 64k rc=$?"; grep -E "MG_CHECK_OK|mg parity" gpurun_out/r2e_mgcheck_fused_64k.log | cut -c1-600 | head -3
for m in 1 3 5; do HOOMD_B200_MG_MODE=$m tr 2952$m tests/mg_check.py 24 300 > gpurun_out/r2e_mgcheck_mode$m.log 2>&1; echo "mg_check mode $m rc=$?"; grep -E "MG_CHECK_OK|RuntimeError|AssertionError" gpurun_out/r2e_mgcheck_mode$m.log | cut -c1-300 | head -3; done
for w in lj1m lj8m; do
  tr 29514 bench.py --gpus $N --workload $w --steps 8 --warmup 3 --no-scale-ref --no-parity > gpurun_out/r2e_${w}_${N}gpu_fused.json 2> gpurun_out/r2e_${w}_${N}gpu_fused.err; echo "$w fused rc=$?"
  HOOMD_B200_MG_MODE=1 tr 29516 bench.py --gpus $N --workload $w --steps 8 --warmup 3 --no-scale-ref --no-parity > gpurun_out/r2e_${w}_${N}gpu_push.json 2> gpurun_out/r2e_${w}_${N}gpu_push.err; echo "$w push-only rc=$?"
  HOOMD_B200_FUSED_GHOSTS=0 tr 29515 bench.py --gpus $N --workload $w --steps 8 --warmup 3 --no-scale-ref --no-parity > gpurun_out/r2e_${w}_${N}gpu_r1.json 2> gpurun_out/r2e_${w}_${N}gpu_r1.err; echo "$w round-1 mode rc=$?"
  python - <<PY
import json
for v in ("fused","push","r1"):
    try:
        d=json.loads(open("gpurun_out/r2e_${w}_${N}gpu_%s.json"%v).read().strip().splitlines()[-1])
        print("${w}", v, "TPS %.1f e2e %.1f kernel_ms %.4f rebuild_ms %.4f" % (d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["roofline"]["nlist_rebuild_ms"]), d.get("parity_checked"), d.get("max_force_err"), d.get("domain"))
    except Exception as e: print("${w}", v, "ERR", e)
PY
done
tail -n 5 gpurun_out/r2e_*err | cut -c1-300
