# 2-GPU: parity vs the reference classes (raw ghost coordinates, global minimum image), then the MG sub-mode A/B
mkdir -p gpurun_out
N=${NGPU:-2}
tr() { port=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
for m in 7 0; do HOOMD_B200_MG_MODE=$m tr 2951$m tests/mg_check.py 24 300 > gpurun_out/r2f_mgcheck_mode$m.log 2>&1; echo "mg_check mode $m rc=$?"; grep -E "MG_CHECK_OK|mg parity|RuntimeError|AssertionError" gpurun_out/r2f_mgcheck_mode$m.log | cut -c1-700 | head -3; done
tr 29513 tests/mg_check.py 40 200 > gpurun_out/r2f_mgcheck_64k.log 2>&1; echo "mg_check 64k rc=$?"; grep -E "MG_CHECK_OK|mg parity|RuntimeError|AssertionError" gpurun_out/r2f_mgcheck_64k.log | cut -c1-700 | head -3
for w in lj1m lj8m; do
  for m in 7 5 3 1 0; do
    if [ $w = lj8m ] && [ $m != 7 ] && [ $m != 0 ] && [ $m != 1 ]; then continue; fi
    P=""; if [ $m != 7 ]; then P="--no-parity"; fi
    HOOMD_B200_MG_MODE=$m tr 2952$m bench.py --gpus $N --workload $w --steps 8 --warmup 3 --no-scale-ref $P > gpurun_out/r2f_${w}_${N}gpu_mode$m.json 2> gpurun_out/r2f_${w}_${N}gpu_mode$m.err; echo "$w mode $m rc=$?"
  done
done
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r2f_*gpu_mode*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        bw=d["domain"]["barrier_wait"]
        print(f.split("/")[-1], "TPS %.1f e2e %.1f kernel_ms %.4f rebuild_ms %.4f step wait %.1f us, rebuild waits %.1f/%.1f/%.1f us, boundary chunks %d/%d" % (d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms"], d["roofline"]["nlist_rebuild_ms"], bw["step"]["mean_us"], bw["migrate"]["mean_us"], bw["ghost_plan"]["mean_us"], bw["overflow_vote"]["mean_us"], bw["boundary_chunks"], bw["chunks"]), d.get("parity_checked"), d.get("max_force_err"))
    except Exception as e: print(f, "ERR", e)
PY
grep -h "Error" gpurun_out/r2f_*err | cut -c1-300 | sort | uniq -c | head
