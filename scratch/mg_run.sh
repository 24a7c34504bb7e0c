timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NP:-2} --master-addr 127.0.0.1 --master-port 29511 tests/mg_check.py ${NS:-24} ${STEPS:-300} > gpurun_out/mg.log 2>&1
echo rc=$?
grep -v "^W1\|^\[W\|warnings.warn" gpurun_out/mg.log | grep -B2 -A25 "Traceback\|Error\|MG_CHECK\|N=" | head -80
