NP=${NP:-2}
timeout ${TMO:-240} python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NP --steps ${STEPS:-3} --warmup ${WARM:-3} --workload ${WL:-lj1m} ${EXTRA} > gpurun_out/mg_bench_${WL:-lj1m}_$NP.json 2> gpurun_out/mg_bench.err
echo rc=$?
grep -v "^W1\|^\[W\|warnings.warn" gpurun_out/mg_bench.err | grep -A12 "Traceback\|Error" | head -40
cat gpurun_out/mg_bench_${WL:-lj1m}_$NP.json
