mkdir -p gpurun_out
N=${NGPU:-2}
tr() { port=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
tr 29581 tests/mg_check.py 24 200 1 kg > gpurun_out/mgk_kg_${N}gpu.log 2>&1; echo "kg rc=$?"; grep -E "MG_CHECK_OK|mg parity|Error|assert" gpurun_out/mgk_kg_${N}gpu.log | cut -c1-1200 | head -8
tail -4 gpurun_out/mgk_kg_${N}gpu.log | cut -c1-300
