# N-GPU (default 2): mg_check with one type (NVE) and with a two-type mixture under Langevin
mkdir -p gpurun_out
N=${NGPU:-2}
tr() { port=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
tr 29551 tests/mg_check.py 24 200 > gpurun_out/mgt_one_${N}gpu.log 2>&1; echo "one type rc=$?"; grep -E "MG_CHECK_OK|mg parity|Error" gpurun_out/mgt_one_${N}gpu.log | cut -c1-700 | head -4
tr 29552 tests/mg_check.py 24 200 2 > gpurun_out/mgt_two_${N}gpu.log 2>&1; echo "two types rc=$?"; grep -E "MG_CHECK_OK|mg parity|Error" gpurun_out/mgt_two_${N}gpu.log | cut -c1-700 | head -6
tail -5 gpurun_out/mgt_two_${N}gpu.log | cut -c1-400
