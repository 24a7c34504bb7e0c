mkdir -p gpurun_out
N=${NGPU:-2}
tr() { port=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
tr 29571 tests/mg_check.py 24 200 1 lj nvt > gpurun_out/mgn_nvt_${N}gpu.log 2>&1; echo "nvt rc=$?"; grep -E "MG_CHECK_OK|mg parity|Error|assert" gpurun_out/mgn_nvt_${N}gpu.log | cut -c1-1500 | head -6
tr 29572 tests/mg_check.py 24 100 > gpurun_out/mgn_nve_${N}gpu.log 2>&1; echo "nve rc=$?"; grep -E "MG_CHECK_OK|Error" gpurun_out/mgn_nve_${N}gpu.log | cut -c1-300 | head -3
