# N-GPU smoke of the domain-decomposed path: mg_check (one type NVE; Kremer-Grest chains) against the reference
mkdir -p gpurun_out
N=${NGPU:-2}
tr() { port=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
tr 29571 tests/mg_check.py 28 200 > gpurun_out/mgm_one_${N}gpu.log 2>&1; echo "one type rc=$?"; grep -E "mg parity|MG_CHECK_OK|Error" gpurun_out/mgm_one_${N}gpu.log | cut -c1-330 | head -3
tr 29572 tests/mg_check.py 24 200 2 > gpurun_out/mgm_two_${N}gpu.log 2>&1; echo "two types rc=$?"; grep -E "mg parity|MG_CHECK_OK|Error" gpurun_out/mgm_two_${N}gpu.log | cut -c1-330 | head -3
tr 29573 tests/mg_check.py 28 200 1 kg > gpurun_out/mgm_kg_${N}gpu.log 2>&1; echo "kg rc=$?"; grep -E "mg parity|MG_CHECK_OK|Error" gpurun_out/mgm_kg_${N}gpu.log | cut -c1-330 | head -3
