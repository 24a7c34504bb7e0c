mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dpd_gpu.py tests/test_pair_evaluators_gpu.py tests/test_abi.py -q -x > gpurun_out/dpd_tests.log 2>&1; echo "tests rc=$?"; tail -25 gpurun_out/dpd_tests.log | cut -c1-300
