timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6
timeout 300 python bench.py --no-cpu-baseline --no-scale-ref > gpurun_out/r1v_lj64k.json 2> gpurun_out/err.log; tail -2 gpurun_out/err.log; cut -c1-120 gpurun_out/r1v_lj64k.json
