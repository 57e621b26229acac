timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v38.log 2>&1; tail -3 gpurun_out/gpu_tests_v38.log
