cd /root/repo
timeout 400 python -m pytest tests -m gpu -x -q -k "ragged or full_size" > gpurun_out/gpu_tests_v49.log 2>&1; tail -5 gpurun_out/gpu_tests_v49.log
