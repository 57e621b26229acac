python -m pytest tests/test_gpu_pool.py -m gpu -q > gpurun_out/r02_gpu_pool_n2.log 2>&1; tail -6 gpurun_out/r02_gpu_pool_n2.log
