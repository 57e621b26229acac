python tools/latency_probe.py f16 > gpurun_out/latency_v35_f16.log 2>&1; cat gpurun_out/latency_v35_f16.log
python tools/latency_probe.py fp32 > gpurun_out/latency_v35_fp32.log 2>&1; cat gpurun_out/latency_v35_fp32.log
