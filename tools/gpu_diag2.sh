cd /root/repo
T=piper-swift_b200/libpiper_b200_trace.so
PIPER_B200_LIB=$T PIPER_B200_MRF_UNIT_STREAM=0 timeout 200 python tools/hang_repro.py 600 3 > gpurun_out/hang_repro_default3.log 2>&1; tail -3 gpurun_out/hang_repro_default3.log
PIPER_B200_LIB=$T PIPER_B200_MRF_UNIT_STREAM=1 timeout 200 python tools/hang_repro.py 600 4 > gpurun_out/hang_repro_stream3.log 2>&1; echo "rc $?"; grep -v "^mrf_tc_kernel" gpurun_out/hang_repro_stream3.log | tail -8
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_diag3.log 2>&1; tail -3 gpurun_out/gpu_tests_diag3.log
timeout 500 python bench.py > gpurun_out/bench_diag3.json 2> gpurun_out/bench_diag3.err; python tools/bench_brief.py gpurun_out/bench_diag3.json
PIPER_B200_MRF_POST_TC=0 timeout 300 python bench.py --no-cpu > gpurun_out/bench_diag3_nopost.json 2>> gpurun_out/bench_diag3.err; python tools/bench_brief.py gpurun_out/bench_diag3_nopost.json
PIPER_B200_LIB=$T PIPER_B200_D2H_CHUNKS=1 PIPER_B200_MRF_TRACE=gpurun_out/mrf_trace3 timeout 300 python tools/trace_config3.py > gpurun_out/trace_config3_3.log 2>&1; tail -1 gpurun_out/trace_config3_3.log
for c in 128 64 32; do python tools/mrf_trace.py gpurun_out/mrf_trace3.$c 6 6 > gpurun_out/mrf_trace3_$c.txt 2>&1; done
