# diagnostic GPU call: full GPU suite, stage-kernel trace of config 3, round-1 hang repro under the mbarrier watchdog,
# compute-sanitizer synccheck / racecheck / memcheck on a short f16 synthesize
cd /root/repo
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/gpu_tests_diag1.log 2>&1; tail -3 gpurun_out/gpu_tests_diag1.log
T=piper-swift_b200/libpiper_b200_trace.so
PIPER_B200_LIB=$T PIPER_B200_D2H_CHUNKS=1 PIPER_B200_MRF_TRACE=gpurun_out/mrf_trace timeout 300 python tools/trace_config3.py > gpurun_out/trace_config3.log 2>&1; tail -2 gpurun_out/trace_config3.log
for c in 128 64 32; do python tools/mrf_trace.py gpurun_out/mrf_trace.$c 6 6 > gpurun_out/mrf_trace_$c.txt 2>&1; done
PIPER_B200_LIB=$T timeout 200 python tools/hang_repro.py 600 3 > gpurun_out/hang_repro_default.log 2>&1; tail -4 gpurun_out/hang_repro_default.log
PIPER_B200_LIB=$T PIPER_B200_MRF_UNIT_STREAM=1 timeout 200 python tools/hang_repro.py 600 4 > gpurun_out/hang_repro_stream.log 2>&1; echo "rc $?"; tail -12 gpurun_out/hang_repro_stream.log
for tool in synccheck memcheck racecheck; do
  timeout 400 compute-sanitizer --tool $tool --print-limit 20 python tools/hang_repro.py 40 1 > gpurun_out/sanitizer_$tool.log 2>&1; echo "$tool rc $?"; tail -4 gpurun_out/sanitizer_$tool.log
done
