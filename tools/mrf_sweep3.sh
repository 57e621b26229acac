cd /root/repo
run() {
  env "$@" PIPER_B200_PROF_DUMP=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/prof_sweep.txt > /dev/null
  echo "$@ :" $(python tools/prof_launches.py --digest gpurun_out/prof_sweep.txt 2>/dev/null | head -1 | cut -c1-150)
}
run X=1
run PIPER_B200_MRF_EG=2
run PIPER_B200_MRF_EG=4
run X=1
timeout 400 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -x -q --timeout 240 2>&1 | tail -2
timeout 240 python bench.py --no-cpu --no-extras > gpurun_out/bench_eg.json 2> gpurun_out/bench_eg.err; python tools/bench_brief.py gpurun_out/bench_eg.json 2>/dev/null | head -4
