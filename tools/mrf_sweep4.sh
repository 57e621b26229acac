cd /root/repo
run() {
  env "$@" PIPER_B200_PROF_DUMP=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/prof_sweep.txt > /dev/null
  echo "$@ :" $(python tools/prof_launches.py --digest gpurun_out/prof_sweep.txt 2>/dev/null | head -1 | cut -c1-150)
}
run X=1
run PIPER_B200_MRF_EG=2
run PIPER_B200_MRF_EG=3
run X=1
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q --timeout 240 -k "oracle or batch or ragged or long_form" 2>&1 | tail -2
