timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v29.log 2>&1; tail -5 gpurun_out/gpu_tests_v29.log
for cfg in "PIPER_B200_MRF_SPLIT=1" "PIPER_B200_MRF_SPLIT=0" "PIPER_B200_MRF_EG=3" "PIPER_B200_MRF_EG=5"; do
  env $cfg timeout 300 python bench.py --no-cpu --steps 5 > gpurun_out/bench_v29_tmp.json 2>gpurun_out/bench_v29.err; tail -2 gpurun_out/bench_v29.err
  echo "$cfg"; python -c "import json,sys; d=json.loads(open('gpurun_out/bench_v29_tmp.json').read()); print(d['value'], d['ms_per_step'], d['stage_ms']['decoder_ms'], d['roofline']['by_class']['mrf_tc_kernel'])"
done
