timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v30.log 2>&1; tail -5 gpurun_out/gpu_tests_v30.log
for cfg in "PIPER_B200_MRF_POSTGROUP=1" "PIPER_B200_MRF_POSTGROUP=0" "PIPER_B200_MRF_ORDER=1" "PIPER_B200_MRF_ORDER=1 PIPER_B200_MRF_POSTGROUP=0"; do
  env $cfg timeout 300 python bench.py --no-cpu --steps 5 > gpurun_out/bench_v30_tmp.json 2>gpurun_out/bench_v30.err; tail -2 gpurun_out/bench_v30.err
  echo "$cfg"; python -c "import json,sys; d=json.loads(open('gpurun_out/bench_v30_tmp.json').read()); print(d['value'], d['ms_per_step'], d['stage_ms']['decoder_ms'], d['roofline']['by_class']['mrf_tc_kernel'])"
done
