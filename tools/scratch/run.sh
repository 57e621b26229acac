timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v31.log 2>&1; tail -5 gpurun_out/gpu_tests_v31.log
timeout 300 python bench.py --no-cpu > gpurun_out/bench_v31.json 2>gpurun_out/bench_v31.err; tail -3 gpurun_out/bench_v31.err
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_v31.json').read()); print(d['value'], d['ms_per_step'], d['stage_ms'], d['roofline']['by_class'])"
PIPER_B200_SDP_TC=0 timeout 300 python bench.py --no-cpu --steps 5 > gpurun_out/bench_v31_nosdp.json 2>>gpurun_out/bench_v31.err
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_v31_nosdp.json').read()); print('SDP_TC=0', d['value'], d['ms_per_step'], d['stage_ms']['front_ms'])"
timeout 600 python tools/front_tc_flips.py --batches 4 > gpurun_out/front_tc_flips_v31.log 2>&1; tail -4 gpurun_out/front_tc_flips_v31.log
