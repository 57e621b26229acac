cd /root/repo
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v57.log 2>&1; tail -3 gpurun_out/gpu_tests_v57.log
timeout 200 python bench.py --no-cpu --steps 6 > gpurun_out/bench_v57.json 2>gpurun_out/bench_v57.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v57.json').read()); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['stage_ms']['front_ms'])"
PIPER_B200_LN_PAIR=0 timeout 200 python bench.py --no-cpu --steps 6 > gpurun_out/bench_v57_scalar.json 2>gpurun_out/bench_v57.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v57_scalar.json').read()); print('scalar LN value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['stage_ms']['front_ms'])"
