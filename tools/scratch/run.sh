timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v39.log 2>&1; tail -3 gpurun_out/gpu_tests_v39.log
timeout 60 python bench.py --voice high --batch 1 --t-lo 2000 --t-hi 2000 --no-cpu > gpurun_out/bench_v39_high.json 2>gpurun_out/bench_v39.err; tail -2 gpurun_out/bench_v39.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v39_high.json').read()); print('high', d['value'], d['ms_per_step'], d['stage_ms'])"
PIPER_B200_ATT_SPLIT=0 timeout 60 python bench.py --voice high --batch 1 --t-lo 2000 --t-hi 2000 --no-cpu > gpurun_out/bench_v39_high_nosplit.json 2>>gpurun_out/bench_v39.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v39_high_nosplit.json').read()); print('high ATT_SPLIT=0', d['value'], d['ms_per_step'], d['stage_ms']['front_ms'])"
timeout 200 python bench.py --no-cpu --steps 5 > gpurun_out/bench_v39.json 2>>gpurun_out/bench_v39.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v39.json').read()); print('medium', d['value'], d['ms_per_step'], d['stage_ms']['front_ms'])"
