cd /root/repo
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v54.log 2>&1; tail -3 gpurun_out/gpu_tests_v54.log
timeout 200 python bench.py --no-cpu --steps 6 > gpurun_out/bench_v54.json 2>gpurun_out/bench_v54.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v54.json').read()); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['by_class'])"
timeout 60 python bench.py --voice high --batch 1 --t-lo 2000 --t-hi 2000 --no-cpu > gpurun_out/bench_v54_high.json 2>>gpurun_out/bench_v54.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v54_high.json').read()); print('high', d['value'], d['ms_per_step'])"
