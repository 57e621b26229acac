timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v37.log 2>&1; tail -3 gpurun_out/gpu_tests_v37.log
timeout 120 python tools/tc_parity_report.py > gpurun_out/tc_parity_v37.log 2>&1; grep -E "high|x_low" gpurun_out/tc_parity_v37.log
timeout 60 python bench.py --voice high --batch 1 --t-lo 2000 --t-hi 2000 --no-cpu > gpurun_out/bench_v37_high.json 2>gpurun_out/bench_v37.err; tail -2 gpurun_out/bench_v37.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v37_high.json').read()); print('high', d['value'], d['ms_per_step'], d['stage_ms'], d['roofline']['by_class'])"
timeout 60 python bench.py --voice high --batch 64 --t-lo 80 --t-hi 120 --no-cpu --steps 5 > gpurun_out/bench_v37_high_64.json 2>>gpurun_out/bench_v37.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v37_high_64.json').read()); print('high64', d['value'], d['ms_per_step'], d['stage_ms'], d['roofline']['by_class'])"
timeout 200 python bench.py --no-cpu --steps 5 > gpurun_out/bench_v37.json 2>>gpurun_out/bench_v37.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v37.json').read()); print('medium', d['value'], d['ms_per_step'], d['roofline']['by_class']['mrf_tc_kernel'])"
