timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v32.log 2>&1; tail -5 gpurun_out/gpu_tests_v32.log
timeout 300 python bench.py --no-cpu > gpurun_out/bench_v32.json 2>gpurun_out/bench_v32.err; tail -3 gpurun_out/bench_v32.err
python -c "import json,sys; d=json.loads(open('gpurun_out/bench_v32.json').read()); print(d['value'], d['ms_per_step'], d['stage_ms'], d['roofline']['by_class'])"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:attention96 -c 6 --csv --log-file gpurun_out/att_v32.csv python bench.py --steps 1 --warmup 0 --no-cpu > /dev/null 2>&1; grep attention gpurun_out/att_v32.csv | tail -2 | cut -c1-50,200-400
