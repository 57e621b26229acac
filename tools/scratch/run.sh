cd /root/repo
timeout 300 python tools/conv_rows_probe.py 2>&1 | grep "tiles/CTA [5]" > gpurun_out/conv_rows_probe_v56.log; cat gpurun_out/conv_rows_probe_v56.log
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v56.log 2>&1; tail -2 gpurun_out/gpu_tests_v56.log
timeout 200 python bench.py --no-cpu --steps 6 > gpurun_out/bench_v56.json 2>gpurun_out/bench_v56.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v56.json').read()); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['by_class'])"
