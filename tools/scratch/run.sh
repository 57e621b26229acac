cd /root/repo
for nb in 4 2; do echo nbuf $nb; PIPER_B200_TC_NBUF=$nb timeout 300 python tools/conv_rows_probe.py 2>&1 | grep "tiles/CTA [5]"; done > gpurun_out/conv_rows_probe_v51.log; cat gpurun_out/conv_rows_probe_v51.log
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v51.log 2>&1; tail -3 gpurun_out/gpu_tests_v51.log
timeout 200 python bench.py --no-cpu --steps 6 > gpurun_out/bench_v51.json 2>gpurun_out/bench_v51.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v51.json').read()); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['by_class'])"
PIPER_B200_TC_NBUF=2 timeout 200 python bench.py --no-cpu --steps 6 > gpurun_out/bench_v51_nbuf2.json 2>gpurun_out/bench_v51.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v51_nbuf2.json').read()); print('nbuf2 value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['by_class']['conv_tc_kernel'])"
