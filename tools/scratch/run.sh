timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v40.log 2>&1; tail -3 gpurun_out/gpu_tests_v40.log
timeout 200 python bench.py --steps 6 > gpurun_out/bench_v40.json 2>gpurun_out/bench_v40.err
python -c "import json; d=json.loads(open('gpurun_out/bench_v40.json').read()); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e'], d.get('gpu_launches'), d['roofline'], d['cpu_baseline'])"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 1 --impl reference --steps 2 --warmup 1 > gpurun_out/bench_v40_reference.json 2>>gpurun_out/bench_v40.err; cat gpurun_out/bench_v40_reference.json | cut -c1-400
