cd /root/repo
nvidia-smi topo -m 2>&1 | head -14 > gpurun_out/topo.log; nproc >> gpurun_out/topo.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29549 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_v55_n8.json 2> gpurun_out/bench_v55_n8.err; python -c "import json; d=json.load(open('gpurun_out/bench_v55_n8.json')); print(d['value'], d['e2e']['value'], d['ms_per_step'])"; grep "bench\]" gpurun_out/bench_v55_n8.err | head -3; cat gpurun_out/topo.log | cut -c1-150
