cd /root/repo
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29548 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_v53_n8.json 2> gpurun_out/bench_v53_n8.err; cut -c1-260 gpurun_out/bench_v53_n8.json; grep -c . gpurun_out/bench_v53_n8.json; tail -3 gpurun_out/bench_v53_n8.err
