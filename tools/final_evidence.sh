# Round-end evidence on the GPU box: tests, smoke, bench (both arms), ncu launch list, ncu --set full of the two
# tensor-core kernel classes.  usage: bash tools/final_evidence.sh TAG   (outputs under gpurun_out/)
TAG=${1:-vX}
cd /root/repo
timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; tail -2 gpurun_out/gpu_tests_$TAG.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; tail -2 gpurun_out/smoke_$TAG.log
timeout 400 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; cut -c1-300 gpurun_out/bench_$TAG.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_${TAG}_reference.json 2>> gpurun_out/bench_$TAG.err; cut -c1-200 gpurun_out/bench_${TAG}_reference.json
timeout 120 python bench.py --voice high --batch 1 --t-lo 2000 --t-hi 2000 --no-cpu > gpurun_out/bench_${TAG}_high.json 2>> gpurun_out/bench_$TAG.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/ncu_launches_$TAG.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_launches_$TAG.log 2>&1
timeout 400 ncu --set full --clock-control none -k regex:mrf_tc_kernel -s 3 -c 3 -o /tmp/mrf_full -f python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_full_mrf_$TAG.log 2>&1
ncu -i /tmp/mrf_full.ncu-rep --page raw --csv > gpurun_out/ncu_full_mrf_$TAG.raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none -k regex:conv_tc_persist -s 78 -c 78 -o /tmp/ctc_full -f python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_full_conv_tc_$TAG.log 2>&1
ncu -i /tmp/ctc_full.ncu-rep --page raw --csv > gpurun_out/ncu_full_conv_tc_$TAG.raw.csv 2>/dev/null
ls -la gpurun_out | tail -12
