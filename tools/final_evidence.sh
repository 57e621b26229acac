# Round evidence on the GPU box: tests, smoke, bench (both arms), ncu launch list, ncu --set full of the two tensor-core
# kernel classes.  usage: bash tools/final_evidence.sh TAG   (outputs under gpurun_out/; copy what is to be judged to profiles/)
TAG=${1:-vX}
cd /root/repo
timeout 700 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/gpu_tests_$TAG.log 2>&1; tail -2 gpurun_out/gpu_tests_$TAG.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; tail -2 gpurun_out/smoke_$TAG.log
timeout 400 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; python tools/bench_brief.py gpurun_out/bench_$TAG.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_${TAG}_reference.json 2>> gpurun_out/bench_$TAG.err; cut -c1-200 gpurun_out/bench_${TAG}_reference.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/ncu_launches_$TAG.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_launches_$TAG.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mrf_tc_kernel -s 3 -c 3 -o /tmp/mrf_full -f python bench.py --steps 1 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_full_mrf_$TAG.log 2>&1
ncu -i /tmp/mrf_full.ncu-rep --page raw --csv > gpurun_out/ncu_full_mrf_$TAG.raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none -k regex:conv_tc_persist -s 78 -c 78 -o /tmp/ctc_full -f python bench.py --steps 1 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_full_conv_tc_$TAG.log 2>&1
ncu -i /tmp/ctc_full.ncu-rep --page raw --csv > gpurun_out/ncu_full_conv_tc_$TAG.raw.csv 2>/dev/null
python tools/ncu_raw_pick.py gpurun_out/ncu_full_mrf_$TAG.raw.csv > gpurun_out/ncu_full_mrf_${TAG}_summary.csv
python tools/ncu_raw_pick.py gpurun_out/ncu_full_conv_tc_$TAG.raw.csv > gpurun_out/ncu_full_conv_tc_${TAG}_summary.csv
ls -la gpurun_out | tail -14
