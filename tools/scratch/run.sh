python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_v36.log 2>&1; tail -2 gpurun_out/gpu_tests_v36.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_v36.log 2>&1; tail -1 gpurun_out/smoke_v36.log
python bench.py > gpurun_out/bench_v36.json 2>gpurun_out/bench_v36.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_v36_reference.json 2>>gpurun_out/bench_v36.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_v36.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_v36.log 2>&1
tail -c 400 gpurun_out/bench_v36.json
