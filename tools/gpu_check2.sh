# smoke + GPU suite + one bench --no-cpu of the default build
cd /root/repo
TAG=${1:-chk}
timeout 150 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python -m pytest tests -m gpu -x -q --timeout 240 > gpurun_out/gpu_tests_$TAG.log 2>&1; tail -3 gpurun_out/gpu_tests_$TAG.log
timeout 240 python bench.py --no-cpu > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; python tools/bench_brief.py gpurun_out/bench_$TAG.json | head -9
