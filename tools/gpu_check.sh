# GPU box: full GPU test suite, smoke, bench (both arms).  usage: bash tools/gpu_check.sh TAG  (outputs under gpurun_out/)
TAG=${1:-vX}
cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; tail -3 gpurun_out/gpu_tests_$TAG.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; tail -3 gpurun_out/smoke_$TAG.log
timeout 500 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; cut -c1-400 gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err
