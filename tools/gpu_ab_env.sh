# smoke, GPU suite, then bench --no-cpu with VAR=1 / VAR=0 / VAR=1   usage: bash tools/gpu_ab_env.sh TAG VAR
cd /root/repo
TAG=$1; VAR=$2
timeout 150 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 || { echo "SMOKE FAILED"; exit 1; }
timeout 500 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/gpu_tests_$TAG.log 2>&1; tail -3 gpurun_out/gpu_tests_$TAG.log
for v in 1 0 1; do
  echo "== $VAR=$v"
  env $VAR=$v timeout 240 python bench.py --no-cpu > gpurun_out/bench_${TAG}_$v.json 2> gpurun_out/bench_${TAG}_$v.err; python tools/bench_brief.py gpurun_out/bench_${TAG}_$v.json | head -9
done
