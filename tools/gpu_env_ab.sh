# same library, one environment switch on / off: usage  bash tools/gpu_env_ab.sh TAG VAR   (bench --no-cpu with VAR=1 and VAR=0)
cd /root/repo
TAG=$1; VAR=$2
for v in 1 0; do
  echo "== $VAR=$v"
  env $VAR=$v timeout 240 python bench.py --no-cpu > gpurun_out/bench_${TAG}_$v.json 2> gpurun_out/bench_${TAG}_$v.err; python tools/bench_brief.py gpurun_out/bench_${TAG}_$v.json | head -12
done
