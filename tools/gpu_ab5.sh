# smoke + suite on the default build, then bench --no-cpu: default, no pre-wait (variant lib), window off, default; trace of the default
cd /root/repo
TAG=${1:-ab5}
timeout 150 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 500 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/gpu_tests_$TAG.log 2>&1; tail -3 gpurun_out/gpu_tests_$TAG.log
run() {
  local name=$1; shift
  echo "== $name"
  env "$@" timeout 240 python bench.py --no-cpu > gpurun_out/bench_${TAG}_$name.json 2> gpurun_out/bench_${TAG}_$name.err
  python tools/bench_brief.py gpurun_out/bench_${TAG}_$name.json | head -9
}
run default X=1
run noprewait PIPER_B200_LIB=piper-swift_b200/libpiper_b200_variant.so
run nowindow PIPER_B200_FLOW_WINDOW=0
run default2 X=1
bash tools/tc_trace_run.sh $TAG X=1 > gpurun_out/trace_$TAG.log 2>&1
