# ups-focused A/B: per-launch times (last 4 conv launches = conv_pre/ups) + bench, default vs variant lib (HEAD's conv_tc.cu); then the tc parity tests
cd /root/repo
TAG=${1:-ab6}
for lib in "" piper-swift_b200/libpiper_b200_variant.so ""; do
  name=$([ -z "$lib" ] && echo default || echo variant)
  PIPER_B200_LIB=$lib PIPER_B200_PROF_DUMP=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/prof_${TAG}_$name.txt > /dev/null
  echo "== $name"; python tools/prof_launches.py --digest gpurun_out/prof_${TAG}_$name.txt | tail -1 | awk '{n=NF; print "ups launches:", $(n-3), $(n-2), $(n-1), $n}'
  python tools/prof_launches.py --digest gpurun_out/prof_${TAG}_$name.txt | sed -n 2p
  PIPER_B200_LIB=$lib timeout 240 python bench.py --no-cpu > gpurun_out/bench_${TAG}_$name.json 2> gpurun_out/bench_${TAG}_$name.err; python tools/bench_brief.py gpurun_out/bench_${TAG}_$name.json | head -3
done
timeout 500 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -x -q --timeout 240 2>&1 | tail -3
