# A/B on the GPU box: default library vs libpiper_b200_variant.so (make variant VARIANT_FLAGS=...), per-launch times + bench
cd /root/repo
TAG=${1:-ab}
for lib in "" piper-swift_b200/libpiper_b200_variant.so; do
  name=$([ -z "$lib" ] && echo default || echo variant)
  PIPER_B200_LIB=$lib PIPER_B200_PROF_DUMP=1 timeout 120 python tools/prof_launches.py 2> gpurun_out/prof_${TAG}_$name.txt > gpurun_out/prof_${TAG}_$name.log
  echo "== $name"; python tools/prof_launches.py --digest gpurun_out/prof_${TAG}_$name.txt | head -2
  PIPER_B200_LIB=$lib timeout 240 python bench.py ${BENCH_FLAGS:---no-cpu} > gpurun_out/bench_${TAG}_$name.json 2> gpurun_out/bench_${TAG}_$name.err; python tools/bench_brief.py gpurun_out/bench_${TAG}_$name.json | head -8
done
