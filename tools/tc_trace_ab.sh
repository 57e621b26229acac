# conv_tc phase trace (CTA 0, first tile) + per-launch event times of one config-3 step, VAR=1 vs VAR=0
cd /root/repo
TAG=$1; VAR=$2
for v in 1 0; do
  env $VAR=$v PIPER_B200_TC_TRACE=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/tctrace_${TAG}_$v.txt > /dev/null
  env $VAR=$v PIPER_B200_PROF_DUMP=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/prof_${TAG}_$v.txt > /dev/null
  echo "== $VAR=$v"; python tools/prof_launches.py --digest gpurun_out/prof_${TAG}_$v.txt | tail -1 | cut -c1-900
  grep "conv_tc trace" gpurun_out/tctrace_${TAG}_$v.txt | tail -78 | sed -n 52,62p
done
