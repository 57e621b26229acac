# conv_tc phase / wait trace of one config-3 step (CTA 0): all 78 launches of the last call
cd /root/repo
TAG=$1; shift
env "$@" PIPER_B200_TC_TRACE=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/tctrace_$TAG.txt > /dev/null
grep "conv_tc trace" gpurun_out/tctrace_$TAG.txt | tail -78 > gpurun_out/tctrace_${TAG}_last.txt
sed -n 1,12p gpurun_out/tctrace_${TAG}_last.txt; sed -n 36,50p gpurun_out/tctrace_${TAG}_last.txt; tail -4 gpurun_out/tctrace_${TAG}_last.txt
