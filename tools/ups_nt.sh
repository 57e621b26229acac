cd /root/repo
for nt in 0 128 64 0; do
  env PIPER_B200_UPS_NT=$nt PIPER_B200_PROF_DUMP=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/prof_upsnt.txt > /dev/null
  echo "UPS_NT=$nt:" $(python tools/prof_launches.py --digest gpurun_out/prof_upsnt.txt 2>/dev/null | tail -1 | awk '{n=NF; print $(n-3), $(n-2), $(n-1), $n}')
done
