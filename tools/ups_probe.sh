# is the merged-polyphase (ConvTranspose) epilogue bound by its global stores?  per-launch times of the last 4 conv launches
cd /root/repo
for d in 0 4; do
  env PIPER_B200_TC_DBG=$d PIPER_B200_PROF_DUMP=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/prof_ups_$d.txt > /dev/null
  echo "== PIPER_B200_TC_DBG=$d"; python tools/prof_launches.py --digest gpurun_out/prof_ups_$d.txt | tail -1 | awk '{n=NF; print $(n-3), $(n-2), $(n-1), $n}'
done
