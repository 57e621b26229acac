# stage-kernel knob sweep at HEAD: per-launch times of the three fused stage launches (config 3)
cd /root/repo
run() {
  env "$@" PIPER_B200_PROF_DUMP=1 timeout 200 python tools/prof_launches.py 2> gpurun_out/prof_sweep.txt > /dev/null
  echo "$@ :" $(python tools/prof_launches.py --digest gpurun_out/prof_sweep.txt | head -1 | cut -c1-150)
}
run X=1
run PIPER_B200_MRF_EG=3
run PIPER_B200_MRF_EG=5
run PIPER_B200_MRF_M64=3
run PIPER_B200_MRF_M=6
run PIPER_B200_MRF_GX=0
run PIPER_B200_MRF_SPLIT=0
run X=1
