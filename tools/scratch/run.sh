cd /root/repo
run() { env "$@" timeout 100 python bench.py --no-cpu --steps 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', round(d['ms_per_step'],3), round(d['roofline']['by_class']['mrf_tc_kernel']['ms_per_step'],3))"; }
run X=0
run PIPER_B200_MRF_EG=3
run PIPER_B200_MRF_EG=5
for m in 2 3 4 5 6 7 8; do run PIPER_B200_MRF_M=$m; done
for m in 2 3 4; do run PIPER_B200_MRF_M64=$m; done
run PIPER_B200_MRF_SPLIT=0
run X=0
