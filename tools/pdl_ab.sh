cd /root/repo
for rep in 1 2; do for v in 1 0; do
  PIPER_B200_PDL=$v timeout 200 python bench.py --no-cpu --no-extras > gpurun_out/bench_pdl2_${v}_$rep.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_pdl2_${v}_$rep.json").read().strip().splitlines()[-1])
print("PDL=$v rep $rep value ms %.3f e2e ms %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]))
PY
done; done
