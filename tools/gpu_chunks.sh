# e2e against the number of row ranges the host-bound audio leaves in
cd /root/repo
for c in 3 4 5 6 3; do
  PIPER_B200_D2H_CHUNKS=$c timeout 200 python bench.py --no-cpu --no-extras > gpurun_out/bench_chunks_$c.json 2> gpurun_out/bench_chunks_$c.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_chunks_$c.json").read().strip().splitlines()[-1])
print("chunks $c value %.0f ms/step %.3f e2e %.0f (%.3f ms)" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"].get("ms_per_step", 0)))
PY
done
