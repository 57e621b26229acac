# latency configurations + parity suite for a small-call change: default lib vs variant lib
cd /root/repo
TAG=${1:-lat}
timeout 400 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -x -q --timeout 240 2>&1 | tail -2
for lib in "" piper-swift_b200/libpiper_b200_variant.so ""; do
  name=$([ -z "$lib" ] && echo default || echo variant)
  echo "== $name"
  PIPER_B200_LIB=$lib timeout 240 python bench.py --no-cpu > gpurun_out/bench_${TAG}_$name.json 2> gpurun_out/bench_${TAG}_$name.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_${TAG}_$name.json").read().strip().splitlines()[-1])
print("value %.0f ms/step %.3f" % (d["value"], d["ms_per_step"]), "config1", d.get("config1"), "config2", d.get("config2"))
PY
done
