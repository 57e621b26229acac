"""One-screen digest of a bench.py JSON line.  usage: python tools/bench_brief.py <file>"""
import json
import sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r = d.get("roofline", {})
print("value %.0f %s  ms/step %.3f  e2e %.0f  frac %.3f (%s)" % (d["value"], d["unit"], d["ms_per_step"], d.get("e2e", {}).get("value", 0),
      r.get("frac", 0), r.get("kernel", "")[:24]))
for k, v in r.get("by_class", {}).items():
    print("   %-18s %.3f ms  %.0f TFLOP/s  x%d" % (k, v["ms_per_step"], v["tflops"], v["launches_per_step"]))
print("   stage_ms", {k: round(v, 3) for k, v in d.get("stage_ms", {}).items()})
if "parity" in d: print("   parity", {k: d["parity"][k] for k in ("utterances", "duration_mismatches", "snr_db_min", "max_abs", "ok")})
for k in ("config1", "config4", "config5"):
    if k in d: print("   %s" % k, {kk: (round(vv, 3) if isinstance(vv, float) else vv) for kk, vv in d[k].items() if kk in ("ms_mean", "ms_per_step", "value")})
