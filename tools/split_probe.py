"""fp32 emulation on the tensor cores (bf16x6 split, conv_tc.h) vs the fp32 CUDA-core kernel: error against a
float64 host reference and kernel time, per front-end convolution shape.  Run on the GPU box:
    python tools/split_probe.py [rows]"""
import json
import sys

sys.path.insert(0, "piper-swift_b200")
sys.path.insert(0, "tests")
from common import voice_files          # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options   # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 28672
mp, cp = voice_files("medium")
r = PiperCUDARuntime(mp, cp, Options(precision="fp32"))
for name, cin, cout, k in [("qkv", 192, 576, 1), ("o", 192, 192, 1), ("ffn1", 192, 768, 3), ("ffn2", 768, 192, 3),
                           ("x_low ffn1", 96, 384, 3)]:
    a = r.selftest_conv(4, cin, cout, k, rows=rows)
    b = r.selftest_conv(5, cin, cout, k, rows=rows)
    fl = 2.0 * rows * cin * cout * k
    print(json.dumps({"conv": name, "cin": cin, "cout": cout, "k": k, "rows": rows,
                      "rms_err_tc": a["max_abs_diff"], "rms_err_simt": a["max_abs_ref"],
                      "max_err_tc": b["max_abs_diff"], "max_err_simt": b["max_abs_ref"],
                      "tc_ms": b["tc_ms"], "simt_ms": b["f32_ms"],
                      "tc_tflops_fp32_equiv": fl / b["tc_ms"] / 1e9, "simt_tflops": fl / b["f32_ms"] / 1e9}))
