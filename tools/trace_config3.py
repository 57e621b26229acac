"""Per-role clock trace of the three fused stage launches of BASELINE config 3 (256 utterances, medium voice, f16).
    make -C piper-swift_b200 trace
    PIPER_B200_LIB=piper-swift_b200/libpiper_b200_trace.so PIPER_B200_MRF_TRACE=gpurun_out/mrf_trace python tools/trace_config3.py
    python tools/mrf_trace.py gpurun_out/mrf_trace.128   (.64 / .32)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from common import S, voice_files                                   # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options           # noqa: E402
from piper_b200.synth_voice import synthetic_utterances            # noqa: E402

if __name__ == "__main__":
    mp, cp = voice_files("medium")
    rt = PiperCUDARuntime(mp, cp, Options(precision="f16"))
    utts = synthetic_utterances(256, 256)
    r = rt.synthesize_batch(utts, *S.FIXTURE_SCALES, copy=False)
    print("gpu_ms", rt.last_run_timings["gpu_ms"], "samples", r.total_samples)
    rt.close()
