"""One selftest launch of the tcgen05 convolution kernel (for ncu): python tools/conv_one.py mode cin cout k rows"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "piper-swift_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from common import voice_files
from piper_b200.runtime import PiperCUDARuntime, Options
mode, cin, cout, k, rows = (int(x) for x in sys.argv[1:6])
mp, cp = voice_files("tiny")
rt = PiperCUDARuntime(mp, cp, Options(precision="fp32"))
print(rt.selftest_conv(mode, cin, cout, k, 1, 1, rows=rows, precision="f16"))
