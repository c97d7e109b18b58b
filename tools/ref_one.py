"""Reference-audio path (WavLM + global encoder), config 4 shape, for ncu: python tools/ref_one.py [batch] [seconds] [iters]"""
import os
import sys
sys.path.insert(0, ".")
sys.path.insert(0, "tools")
from make_synth_gguf import write_gguf, write_wavlm_gguf
from mio_tts_cpp_b200 import MioTTS, benchutil as U
a = [float(x) for x in sys.argv[1:]] + [32, 10.0, 3][len(sys.argv) - 1:]
dec, wl = "/tmp/mio_bench_tiny_ge.gguf", "/tmp/mio_bench_wavlm_base.gguf"
if not os.path.exists(dec):
    write_gguf(dec, "tiny", 0, "default")
if not os.path.exists(wl):
    write_wavlm_gguf(wl, "base", 0)
m = MioTTS(dec, wl)
clip = U.reference_clip(a[1], 24000, 3)
print("ms per batch:", m.reference_bench(clip, 24000, int(a[0]), int(a[2])))
