"""One iSTFT launch shape for ncu: python tools/istft_one.py [n_utt] [frames] [n_fft] [hop]"""
import sys
sys.path.insert(0, ".")
from mio_tts_cpp_b200 import MioTTS, benchutil as U
a = [int(x) for x in sys.argv[1:]] + [64, 800, 1920, 480][len(sys.argv) - 1:]
m = MioTTS(U.synth_gguf("tiny"))
print(m.istft_bench(a[0], a[1], a[2], a[3], 2))
