import sys
sys.path.insert(0, ".")
from mio_tts_cpp_b200 import MioTTS, benchutil as U
m = MioTTS(U.synth_gguf("tiny"))
for (nu, fr) in ((64, 800), (1000, 1000)):
    for nfft, hop in ((1920, 480), (2048, 512), (1280, 320)):
        ms, cs = m.istft_bench(nu, fr, nfft, hop, 10)
        by = nu * fr * ((nfft + 2) * 4 + hop * 4)
        print(nu, fr, nfft, "ms", round(ms, 4), "GB/s", round(by / ms / 1e6, 1), "checksum", cs)
