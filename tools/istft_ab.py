import sys, os, json
sys.path.insert(0, '.')
from mio_tts_cpp_b200 import MioTTS, benchutil as U
m = MioTTS(U.synth_gguf("tiny"))
ms, _ = m.istft_bench(1000, 1000, 1920, 480, 5)
ms2, _ = m.istft_bench(64, 800, 1920, 480, 20)
print(os.environ.get("MIO_TTS_B200_ISTFT_FW", "8"), "1M frames ms", ms, "GB/s", 1e6*9608/(ms*1e-3)/1e9, "64x800 ms", ms2)
