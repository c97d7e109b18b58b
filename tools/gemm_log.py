"""Per-GEMM timings of one config-2 step (CUDA events around every launch): MIO_TTS_B200_GEMM_LOG=1 python tools/gemm_log.py"""
import os
import sys
os.environ["MIO_TTS_B200_GEMM_LOG"] = "1"
sys.path.insert(0, ".")
from mio_tts_cpp_b200 import MioTTS, benchutil as U
m = MioTTS(U.synth_gguf("default"))
codes = U.config2_workload(0, 64, 400)
m.prepare(codes, [U.speaker_embedding()] * 64)
for _ in range(3):
    m.run_resident(1)
m.profile_resident()
sys.stderr.flush()
sys.stderr.write("=====\n")
p = m.profile_resident()
print(p)
