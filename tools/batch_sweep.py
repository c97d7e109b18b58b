#!/usr/bin/env python3
"""Resident-throughput sweep over the batch size (utterances per step) and per-stage / per-GEMM breakdown.

  python tools/batch_sweep.py [--batches 8,16,32,64,128] [--codes 400] [--gemm-log]

Used to decide the utterance chunk size of a step (does a smaller batch whose activations fit the 126 MB L2
beat one big batch?) and to read the per-shape GEMM table (MIO_TTS_B200_GEMM_LOG=1).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="8,16,32,64,128")
    ap.add_argument("--codes", type=int, default=400)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--gemm-log", action="store_true")
    ap.add_argument("--preset", default="default", help="synthetic-weight preset (tools/make_synth_gguf.py), e.g. upsampler")
    args = ap.parse_args()
    if args.gemm_log:
        os.environ["MIO_TTS_B200_GEMM_LOG"] = "1"
    from mio_tts_cpp_b200 import MioTTS
    from mio_tts_cpp_b200 import benchutil as U

    m = MioTTS(U.synth_gguf(args.preset))
    emb = U.speaker_embedding()
    sr = m.sample_rate()
    for b in [int(x) for x in args.batches.split(",")]:
        codes = U.config2_workload(0, b, args.codes)
        m.prepare(codes, [emb] * b)
        for _ in range(3):
            m.run_resident(1)
        ms = m.run_resident(args.iters)
        secs = m.resident_samples() / sr
        p = m.profile_resident()
        print(json.dumps({"preset": args.preset, "batch": b, "codes": args.codes, "ms_per_step": ms, "audio_s_per_s": secs / (ms * 1e-3),
                          "stages": {k: round(v, 3) for k, v in p.items() if k.endswith("_ms")},
                          "gemm_tflops": p["gemm_flops"] / (p["gemm_ms"] * 1e-3) / 1e12}), flush=True)
    m.free()


if __name__ == "__main__":
    main()
