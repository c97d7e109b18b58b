#!/usr/bin/env python3
"""GPU-side check of the reference-audio path against the numpy oracle (development tool, runs on the GPU box)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import make_synth_gguf as M  # noqa: E402
from mio_tts_cpp_b200 import MioTTS  # noqa: E402
from oracle import miocodec_np as O  # noqa: E402
from oracle import wavlm_np as W  # noqa: E402


def snr(a, b):
    a = a.astype(np.float64)
    b = b.astype(np.float64)
    return 10 * np.log10((a ** 2).sum() / max(1e-300, ((a - b) ** 2).sum()))


def synth_pcm(seconds, sr=24000, seed=3):
    rng = np.random.default_rng(seed)
    n = int(seconds * sr)
    t = np.arange(n) / sr
    x = np.zeros(n)
    for _ in range(12):
        f, a, ph = rng.uniform(80, 3000), rng.uniform(0.05, 0.3), rng.uniform(0, 6.28)
        env = 0.5 * (1 + np.sin(2 * np.pi * rng.uniform(0.3, 3) * t + rng.uniform(0, 6.28)))
        x += a * env * np.sin(2 * np.pi * f * t + ph)
    x += 0.02 * rng.standard_normal(n)
    x = x / np.abs(x).max() * 0.8
    return (np.round(x * 32767) / 32768.0).astype(np.float32)


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "tiny"
    wl_preset, ge_preset, secs = ("tiny", "tiny", 2.0) if which == "tiny" else ("base", "default", 10.0)
    dec, wl = f"/tmp/dbg_dec_{which}.gguf", f"/tmp/dbg_wavlm_{which}.gguf"
    M.write_gguf(dec, "tiny", 0, ge_preset)
    M.write_wavlm_gguf(wl, wl_preset, 0)
    m = MioTTS(dec, wl)
    pcm = synth_pcm(secs)
    wm, om = W.WavLM(wl), O.Model(dec)
    t0 = time.time()
    ssl_ref = W.extract_ssl_from_pcm(wm, pcm, 24000)
    emb_ref = W.global_embedding(om, ssl_ref)
    print("oracle time", time.time() - t0)
    ssl = m.extract_ssl(pcm, 24000)
    print("ssl", ssl.shape, ssl_ref.shape, "snr", snr(ssl_ref, ssl), "maxabs", np.abs(ssl_ref - ssl).max())
    emb1 = m.ssl_to_embedding(ssl_ref)
    print("global encoder on oracle ssl: snr", snr(emb_ref, emb1), "maxabs", np.abs(emb_ref - emb1).max())
    embs = m.reference_pcm_to_embedding_batch([pcm, pcm[: len(pcm) * 3 // 4], pcm], 24000)
    print("end to end: snr", snr(emb_ref, embs[0]), "maxabs", np.abs(emb_ref - embs[0]).max(),
          "batch consistency", np.abs(embs[0] - embs[2]).max())
    emb_short = W.global_embedding(om, W.extract_ssl_from_pcm(wm, pcm[: len(pcm) * 3 // 4], 24000))
    print("ragged clip: snr", snr(emb_short, embs[1]))
    for b in (1, 8, 32):
        ms = m.reference_bench(pcm, 24000, b, 5)
        print(f"bench batch {b}: {ms:.3f} ms per batch, {b * secs / (ms * 1e-3):.0f} audio-s/s, {b / (ms * 1e-3):.0f} clips/s")
    m.free()


if __name__ == "__main__":
    main()
