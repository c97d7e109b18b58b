#!/usr/bin/env python3
"""Stage-by-stage check of the CUDA path against the numpy oracle and the committed golden vectors.

Development aid for the GPU box (`gpurun -- python tools/gpu_debug.py`); the real parity tests live in tests/.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

from make_synth_gguf import write_gguf  # noqa: E402
from oracle import miocodec_np as O  # noqa: E402
from mio_tts_cpp_b200 import MioTTS  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def stats(name, got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    if got.shape != ref.shape:
        print(f"  {name}: SHAPE MISMATCH got {got.shape} ref {ref.shape}")
        return
    err = got - ref
    den = float((ref * ref).sum())
    snr = 10 * np.log10(den / max(float((err * err).sum()), 1e-300)) if den > 0 else float("nan")
    print(f"  {name}: max_abs={np.abs(err).max():.3e} ref_max={np.abs(ref).max():.3e} snr={snr:.1f} dB "
          f"nan={int(np.isnan(got).sum())}")


def main():
    presets = sys.argv[1:] or ["tiny", "default"]
    emb = np.load(os.path.join(GOLD, "en_female_emb.npy"))
    for n_fft, hop in ((120, 30), (392, 98), (1280, 320), (1920, 480), (2048, 512)):
        pass
    ctx_for_istft = None
    for preset in presets:
        path = f"/tmp/mio_dbg_{preset}.gguf"
        t0 = time.time()
        write_gguf(path, preset, 0)
        print(f"[{preset}] wrote gguf in {time.time() - t0:.1f}s")
        t0 = time.time()
        m = MioTTS(path)
        print(f"[{preset}] loaded in {time.time() - t0:.1f}s weights={m.weight_bytes() / 1e6:.1f} MB")
        om = O.Model(path)
        if ctx_for_istft is None:
            ctx_for_istft = m
            for f in sorted(os.listdir(GOLD)):
                if f.startswith("istft_"):
                    z = np.load(os.path.join(GOLD, f))
                    a = m.istft(z["spec"], int(z["n_fft"]), int(z["hop"]))
                    stats(f, a, z["audio"])
        fixtures = {"tiny": [("tiny_T50", 50, 42), ("tiny_T1", 1, 7), ("tiny_T133", 133, 9)],
                    "default": [("default_T24", 24, 42)],
                    "upsampler_tiny": [("upsampler_tiny_T20", 20, 42)],
                    "upsampler": [("upsampler_T12", 12, 42)]}[preset]
        for name, T, seed in fixtures:
            z = np.load(os.path.join(GOLD, name + ".npz"))
            codes = z["codes"]
            print(f"[{preset}] {name}: T={T}")
            g = m.gather(codes)
            print("  gather bit-exact:", bool(np.array_equal(g, O.gather_codebook(om, codes))))
            spec_o, st = O.decode_tokens(om, codes, emb, return_stages=True)
            for i, key in enumerate(["prenet", "upsample", "prior", "decoder", "post"], start=1):
                try:
                    stats(f"stage{i}:{key}", m.debug_stage(codes, emb, i), st[key])
                except Exception as e:  # noqa: BLE001
                    print(f"  stage{i}:{key}: ERROR {e}")
            spec = m.decode_spec(codes, emb)
            stats("spec vs golden(ref)", spec, z["spec"])
            stats("spec vs numpy oracle", spec, spec_o)
            m.set_generic_attention(True)
            stats("spec (generic attn) vs golden", m.decode_spec(codes, emb), z["spec"])
            m.set_generic_attention(False)
            audio, sr = m.synthesize(codes, emb)
            stats("audio vs golden(ref)", audio, z["audio"])
        if preset == "default":
            rng = np.random.default_rng(1)
            codes_l = [rng.integers(0, 12800, 400).astype(np.int32) for _ in range(8)]
            t0 = time.time()
            res, sr = m.synthesize_batch(codes_l, [emb] * 8)
            print(f"  batch 8x400 e2e {1e3 * (time.time() - t0):.1f} ms (first call)")
            t0 = time.time()
            res, sr = m.synthesize_batch(codes_l, [emb] * 8)
            print(f"  batch 8x400 e2e {1e3 * (time.time() - t0):.1f} ms")
            one, _ = m.synthesize(codes_l[3], emb)
            stats("batch[3] vs single", res[3], one)
            m.prepare(codes_l, [emb] * 8)
            print("  resident ms/pass:", m.run_resident(3), m.profile_resident())
        print(f"[{preset}] launches so far: {m.launch_count()}")


if __name__ == "__main__":
    main()
