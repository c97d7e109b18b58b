#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref) in this container.

/root/reference does not exist on the GPU box, so the reference's outputs on seeded inputs are
committed as small fixtures.  Inputs are reproducible: weights = tools/make_synth_gguf.py (preset, seed 0),
codes = default_rng(seed).integers(0, 12800, T), embedding = models/en_female.emb.gguf (copied into the
fixture; 128 floats).  Each fixture records the weight sha256 so a drifting generator is detected.

    python tools/gen_golden.py            # needs oracle/_ref built and /root/reference mounted
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

from make_synth_gguf import write_gguf, write_wavlm_gguf  # noqa: E402
from mio_tts_cpp_b200.benchutil import reference_clip, write_wav16  # noqa: E402
from oracle.refbind import RefLib  # noqa: E402

# reference-audio path: (fixture, wavlm preset, global-encoder preset, seconds, clip seed, ssl stride kept)
REF_CASES = [
    ("refemb_tiny", "tiny", "tiny", 2.0, 3, 1),
    ("refemb_tiny_short", "tiny", "tiny", 0.31, 5, 1),
    ("refemb_base", "base", "default", 3.0, 3, 4),
]

CASES = [  # (fixture, preset, T, codes seed)
    ("tiny_T50", "tiny", 50, 42),
    ("tiny_T1", "tiny", 1, 7),
    ("tiny_T133", "tiny", 133, 9),
    ("upsampler_tiny_T20", "upsampler_tiny", 20, 42),
    ("default_T24", "default", 24, 42),
    ("upsampler_T12", "upsampler", 12, 42),
    # frames != 2T: the lerp branch of ggml_interpolate (three frames per token; 1.5 with an odd T)
    ("tiny_lerp3_T37", "tiny_lerp3", 37, 42),
    ("tiny_lerp15_T41", "tiny_lerp15", 41, 42),
    # dynamic_global = 0: plain-LayerNorm decoder, no embedding accepted
    ("tiny_static_T50", "tiny_static", 50, 42),
]
STATIC_PRESETS = {"tiny_static"}

# BASELINE configs 1 / 2 at full size (default hyper-parameters, 400 codes, en_female): utterances 0 / 17 / 63 of the
# bench batch (codes seed 42 + i).  Reduced: every SPEC_STRIDE-th spectrogram frame; the whole waveform for utterance 0,
# three 1-second windows for the others.
FULL_CASES = [("default_T400", 42, None), ("default_T400_u17", 59, 24000), ("default_T400_u63", 105, 24000)]
SPEC_STRIDE = 8


def main():
    import gguf

    R = RefLib()
    emb = np.array(gguf.GGUFReader("/root/reference/models/en_female.emb.gguf").tensors[0].data, dtype=np.float32)
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    np.save(os.path.join(out_dir, "en_female_emb.npy"), emb)
    cache = {}
    for name, preset, T, seed in CASES:
        if preset not in cache:
            path = f"/tmp/mio_golden_{preset}.gguf"
            cache[preset] = (path, write_gguf(path, preset, 0)[0])
        path, sha = cache[preset]
        codes = np.random.default_rng(seed).integers(0, 12800, T).astype(np.int32)
        e = None if preset in STATIC_PRESETS else emb
        h = R.open_decoder(path)
        spec = R.decode_spec(h, codes, e, n_threads=1, flash=0)
        R.close_decoder(h)
        ctx = R.open_context(path)
        audio, sr = R.synthesize(ctx, codes, e, n_threads=1, flash=0)
        R.close_context(ctx)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), preset=preset, weights_sha256=sha, codes=codes,
                            spec=spec, audio=audio, sample_rate=sr)
        print(name, spec.shape, audio.shape, sr, float(np.abs(audio).max()))
    if "default" not in cache:
        cache["default"] = ("/tmp/mio_golden_default.gguf", write_gguf("/tmp/mio_golden_default.gguf", "default", 0)[0])
    path, sha = cache["default"]
    ctx = R.open_context(path)
    h = R.open_decoder(path)
    for name, seed, win in FULL_CASES:
        codes = np.random.default_rng(seed).integers(0, 12800, 400).astype(np.int32)
        spec = R.decode_spec(h, codes, emb, n_threads=8, flash=0)
        audio, sr = R.synthesize(ctx, codes, emb, n_threads=8, flash=0)
        if win is None:
            starts, kept = np.array([0]), audio
        else:
            starts = np.array([0, (audio.size // 2 // win) * win, audio.size - win])
            kept = np.concatenate([audio[s:s + win] for s in starts])
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), preset="default", weights_sha256=sha, codes=codes,
                            spec_stride=SPEC_STRIDE, n_frames=spec.shape[0], spec=spec[::SPEC_STRIDE],
                            audio_starts=starts, audio_window=win if win else audio.size, audio=kept,
                            n_audio=audio.size, sample_rate=sr, audio_peak=float(np.abs(audio).max()))
        print(name, spec.shape, audio.shape, sr, float(np.abs(audio).max()))
    R.close_decoder(h)
    R.close_context(ctx)
    # iSTFT-only fixture: the reference's embd_to_audio on BASELINE config-3 style spectra
    rng = np.random.default_rng(7)
    for n_fft, hop, frames in ((1920, 480, 12), (392, 98, 30), (120, 30, 5), (2048, 512, 3), (1280, 320, 4)):
        nf = n_fft // 2 + 1
        spec = np.concatenate([rng.normal(-2, 1, (frames, nf)), rng.uniform(-np.pi, np.pi, (frames, nf))], 1)
        spec = spec.astype(np.float32)
        spec[0, 0] = 6.0  # exercises the magnitude clip at 100
        audio = R.istft(spec, n_fft, hop, 1)
        np.savez_compressed(os.path.join(out_dir, f"istft_{n_fft}_{hop}.npz"), spec=spec, audio=audio,
                            n_fft=n_fft, hop=hop)
        print("istft", n_fft, hop, audio.shape)


def main_reference_path():
    """wavlm_extractor::extract_ssl_features + encode_global_embedding_from_ssl + mio_tts_reference_to_embedding of the
    compiled reference on seeded clips (written as 24 kHz PCM16 WAV, the codec rate, so miniaudio does not resample)."""
    R = RefLib()
    out_dir = os.path.join(ROOT, "tests", "golden")
    for name, wl_preset, ge_preset, secs, seed, stride in REF_CASES:
        dec, wl, wav = f"/tmp/mio_golden_dec_{ge_preset}.gguf", f"/tmp/mio_golden_wavlm_{wl_preset}.gguf", f"/tmp/mio_golden_{name}.wav"
        write_gguf(dec, "tiny", 0, ge_preset)
        write_wavlm_gguf(wl, wl_preset, 0)
        write_wav16(wav, reference_clip(secs, 24000, seed), 24000)
        h = R.open_wavlm(wl)
        ssl = R.wavlm_extract(h, wav, 24000, 20.0, 4, 0)
        R.close_wavlm(h)
        d = R.open_decoder(dec)
        emb_from_ssl = R.encode_global(d, ssl, 4)
        R.close_decoder(d)
        ctx = R.open_context(dec, wl)
        emb = R.reference_to_embedding(ctx, wav, 4)
        R.close_context(ctx)
        assert np.array_equal(emb, emb_from_ssl)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), wavlm_preset=wl_preset, ge_preset=ge_preset, seconds=secs,
                            clip_seed=seed, ssl_stride=stride, n_frames=ssl.shape[0], ssl=ssl[::stride].astype(np.float32),
                            ssl_row_norm=np.linalg.norm(ssl, axis=1).astype(np.float32), embedding=emb)
        print(name, ssl.shape, float(ssl.std()), emb[:4])


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "reference":
        main_reference_path()
    else:
        main()
        main_reference_path()
