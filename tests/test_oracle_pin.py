"""Pins the numpy oracle (oracle/miocodec_np.py) against the reference ITSELF.

The reference has no golden vectors for this path, so the pins are (a) the committed outputs of
the unmodified reference compiled into oracle/_ref (tests/golden/*.npz, made by tools/gen_golden.py) and (b), when
oracle/_ref is present in this checkout, the live reference on fresh seeded inputs.
"""
import os

import numpy as np
import pytest

from conftest import FULL_CASES, GOLDEN_CASES, ISTFT_CASES, case_embedding, full_case_audio_windows, load_golden, snr_db
from oracle import miocodec_np as O


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_decode_matches_reference_fixture(name, gguf_paths, emb):
    z = load_golden(name)
    path, sha = gguf_paths(str(z["preset"]))
    assert sha == str(z["weights_sha256"]), "synthetic weight generator drifted from the fixture"
    m = O.Model(path)
    emb = case_embedding(z, emb)
    spec = O.decode_tokens(m, z["codes"], emb)
    assert spec.shape == z["spec"].shape
    # fp32 re-association (numpy BLAS vs ggml) plus fp16 rounding-boundary flips in the conv operands: measured
    # 66-88 dB over the six fixtures, far tighter than any reduced-precision GPU mode (41-47 dB)
    assert snr_db(spec, z["spec"]) > 60.0
    assert np.abs(spec - z["spec"]).max() < 2e-2
    audio = O.synthesize(m, z["codes"], emb)
    assert audio.shape == z["audio"].shape
    assert int(z["sample_rate"]) == m.hp["sample_rate"]
    assert snr_db(audio, z["audio"]) > 50.0


@pytest.mark.parametrize("name", FULL_CASES)
def test_full_size_decode_matches_reference_fixture(name, gguf_paths, emb):
    """Default hyper-parameters, 400 codes (BASELINE configs 1 / 2): window-65 band mask with real masking over 800
    frames, RoPE positions to 799, GroupNorm over 800 x 512."""
    z = load_golden(name)
    path, sha = gguf_paths("default")
    assert sha == str(z["weights_sha256"])
    m = O.Model(path)
    spec = O.decode_tokens(m, z["codes"], emb)
    assert spec.shape[0] == int(z["n_frames"])
    assert snr_db(spec[::int(z["spec_stride"])], z["spec"]) > 60.0
    audio = O.peak_normalize(O.istft(spec, m.hp["n_fft"], m.hp["hop_length"]))
    assert audio.size == int(z["n_audio"])
    assert snr_db(full_case_audio_windows(z, audio), z["audio"]) > 50.0


@pytest.mark.parametrize("name", ISTFT_CASES)
def test_istft_matches_reference_fixture(name):
    z = load_golden(name)
    a = O.istft(z["spec"], int(z["n_fft"]), int(z["hop"]))
    assert a.shape == z["audio"].shape
    assert np.abs(a - z["audio"]).max() < 1e-5
    assert snr_db(a, z["audio"]) > 100.0


def test_gather_is_a_row_copy(gguf_paths):
    path, _ = gguf_paths("tiny")
    m = O.Model(path)
    codes = np.array([0, 12799, 5, 5, 77], np.int32)
    g = O.gather_codebook(m, codes)
    assert g.dtype == np.float32 and np.array_equal(g, m.t["token_embd"][codes])


def test_peak_normalize_rule():
    a = np.array([0.5, -0.99, np.inf, 0.1], np.float32)
    b = O.peak_normalize(a)
    assert np.isclose(b[1], -0.95, atol=1e-6)          # 0.95 / peak over finite samples only
    assert O.peak_normalize(np.array([0.5, -0.97], np.float32))[1] == np.float32(-0.97)


def _live_ref():
    from oracle.refbind import RefLib, ref_lib_path

    if ref_lib_path() is None:
        pytest.skip("oracle/_ref not built in this checkout")
    return RefLib()


def test_oracle_vs_live_reference(gguf_paths, emb):
    R = _live_ref()
    path, _ = gguf_paths("tiny")
    m = O.Model(path)
    codes = np.random.default_rng(123).integers(0, 12800, 37).astype(np.int32)
    h = R.open_decoder(path)
    spec_ref = R.decode_spec(h, codes, emb, n_threads=1, flash=0)
    R.close_decoder(h)
    assert snr_db(O.decode_tokens(m, codes, emb), spec_ref) > 60.0
    rng = np.random.default_rng(5)
    spec = np.concatenate([rng.normal(-2, 1, (9, 61)), rng.uniform(-np.pi, np.pi, (9, 61))], 1).astype(np.float32)
    assert np.abs(O.istft(spec, 120, 30) - R.istft(spec, 120, 30, 1)).max() < 1e-5
