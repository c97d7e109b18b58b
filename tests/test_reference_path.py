"""Reference-audio -> speaker-embedding path: WavLM SSL extractor + global encoder.

CPU part: the numpy restatement (oracle/wavlm_np.py) is pinned against the committed outputs of the unmodified
reference (tests/golden/refemb_*.npz, tools/gen_golden.py reference) and, when oracle/_ref is present, against the
live reference.  GPU part (-m gpu): the CUDA path through the C ABI against the oracle and the same fixtures.

Tolerances (floating point, measured): oracle vs reference 64-67 dB on the SSL features (fp16 GELU-table and conv
rounding-boundary flips), 94 dB on the embedding.  CUDA (bf16 linears, fp16 convs, fp32 accumulate) vs reference:
48-49 dB SSL, 51-52 dB embedding, 60+ dB for the global encoder alone.  Gates are set ~8 dB under the measurements.
"""
import os

import numpy as np
import pytest

from conftest import load_golden, snr_db
from mio_tts_cpp_b200.benchutil import reference_clip, write_wav16
from oracle import miocodec_np as O
from oracle import wavlm_np as W

REF_CASES = ["refemb_tiny", "refemb_tiny_short", "refemb_base"]


@pytest.fixture(scope="session")
def ref_models(tmp_path_factory):
    """(decoder GGUF with global-encoder tensors, WavLM GGUF) per (wavlm preset, ge preset), synthetic seed 0."""
    from make_synth_gguf import write_gguf, write_wavlm_gguf

    d = tmp_path_factory.mktemp("refgguf")
    cache = {}

    def get(wl_preset, ge_preset):
        key = (wl_preset, ge_preset)
        if key not in cache:
            dec, wl = str(d / f"dec_{ge_preset}.gguf"), str(d / f"wavlm_{wl_preset}.gguf")
            write_gguf(dec, "tiny", 0, ge_preset)
            write_wavlm_gguf(wl, wl_preset, 0)
            cache[key] = (dec, wl)
        return cache[key]

    return get


def _case(name, ref_models):
    z = load_golden(name)
    dec, wl = ref_models(str(z["wavlm_preset"]), str(z["ge_preset"]))
    pcm = reference_clip(float(z["seconds"]), 24000, int(z["clip_seed"]))
    return z, dec, wl, pcm


# --------------------------------------------------------------------------------------------------------
# CPU: oracle pins
# --------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", REF_CASES)
def test_oracle_matches_reference_fixture(name, ref_models):
    z, dec, wl, pcm = _case(name, ref_models)
    ssl = W.extract_ssl_from_pcm(W.WavLM(wl), pcm, 24000)
    assert ssl.shape[0] == int(z["n_frames"])
    st = int(z["ssl_stride"])
    assert snr_db(ssl[::st], z["ssl"]) > 58.0
    assert np.allclose(np.linalg.norm(ssl, axis=1), z["ssl_row_norm"], rtol=2e-3)
    emb = W.global_embedding(O.Model(dec), ssl)
    assert emb.shape == z["embedding"].shape
    assert snr_db(emb, z["embedding"]) > 60.0
    assert np.abs(emb - z["embedding"]).max() < 5e-3


def test_wav_reader_and_front_end(tmp_path):
    pcm = reference_clip(0.5, 24000, 11)
    p = str(tmp_path / "clip.wav")
    write_wav16(p, pcm, 24000)
    x, sr = W.read_wav_mono(p)
    assert sr == 24000 and np.array_equal(x, pcm)
    y = W.prepare_waveform(x, 24000, 16000, 20.0)
    assert y.size == 8000 and abs(float(np.abs(W.normalize_audio(x)).max()) - 1.0) < 1e-6
    assert W.prepare_waveform(x, 24000, 16000, 0.25).size == 4000           # max_reference_seconds truncates first
    # bucket function: exact below max_exact, log-spaced above, sign selects the upper half (wavlm-extractor.cpp:257-279)
    b = W.relative_position_bucket(np.array([0, -1, 1, -79, 79, -80, 80, -5000, 5000]), 320, 800)
    assert list(b[:5]) == [0, 1, 161, 79, 239] and b[5] == 80 and b[6] == 240 and b[7] == 159 and b[8] == 319


def test_oracle_vs_live_reference_path(ref_models, tmp_path):
    from oracle.refbind import RefLib, ref_lib_path

    if ref_lib_path() is None:
        pytest.skip("oracle/_ref not built in this checkout")
    R = RefLib()
    if not hasattr(R.lib, "ref_wavlm_open"):
        pytest.skip("oracle/_ref predates the reference-path hooks")
    dec, wl = ref_models("tiny", "tiny")
    pcm = reference_clip(1.3, 24000, 21)
    wav = str(tmp_path / "live.wav")
    write_wav16(wav, pcm, 24000)
    h = R.open_wavlm(wl)
    ssl_ref = R.wavlm_extract(h, wav, 24000, 20.0, 2, 0)
    R.close_wavlm(h)
    ssl = W.extract_ssl_from_pcm(W.WavLM(wl), pcm, 24000)
    assert snr_db(ssl, ssl_ref) > 58.0
    d = R.open_decoder(dec)
    emb_ref = R.encode_global(d, ssl_ref, 2)
    R.close_decoder(d)
    assert snr_db(W.global_embedding(O.Model(dec), ssl_ref), emb_ref) > 80.0


# --------------------------------------------------------------------------------------------------------
# GPU: parity through the C ABI
# --------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ref_ctxs(ref_models):
    from mio_tts_cpp_b200 import MioTTS

    made = {}

    def get(wl_preset, ge_preset):
        key = (wl_preset, ge_preset)
        if key not in made:
            dec, wl = ref_models(wl_preset, ge_preset)
            made[key] = MioTTS(dec, wl)
        return made[key]

    yield get
    for m in made.values():
        m.free()


@pytest.mark.gpu
@pytest.mark.parametrize("name", REF_CASES)
def test_gpu_reference_path_vs_reference_fixture(name, ref_models, ref_ctxs, tmp_path):
    z, dec, wl, pcm = _case(name, ref_models)
    m = ref_ctxs(str(z["wavlm_preset"]), str(z["ge_preset"]))
    n0 = m.launch_count()
    ssl = m.extract_ssl(pcm, 24000)
    assert m.launch_count() > n0
    assert ssl.shape[0] == int(z["n_frames"])
    st = int(z["ssl_stride"])
    assert snr_db(ssl[::st], z["ssl"]) > 40.0                      # bf16 linears / fp16 convs: measured 48-49 dB
    # global encoder alone, fed the reference's own features (reconstructed where the fixture keeps every frame)
    if st == 1:
        emb_ge = m.ssl_to_embedding(z["ssl"])
        assert snr_db(emb_ge, z["embedding"]) > 50.0               # measured 60-62 dB
    # the drop-in call: WAV file in, malloc'd embedding out (mio_tts_reference_to_embedding, lib.cpp:1048)
    wav = str(tmp_path / (name + ".wav"))
    write_wav16(wav, pcm, 24000)
    emb = m.reference_to_embedding(wav)
    assert emb.shape == z["embedding"].shape
    assert snr_db(emb, z["embedding"]) > 42.0                      # measured 51-52 dB
    assert np.abs(emb - z["embedding"]).max() < 3e-2


@pytest.mark.gpu
def test_gpu_reference_batch_ragged_and_resampled(ref_models, ref_ctxs):
    dec, wl = ref_models("tiny", "tiny")
    m = ref_ctxs("tiny", "tiny")
    wm, om = W.WavLM(wl), O.Model(dec)
    clips = [reference_clip(s, 24000, 30 + i) for i, s in enumerate((0.9, 2.0, 0.05, 1.37))]
    embs = m.reference_pcm_to_embedding_batch(clips, 24000)
    assert len(embs) == 4
    for pcm, e in zip(clips, embs):
        ref = W.global_embedding(om, W.extract_ssl_from_pcm(wm, pcm, 24000))
        assert snr_db(e, ref) > 40.0
        single = m.reference_pcm_to_embedding_batch([pcm], 24000)[0]
        assert np.array_equal(single, e), "a clip's embedding must not depend on its batch neighbours"
    # max_reference_seconds truncates the source before normalisation (wavlm-extractor.cpp:562-580)
    p = m.default_params()
    p.max_reference_seconds = 0.5
    e_cut = m.reference_pcm_to_embedding_batch([clips[1]], 24000, p)[0]
    assert np.array_equal(e_cut, m.reference_pcm_to_embedding_batch([clips[1][:12000]], 24000)[0])
    # a clip already at the WavLM rate skips the resampler; one at 48 kHz goes through it
    for sr in (16000, 48000):
        pcm = reference_clip(0.8, sr, 44)
        ref = W.global_embedding(om, W.extract_ssl_from_pcm(wm, pcm, sr))
        assert snr_db(m.reference_pcm_to_embedding_batch([pcm], sr)[0], ref) > 40.0


@pytest.mark.gpu
def test_gpu_reference_errors(ref_models, gguf_paths, tmp_path):
    from mio_tts_cpp_b200 import MioTTS, MioTtsError

    dec, wl = ref_models("tiny", "tiny")
    pcm = reference_clip(0.5, 24000, 2)
    wav = str(tmp_path / "e.wav")
    write_wav16(wav, pcm, 24000)
    m = MioTTS(dec)                                                 # no WavLM GGUF
    with pytest.raises(MioTtsError, match="WavLM model is not loaded"):
        m.reference_to_embedding(wav)
    with pytest.raises(MioTtsError, match="WavLM model is not loaded"):
        m.reserve_reference_workspace(5.0)
    m.free()
    plain, _ = gguf_paths("tiny")                                   # no global-encoder tensors
    m = MioTTS(plain, wl)
    with pytest.raises(MioTtsError, match="requires global_encoder tensors"):
        m.reference_to_embedding(wav)
    m.free()
    m = MioTTS(dec, wl)
    with pytest.raises(MioTtsError, match="reference_audio path is empty"):
        m.reference_to_embedding("")
    with pytest.raises(MioTtsError, match="failed to extract WavLM SSL features: failed to decode reference audio"):
        m.reference_to_embedding(str(tmp_path / "missing.wav"))
    with pytest.raises(MioTtsError, match="too short"):
        m.reference_pcm_to_embedding_batch([pcm[:100]], 24000)
    m.reserve_reference_workspace(2.0)
    assert m.estimate_reference_workspace_bytes(2.0) > 0
    c = m.clone()                                                   # clones share the weights, own their workspace
    assert np.array_equal(c.reference_to_embedding(wav), m.reference_to_embedding(wav))
    c.free()
    with pytest.raises(MioTtsError, match="failed to load WavLM GGUF"):
        MioTTS(dec, str(tmp_path / "nope.gguf"))
    m.free()
