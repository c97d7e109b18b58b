"""Shapes far from the fixtures, through the C ABI on the GPU: one very long utterance, many very short ones, and both
together -- for the default 24 kHz path and for the wave-upsampler branch.  No oracle at these sizes (the numpy restatement
would take minutes); the checks are the size-independent properties the path guarantees: sample counts
(mio-tts-lib.cpp:638-641, 742-744), finite output, peak <= 0.98 after the normaliser (:1299-1310), and bit-identity of an
utterance decoded alone and inside the batch (no kernel mixes utterances: contiguous item ranges of the A-stationary GEMM,
tail-split tiles, polyphase taps across gap rows, iSTFT chunks that straddle CTA boundaries).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def models(gguf_paths):
    from mio_tts_cpp_b200 import MioTTS

    cache = {}

    def get(preset):
        if preset not in cache:
            cache[preset] = MioTTS(gguf_paths(preset)[0])
        return cache[preset]

    yield get
    for m in cache.values():
        m.free()


def _check(m, codes, embs, probe):
    batch, sr = m.synthesize_batch(codes, embs)
    assert sr == m.sample_rate() and len(batch) == len(codes)
    spt = m.samples_per_token()
    for c, a in zip(codes, batch):
        assert a.shape == (c.size * spt,)
        assert np.isfinite(a).all()
        assert float(np.abs(a).max()) <= 0.98 + 1e-6
    for u in probe:
        single, _ = m.synthesize(codes[u], embs[u])
        assert single.tobytes() == batch[u].tobytes(), f"utterance {u} ({codes[u].size} codes) differs alone / in the batch"


@pytest.mark.parametrize("preset,long_codes", [("default", 3000), ("upsampler", 1200)])
def test_one_long_utterance_among_short_ones(preset, long_codes, models, emb):
    m = models(preset)
    rng = np.random.default_rng(long_codes)
    lens = [3, long_codes, 1, 17, 2]
    codes = [rng.integers(0, 12800, n).astype(np.int32) for n in lens]
    embs = [emb * (1.0 + 0.2 * i) for i in range(len(lens))]
    _check(m, codes, embs, probe=[1, 2, 4])


@pytest.mark.parametrize("preset,n_utt", [("default", 300), ("upsampler", 150)])
def test_many_short_utterances(preset, n_utt, models, emb):
    m = models(preset)
    rng = np.random.default_rng(n_utt)
    codes = [rng.integers(0, 12800, int(rng.integers(1, 40))).astype(np.int32) for _ in range(n_utt)]
    embs = [emb] * n_utt
    _check(m, codes, embs, probe=[0, n_utt // 2, n_utt - 1])
