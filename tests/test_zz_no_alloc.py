"""mio_tts_b200_pcm16_static: the allocation-free WAV16 conversion (peak rule + clamp + round; static __device__ scratch,
csrc/static_scratch.cu), through the C ABI, against the pooled entry point mio_tts_b200_audio_to_pcm16 -- bit-exact, on
two calls in a row (the second one smaller and quieter, so peak words left over from the first call would show).

The entry point has no counterpart in the reference.  Collected last on
purpose (file name).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _waves(seed, n_utt, loud):
    rng = np.random.default_rng(seed)
    out = []
    for u in range(n_utt):
        n = int(rng.integers(1, 40000))
        a = rng.standard_normal(n).astype(np.float32) * np.float32(0.2)
        if loud and u % 2 == 0:
            a *= np.float32(7.0)          # peak well above 0.98: the gain branch
        if u == 1 and n > 8:
            a[3] = np.inf                 # non-finite samples do not take part in the peak
            a[5] = np.nan
        out.append(a)
    return out


def _run_static(waves):
    import torch

    from mio_tts_cpp_b200 import lib as L

    off, total = [], 0
    for a in waves:
        off.append(total)
        total += a.size + 3               # deliberately unaligned offsets
    host = np.zeros(total, np.float32)
    for o, a in zip(off, waves):
        host[o:o + a.size] = a
    d_audio = torch.from_numpy(host).to("cuda:0")
    d_pcm = torch.full((total,), 12345, dtype=torch.int16, device="cuda:0")
    torch.cuda.synchronize()
    L.pcm16_static(0, d_audio.data_ptr(), off, [a.size for a in waves], d_pcm.data_ptr())
    pcm = d_pcm.cpu().numpy()
    return [pcm[o:o + a.size] for o, a in zip(off, waves)], pcm, off


def test_static_scratch_pcm16_matches_pooled_entry_point_twice(gguf_paths):
    from mio_tts_cpp_b200 import MioTTS
    from mio_tts_cpp_b200 import lib as L

    m = MioTTS(gguf_paths("tiny")[0])
    try:
        for seed, n_utt, loud in ((11, 9, True), (12, 4, False)):
            waves = _waves(seed, n_utt, loud)
            want = m.pcm16_from_audio(waves)
            n0 = m.launch_count()
            got, raw, off = _run_static(waves)
            assert m.launch_count() >= n0 + 2, "the static entry point launched no kernels"
            for u, (g, w) in enumerate(zip(got, want)):
                assert np.array_equal(g, w), f"seed {seed} utterance {u}: static-scratch result differs"
            # samples between utterances are not touched
            for u in range(n_utt):
                gap = off[u] + waves[u].size
                assert np.all(raw[gap:gap + 3] == 12345)
    finally:
        m.free()
    # sizes beyond the static maximum are refused with a clear message, before anything is launched
    n = 2048 + 1
    with pytest.raises(L.MioTtsError, match="exceeds the static scratch"):
        L.pcm16_static(0, 256, np.zeros(n, np.int64), np.ones(n, np.int64), 256)
