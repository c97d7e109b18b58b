"""Server-side rows either side of the decode, through the C ABI on the GPU:

  * WAV16 emission (mio_tts_b200_synthesize_wav16_batch): the bytes the reference server builds from the float waveform
    (/root/reference/src/tts-mio-server.cpp:3965-3973 + build_wav_header :2013-2034; same rule as save_wav16,
    src/mio-tts-lib.cpp:758-791) -- bit-exact against that rule applied on the host to the library's own fp32 output.
  * micro-batcher (mio_tts_b200_context_enable_batching): concurrent mio_tts_synthesize calls from slot threads, each on
    its own clone like the reference server's workers (:2724-2736), are coalesced; every caller must get exactly the
    samples an un-batched call returns (bit-exact) and a bad request must fail alone with its usual message.
"""
import struct
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _wav16_rule(audio, sample_rate):
    """The reference's conversion, restated: clamp, x * 32767 in fp32, lrintf (round half to even), RIFF header."""
    x = np.clip(audio.astype(np.float32), np.float32(-1.0), np.float32(1.0))
    pcm = np.rint(x * np.float32(32767.0)).astype(np.int16)
    n = pcm.size * 2
    hdr = b"RIFF" + struct.pack("<I", 36 + n) + b"WAVE" + b"fmt " + struct.pack("<IHHIIHH", 16, 1, 1, sample_rate, sample_rate * 2, 2, 16)
    return hdr + b"data" + struct.pack("<I", n) + pcm.tobytes()


@pytest.fixture(scope="module")
def tiny(gguf_paths):
    from mio_tts_cpp_b200 import MioTTS

    m = MioTTS(gguf_paths("tiny")[0])
    yield m
    m.free()


def _workload(n_utt, seed, lo=3, hi=120):
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 12800, int(rng.integers(lo, hi))).astype(np.int32) for _ in range(n_utt)]


def test_wav16_bytes_match_the_reference_rule(tiny, emb, tmp_path):
    codes = _workload(5, 1) + [np.array([7], np.int32)]
    embs = [emb * s for s in (1.0, 0.5, 2.0, 1.0, 3.0, 1.0)]          # different loudness: some cross the 0.98 peak rule
    audio, sr = tiny.synthesize_batch(codes, embs)
    n0 = tiny.launch_count()
    wavs = tiny.synthesize_wav16_batch(codes, embs)
    assert tiny.launch_count() > n0
    for a, w in zip(audio, wavs):
        assert len(w) == 44 + 2 * a.size
        assert w == _wav16_rule(a, sr)
    # and the library's own file writer (mio_tts_audio_save_wav16, lib.cpp:758-791) produces the same bytes
    from mio_tts_cpp_b200 import lib as L

    p = str(tmp_path / "a.wav")
    L.audio_save_wav16(p, audio[0], sr)
    assert open(p, "rb").read() == wavs[0]
    # single call == batch member
    assert tiny.synthesize_wav16_batch([codes[2]], [embs[2]])[0] == wavs[2]


def test_wav16_synthetic_extremes(tiny):
    """The conversion kernel alone over values that exercise clamp, ties-to-even and the 0.95/peak gain."""
    x = np.array([0.0, 1.0, -1.0, 1.5, -1.5, 0.5 / 32767, 1.5 / 32767, 2.5 / 32767, -0.5 / 32767, 0.999999, 1e-9, -0.3333333], np.float32)
    got = tiny.pcm16_from_audio([x, x * np.float32(0.5)])
    for a, g in zip([x, x * np.float32(0.5)], got):
        peak = float(np.abs(a).max())
        ref = a * np.float32(0.95 / peak) if peak > 0.98 else a      # lib.cpp:1299-1310
        assert g.tobytes() == _wav16_rule(ref, 24000)[44:]


def test_micro_batcher_coalesces_and_is_bit_exact(tiny, emb):
    n_slots, rounds = 8, 3
    work = [_workload(rounds, 100 + s) for s in range(n_slots)]
    expect = [[tiny.synthesize(c, emb)[0] for c in w] for w in work]              # un-batched, one by one
    tiny.enable_batching(64, 2000, 2)
    try:
        slots = [tiny.clone() for _ in range(n_slots)]                            # made AFTER enabling: they share the batcher
        got = [[None] * rounds for _ in range(n_slots)]
        errs = []
        gate = threading.Barrier(n_slots)

        def slot(i):
            try:
                for r in range(rounds):
                    gate.wait()
                    got[i][r] = slots[i].synthesize(work[i][r], emb)[0]
            except Exception as e:  # noqa: BLE001
                errs.append(e)
                gate.abort()

        th = [threading.Thread(target=slot, args=(i,)) for i in range(n_slots)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errs, errs
        for i in range(n_slots):
            for r in range(rounds):
                assert np.array_equal(got[i][r], expect[i][r]), (i, r)
        st = tiny.batching_stats()
        assert st["requests"] == n_slots * rounds
        assert st["batches"] < st["requests"] and st["largest"] > 1, st
        for s in slots:
            s.free()
    finally:
        tiny.enable_batching(0)
    assert tiny.batching_stats() is None


def test_micro_batcher_isolates_a_bad_request(tiny, emb):
    from mio_tts_cpp_b200 import MioTtsError

    good = _workload(5, 7)
    expect = [tiny.synthesize(c, emb)[0] for c in good]
    with pytest.raises(MioTtsError) as plain:
        tiny.synthesize(np.array([5, 12800, 9], np.int32), emb)
    tiny.enable_batching(64, 5000, 1)
    try:
        slots = [tiny.clone() for _ in range(6)]
        out, bad = [None] * 5, []
        gate = threading.Barrier(6)

        def ok_slot(i):
            gate.wait()
            out[i] = slots[i].synthesize(good[i], emb)[0]

        def bad_slot():
            gate.wait()
            try:
                slots[5].synthesize(np.array([5, 12800, 9], np.int32), emb)
            except MioTtsError as e:
                bad.append(str(e))

        th = [threading.Thread(target=ok_slot, args=(i,)) for i in range(5)] + [threading.Thread(target=bad_slot)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert bad == [str(plain.value)]
        for i in range(5):
            assert np.array_equal(out[i], expect[i])
        # mixed formats never share a launch, but both complete
        w = slots[0].synthesize_wav16_batch([good[0]], [emb])[0]
        assert w == _wav16_rule(expect[0], tiny.sample_rate())
        for s in slots:
            s.free()
    finally:
        tiny.enable_batching(0)


def test_graph_replay_is_bit_exact(tiny, emb):
    """Small repeated batch shapes are replayed from a captured CUDA graph (engine::run): the first call launches the
    kernels one by one, the second captures them, later ones replay -- all must return the same samples, the launch
    counter must keep counting the kernels inside the graph, and a different shape in between must not disturb it."""
    c = _workload(1, 55, 40, 41)[0]
    other = _workload(1, 56, 90, 91)[0]
    outs, counts = [], []
    for i in range(5):
        n0 = tiny.launch_count()
        outs.append(tiny.synthesize(c, emb)[0])
        counts.append(tiny.launch_count() - n0)
        if i == 2:
            tiny.synthesize(other, emb)
    for o in outs[1:]:
        assert np.array_equal(o, outs[0])
    assert len(set(counts)) == 1 and counts[0] > 20, counts
    # the same utterance inside a batch (a different shape, plain launches) still matches
    both, _ = tiny.synthesize_batch([other, c], [emb, emb])
    assert np.array_equal(both[1], outs[0])
    # WAV16 output is a different graph of the same shape
    w1 = tiny.synthesize_wav16_batch([c], [emb])[0]
    w2 = tiny.synthesize_wav16_batch([c], [emb])[0]
    w3 = tiny.synthesize_wav16_batch([c], [emb])[0]
    assert w1 == w2 == w3 == _wav16_rule(outs[0], tiny.sample_rate())
