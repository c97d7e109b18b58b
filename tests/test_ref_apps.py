"""The reference's OWN programs on top of libmio_tts_b200.so (BASELINE config 5).

tools/build_ref_apps.sh compiles /root/reference/src/tts-mio-cli.cpp and tts-mio-server.cpp UNMODIFIED against the
reference's own header and links them to this repo's library instead of the reference's static mio-tts-lib
(INTEGRATION.md section 1).  The binaries are built in the authoring container (the reference tree does not travel) and
ride to the GPU box under tools/ref_apps/_build/bin; every test skips when they are absent.

  * CLI, codes-file mode (/root/reference/src/tts-mio-cli.cpp:1122-1127, 1298-1317): WAV bytes equal this library's
    own mio_tts_synthesize + mio_tts_audio_save_wav16 bit for bit, and agree with the WAV written by the reference CLI
    on the reference's CPU library (llama-tts-mio-ref) within the bf16 decoder tolerance, stated below in int16 steps.
  * server (/root/reference/src/tts-mio-server.cpp:2657-2746 slot set-up, :2153-2453 request path, inline "codes"
    :2105-2110): /health, concurrent POST /mio/tts with inline codes, bodies equal the library's WAV16 rule bit for bit,
    error isolation (one bad request fails alone), with and without MIO_TTS_B200_BATCHING.
"""
import json
import os
import socket
import subprocess
import sys
import threading
import time
import urllib.error
import urllib.request
import wave

import numpy as np
import pytest

from conftest import ROOT, load_golden, snr_db

BIN = os.path.join(ROOT, "tools", "ref_apps", "_build", "bin")
CLI, CLI_REF, SERVER = (os.path.join(BIN, n) for n in ("llama-tts-mio", "llama-tts-mio-ref", "mio-tts-server"))


def _env(**extra):
    ld = os.pathsep.join([os.path.join(ROOT, "mio_tts_cpp_b200"), BIN, os.environ.get("LD_LIBRARY_PATH", "")])
    return dict(os.environ, LD_LIBRARY_PATH=ld, **extra)


def _need(*paths):
    for p in paths:
        if not os.path.exists(p):
            pytest.skip(f"{os.path.relpath(p, ROOT)} not built (tools/build_ref_apps.sh needs /root/reference)")


def _read_wav(path):
    with wave.open(path) as w:
        assert w.getnchannels() == 1 and w.getsampwidth() == 2
        return np.frombuffer(w.readframes(w.getnframes()), np.int16), w.getframerate()


def _inputs(tmp_path, emb, codes):
    from mio_tts_cpp_b200 import lib as L

    e, c = str(tmp_path / "spk.emb.gguf"), str(tmp_path / "codes.txt")
    L.embedding_save_gguf(e, emb)
    L.codes_save(c, np.asarray(codes, np.int32))
    return e, c


# ------------------------------------------------------------------------------------------------- CPU box
def test_swapped_cli_resolves_every_symbol_and_fails_loudly_without_a_gpu(tmp_path, gguf_paths, emb):
    """Dynamic linking succeeds (all mio_tts_* imports of the unmodified CLI are exported by the library); with no
    device the program reports the library's message instead of falling back to a CPU path."""
    _need(CLI)
    r = subprocess.run([CLI, "--help"], env=_env(), capture_output=True, text=True, timeout=120)
    assert "model-vocoder" in r.stdout + r.stderr
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    e, c = _inputs(tmp_path, emb, [1, 2, 3])
    r = subprocess.run([CLI, "-mv", gguf_paths("tiny")[0], "--tts-mio-codes-in", c, "-emb", e, "-o", str(tmp_path / "o.wav")],
                       env=_env(), capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "no CUDA device available" in r.stderr


def test_reference_cli_on_its_own_library_matches_the_fixture(tmp_path, gguf_paths, emb):
    """llama-tts-mio-ref (reference CLI + reference CPU library, same build recipe) reproduces the committed fixture:
    the comparison partner of the GPU test below is the real thing."""
    _need(CLI_REF)
    z = load_golden("tiny_T50")
    e, c = _inputs(tmp_path, emb, z["codes"])
    out = str(tmp_path / "ref.wav")
    r = subprocess.run([CLI_REF, "-mv", gguf_paths("tiny")[0], "--tts-mio-codes-in", c, "-emb", e, "-o", out],
                       env=_env(), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    pcm, sr = _read_wav(out)
    assert sr == 24000 and pcm.size == z["audio"].size
    assert np.abs(pcm.astype(np.int32) - np.rint(np.clip(z["audio"], -1, 1) * 32767).astype(np.int32)).max() <= 2


# ------------------------------------------------------------------------------------------------- GPU box
@pytest.mark.gpu
@pytest.mark.parametrize("preset,fixture", [("tiny", "tiny_T50"), ("default", "default_T400")])
def test_swapped_cli_codes_file_mode(preset, fixture, tmp_path, gguf_paths, emb):
    _need(CLI, CLI_REF)
    from mio_tts_cpp_b200 import MioTTS
    from mio_tts_cpp_b200 import lib as L

    z = load_golden(fixture)
    path = gguf_paths(preset)[0]
    e, c = _inputs(tmp_path, emb, z["codes"])
    out, ref_out, lib_out = (str(tmp_path / n) for n in ("b200.wav", "ref.wav", "lib.wav"))
    r = subprocess.run([CLI, "-mv", path, "--tts-mio-codes-in", c, "-emb", e, "-o", out], env=_env(), capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "audio written to file" in r.stderr
    m = MioTTS(path)
    audio, sr = m.synthesize(z["codes"], emb)
    m.free()
    L.audio_save_wav16(lib_out, audio, sr)
    assert open(out, "rb").read() == open(lib_out, "rb").read()          # the program adds nothing of its own
    r = subprocess.run([CLI_REF, "-mv", path, "--tts-mio-codes-in", c, "-emb", e, "-o", ref_out, "--threads", "16"],
                       env=_env(), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    a, sr_a = _read_wav(out)
    b, sr_b = _read_wav(ref_out)
    assert sr_a == sr_b == 24000 and a.size == b.size == z["codes"].size * 960 // (16 if preset == "tiny" else 1)
    db = snr_db(a.astype(np.float64), b.astype(np.float64))
    step = np.abs(a.astype(np.int32) - b.astype(np.int32)).max()
    print(f"{fixture}: swapped CLI vs reference CLI: {db:.1f} dB, max {step} int16 steps (reference peak {np.abs(b).max()})")
    # fp16 operands against the reference's fp32 CPU decode, in WAV terms: >= 44 dB and no sample further than 33 int16
    # steps (1e-3 of full scale) from the reference's (measured: 51.6-54.9 dB, 9-15 steps)
    assert db >= 44.0 and step <= 33


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class Server:
    def __init__(self, model, emb_path, parallel, devices, log_path, **env):
        self.port = _free_port()
        self.log = open(log_path, "w")
        cmd = [SERVER, "-mv", model, "--host", "127.0.0.1", "--port", str(self.port), "--parallel", str(parallel),
               "--reference-file-json", json.dumps({"key": "spk", "path": emb_path})]
        if devices:
            cmd += ["--mio-backend-devices", devices]
        self.p = subprocess.Popen(cmd, env=_env(**env), stdout=self.log, stderr=subprocess.STDOUT)
        t0 = time.time()
        while True:
            if self.p.poll() is not None:
                raise RuntimeError("server exited: " + open(log_path).read()[-2000:])
            try:
                h = json.load(urllib.request.urlopen(f"http://127.0.0.1:{self.port}/health", timeout=2))
                if h.get("status") == "ok":
                    self.health = h
                    return
            except (urllib.error.URLError, ConnectionError, socket.timeout):
                pass
            if time.time() - t0 > 120:
                self.close()
                raise RuntimeError("server did not come up: " + open(log_path).read()[-2000:])
            time.sleep(0.2)

    def tts(self, codes, key="spk", route="/mio/tts/stream"):
        req = urllib.request.Request(f"http://127.0.0.1:{self.port}{route}", method="POST",
                                     data=json.dumps({"codes": [int(c) for c in codes], "reference_key": key}).encode(),
                                     headers={"Content-Type": "application/json"})
        try:
            with urllib.request.urlopen(req, timeout=120) as r:
                return r.status, r.read(), dict(r.headers)
        except urllib.error.HTTPError as e:
            return e.code, e.read(), dict(e.headers)

    def close(self):
        self.p.terminate()
        try:
            self.p.wait(timeout=20)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.log.close()


@pytest.mark.gpu
@pytest.mark.parametrize("batching", [None, "16,300,2"])
def test_swapped_server_inline_codes(batching, tmp_path, gguf_paths, emb):
    """mio-tts-server --parallel 4 on one device: slot 0 owns the context, slots 1..3 are clones
    (tts-mio-server.cpp:2724-2736).  24 concurrent requests of ragged lengths on POST /mio/tts/stream (stream_tokens
    off: the binary audio/wav response, :3901-4006); every body is the 44-byte header + the server's own float -> int16
    rule (:3965-3973) applied to what mio_tts_synthesize returns for those codes.  POST /mio/tts (:3630-3703) writes the
    WAV through mio_tts_audio_save_wav16 and answers with JSON naming the file."""
    _need(SERVER)
    from mio_tts_cpp_b200 import MioTTS

    path = gguf_paths("tiny")[0]
    e, _ = _inputs(tmp_path, emb, [1])
    env = {"MIO_TTS_B200_BATCHING": batching} if batching else {}
    srv = Server(path, e, 4, "CUDA0", str(tmp_path / "server.log"), **env)
    try:
        assert srv.health["parallel"] == 4 and srv.health["reference_cache"] >= 1
        rng = np.random.default_rng(5)
        reqs = [rng.integers(0, 12800, int(n)).astype(np.int32) for n in rng.integers(5, 300, 24)]
        reqs[7] = np.array([5, 12800, 7], np.int32)          # out of range: must fail alone
        out = [None] * len(reqs)

        def work(i):
            out[i] = srv.tts(reqs[i])

        th = [threading.Thread(target=work, args=(i,)) for i in range(len(reqs))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        m = MioTTS(path)
        for i, (status, body, hdr) in enumerate(out):
            if i == 7:
                assert status == 400 and b"out of range" in body
                continue
            assert status == 200, body[:300]
            audio, sr = m.synthesize(reqs[i], emb)
            pcm = np.rint(np.clip(audio, -1.0, 1.0) * np.float32(32767.0)).astype(np.int16)
            assert len(body) == 44 + 2 * audio.size and body[:4] == b"RIFF"
            assert int(hdr.get("X-Sample-Rate", 0)) == sr and int(hdr.get("X-Audio-Samples", 0)) == audio.size
            assert body[44:] == pcm.tobytes(), i
        status, body, _ = srv.tts(reqs[0], route="/mio/tts")
        assert status == 200, body[:300]
        j = json.loads(body)
        assert j["ok"] and j["n_audio"] == reqs[0].size * m.samples_per_token() and j["sample_rate"] == 24000
        from mio_tts_cpp_b200 import lib as L

        audio, sr = m.synthesize(reqs[0], emb)
        L.audio_save_wav16(str(tmp_path / "lib.wav"), audio, sr)
        assert open(j["output_file"], "rb").read() == open(str(tmp_path / "lib.wav"), "rb").read()
        os.remove(j["output_file"])
        m.free()
        status, body, _ = srv.tts(reqs[0], key="nobody")
        assert status == 400
    finally:
        srv.close()
    log = open(str(tmp_path / "server.log")).read()
    assert "ok=true" in log and "ok=false" in log
