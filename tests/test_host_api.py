"""CPU-side tests of the C ABI: symbol coverage of include/mio_tts_b200.h and the host logic that mirrors
/root/reference/src/mio-tts-lib.cpp (codes text files, embedding GGUF, WAV16, error strings).  No GPU compute."""
import os
import re
import struct

import numpy as np
import pytest

from conftest import ROOT
from mio_tts_cpp_b200 import MioTTS, MioTtsError, lib as L


def header_functions():
    src = open(os.path.join(ROOT, "include", "mio_tts_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mio_tts_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = L.load_library()
    names = header_functions()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"libmio_tts_b200.so does not export {n}"
    # and the Python mirror binds every one of them
    assert set(names) == set(lib._signatures)


def test_reference_api_names_are_all_present():
    """Every function of the reference header (lib.h:27-170) is exported under the same name."""
    ref = ["mio_tts_default_params", "mio_tts_init_from_file", "mio_tts_context_clone", "mio_tts_free",
           "mio_tts_context_is_dynamic_global", "mio_tts_context_sample_rate", "mio_tts_context_samples_per_token",
           "mio_tts_context_set_backend_device", "mio_tts_vocab_map_init", "mio_tts_vocab_map_free",
           "mio_tts_tokens_to_codes", "mio_tts_token_to_code", "mio_tts_codes_load", "mio_tts_codes_save",
           "mio_tts_codes_free", "mio_tts_reference_to_embedding", "mio_tts_embedding_load_gguf",
           "mio_tts_embedding_save_gguf", "mio_tts_embedding_free", "mio_tts_synthesize", "mio_tts_reserve_workspace",
           "mio_tts_estimate_workspace_bytes", "mio_tts_reserve_reference_workspace",
           "mio_tts_estimate_reference_workspace_bytes", "mio_tts_audio_save_wav16", "mio_tts_audio_free"]
    assert set(ref) <= set(header_functions())


def test_default_params():
    p = L.load_library().mio_tts_default_params()   # lib.cpp:805-812
    assert (p.n_threads, p.max_reference_seconds, p.wavlm_flash_attn_type, p.miocodec_flash_attn_type) == (0, 20.0, -1, -1)


def test_codes_roundtrip_and_token_forms(tmp_path):
    p = str(tmp_path / "codes.txt")
    L.codes_save(p, [0, 17, 12799])
    assert open(p).read() == "0\n17\n12799\n"
    assert L.codes_load(p).tolist() == [0, 17, 12799]
    open(p, "w").write("<|s_5|> 7, <|s_12|>.\n  42\n")
    assert L.codes_load(p).tolist() == [5, 7, 12, 42]


# token -> parsed value (None: "failed to parse code token").  The accepted language is the reference's
# (lib.cpp:79-113): ASCII punctuation stripped on both sides except a leading '<' / '-' and a trailing '>', optional
# <|s_N|> wrapper, then exactly what std::stoll takes in full, within int32.
TOKEN_FORMS = [("17", 17), ("+17", 17), ("0017", 17), ("<|s_17|>", 17), ("(17)", 17), ('"17",', 17), ("[<|s_9|>]", 9),
               ("<|s_+4|>", 4), ("<|s_|>", None), ("<|s_1x|>", None), ("1x", None), ("x1", None), ("--1", None),
               ("<17>", None), ("1.5", None), ("99999999999", None), ("9223372036854775808", None), ("+", None),
               ("!!!", None), ("<|s_12799|>;", 12799), ("#12", 12), ("12#", 12), ("12_", 12), ("_12", 12)]


def _load_one(tmp_path, loader, tok):
    p = str(tmp_path / "t.txt")
    open(p, "w").write(tok + "\n")
    try:
        return [int(x) for x in loader(p)]
    except Exception as e:
        return str(e).split(": ")[0] if "failed to parse" in str(e) else str(e)


def test_code_token_language(tmp_path):
    for tok, want in TOKEN_FORMS:
        got = _load_one(tmp_path, L.codes_load, tok)
        if tok == "!!!":
            assert got == "codes input file is empty" or got == "failed to parse code token", (tok, got)
        else:
            assert got == ([want] if want is not None else "failed to parse code token"), (tok, got)


def test_code_token_language_matches_live_reference(tmp_path):
    """Same strings through the compiled reference's own mio_tts_codes_load when oracle/_ref is in this checkout."""
    import ctypes as C

    from oracle.refbind import RefLib, ref_lib_path

    if ref_lib_path() is None:
        pytest.skip("oracle/_ref not built in this checkout")
    R = RefLib().lib
    R.mio_tts_codes_load.restype = C.c_bool
    R.mio_tts_codes_load.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_size_t), C.c_char_p, C.c_size_t]
    R.mio_tts_codes_free.argtypes = [C.POINTER(C.c_int32)]

    def ref_load(path):
        out, n, err = C.POINTER(C.c_int32)(), C.c_size_t(0), C.create_string_buffer(256)
        if not R.mio_tts_codes_load(path.encode(), C.byref(out), C.byref(n), err, 256):
            raise RuntimeError(err.value.decode())
        v = [out[i] for i in range(n.value)]
        R.mio_tts_codes_free(out)
        return v

    extra = ["-0", "<|s_-0|>", "0x10", "1e3", "<|s_5|", "|s_5|>", "<|s_ 5|>", "12-", "-12-", "'<|s_3|>'", "<<|s_3|>>"]
    for tok in [t for t, _ in TOKEN_FORMS] + extra:
        assert _load_one(tmp_path, L.codes_load, tok) == _load_one(tmp_path, ref_load, tok), tok


@pytest.mark.parametrize("text,msg", [("", "codes input file is empty"), ("12800\n", "code id out of range in input file"),
                                      ("-1\n", "code id out of range in input file"),
                                      ("abc\n", "failed to parse code token: abc")])
def test_codes_load_errors(tmp_path, text, msg):
    p = str(tmp_path / "c.txt")
    open(p, "w").write(text)
    with pytest.raises(MioTtsError, match=re.escape(msg)):
        L.codes_load(p)


def test_codes_path_errors(tmp_path):
    with pytest.raises(MioTtsError, match="codes input path is empty"):
        L.codes_load("")
    with pytest.raises(MioTtsError, match="failed to open codes input file"):
        L.codes_load(str(tmp_path / "missing.txt"))
    with pytest.raises(MioTtsError, match="codes output path is empty"):
        L.codes_save("", [1])
    with pytest.raises(MioTtsError, match="codes are empty"):
        L.codes_save(str(tmp_path / "x.txt"), [])


def test_embedding_gguf_roundtrip_and_layout(tmp_path, emb):
    import gguf

    p = str(tmp_path / "e.emb.gguf")
    L.embedding_save_gguf(p, emb)
    assert np.array_equal(L.embedding_load_gguf(p), emb)
    r = gguf.GGUFReader(p)   # same KVs / tensor name as save_embedding_gguf_vec, lib.cpp:268-286
    assert [t.name for t in r.tensors] == ["mio.global_embedding"]
    assert np.array_equal(np.array(r.tensors[0].data, np.float32), emb)
    assert bytes(r.fields["general.architecture"].parts[-1]).decode() == "mio-embedding"
    assert int(r.fields["mio.embedding.dim"].parts[-1][0]) == emb.size
    # a file written by the gguf python package (as scripts/convert_preset_embedding_to_gguf.py does) loads too
    p2 = str(tmp_path / "w.gguf")
    w = gguf.GGUFWriter(p2, "mio-embedding")
    w.add_tensor("mio.global_embedding", emb)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    assert np.array_equal(L.embedding_load_gguf(p2), emb)


def test_embedding_errors(tmp_path):
    with pytest.raises(MioTtsError, match="embedding input path is empty"):
        L.embedding_load_gguf("")
    with pytest.raises(MioTtsError, match="failed to open embedding GGUF"):
        L.embedding_load_gguf(str(tmp_path / "nope.gguf"))
    with pytest.raises(MioTtsError, match="embedding is empty"):
        L.embedding_save_gguf(str(tmp_path / "x.gguf"), np.zeros(0, np.float32))


def test_wav16_bytes(tmp_path):
    p = str(tmp_path / "a.wav")
    audio = np.array([0.0, 0.5, -0.5, 2.0, -2.0, 1e-5], np.float32)
    L.audio_save_wav16(p, audio, 24000)
    b = open(p, "rb").read()
    assert len(b) == 44 + 2 * audio.size and b[:4] == b"RIFF" and b[8:16] == b"WAVEfmt "
    fmt = struct.unpack("<IHHIIHH", b[16:36])
    assert fmt == (16, 1, 1, 24000, 48000, 2, 16)            # lib.cpp:49-63, 770-777
    assert struct.unpack("<I", b[40:44])[0] == 2 * audio.size
    pcm = np.frombuffer(b[44:], np.int16)
    assert pcm.tolist() == [0, 16384, -16384, 32767, -32767, 0]   # clamp + lrintf(x * 32767), lib.cpp:781-784
    with pytest.raises(MioTtsError, match="audio is empty"):
        L.audio_save_wav16(p, np.zeros(0, np.float32), 24000)


def test_init_errors():
    with pytest.raises(MioTtsError, match="model_vocoder path is empty"):
        MioTTS("")
    lib = L.load_library()
    import ctypes as C
    err = C.create_string_buffer(64)
    assert not lib.mio_tts_context_clone(None, err, 64) and err.value == b"source context is null"
    assert lib.mio_tts_context_sample_rate(None) == 0 and lib.mio_tts_context_samples_per_token(None) == 0
    assert not lib.mio_tts_context_is_dynamic_global(None)
    out, n, sr = C.POINTER(C.c_float)(), C.c_size_t(0), C.c_int32(0)
    ok = lib.mio_tts_synthesize(None, None, 0, None, 0, lib.mio_tts_default_params(), C.byref(out), C.byref(n),
                                C.byref(sr), err, 64)
    assert not ok and err.value == b"context is null"
    small = C.create_string_buffer(8)                      # messages are truncated, never overflow (set_err)
    assert not lib.mio_tts_context_clone(None, small, 8) and small.value == b"source "


def test_no_gpu_fails_loudly(gguf_paths):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    path, _ = gguf_paths("tiny")
    with pytest.raises(MioTtsError, match="no CUDA device available"):
        MioTTS(path)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "mio_tts_cpp_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(d, f), errors="ignore").read()
                for needle in ("from oracle", "import oracle", "miocodec_np", "refbind", "libmio_ref", "oracle/"):
                    assert needle not in src, f"{f} references the test oracle ({needle})"
