"""ctypes binding to oracle/_ref/libmio_ref_<isa>.so -- TEST INFRASTRUCTURE ONLY.

The .so is the UNMODIFIED reference (mmnga/mio-tts-cpp + its vendored ggml-cpu) compiled by
oracle/Makefile from the sources under /root/reference, plus oracle/ref_driver.cpp.  It is the
checker and the timed CPU baseline; nothing in the product path (mio_tts_cpp_b200/) may import this.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _cpu_flags():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def ref_lib_path():
    flags = _cpu_flags()
    need512 = {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512_vnni", "avx512_bf16"}
    order = ["avx512", "avx2"] if need512 <= flags else ["avx2"]
    for isa in order:
        p = os.path.join(_HERE, "_ref", f"libmio_ref_{isa}.so")
        if os.path.exists(p):
            return p
    return None


class MioTtsParams(C.Structure):
    # /root/reference/src/mio-tts-lib.h:20-25
    _fields_ = [("n_threads", C.c_int32), ("max_reference_seconds", C.c_float),
                ("wavlm_flash_attn_type", C.c_int), ("miocodec_flash_attn_type", C.c_int)]


class RefLib:
    """The reference library: public mio_tts_* C API + the driver's spectrogram / iSTFT hooks."""

    def __init__(self, path=None):
        path = path or ref_lib_path()
        if path is None:
            raise FileNotFoundError("oracle/_ref/libmio_ref_*.so not built (run `make -C oracle`)")
        self.path = path
        L = self.lib = C.CDLL(path)
        vp, cp, sz = C.c_void_p, C.c_char_p, C.c_size_t
        L.ref_decoder_open.restype = vp
        L.ref_decoder_open.argtypes = [cp, cp, sz]
        L.ref_decoder_close.argtypes = [vp]
        L.ref_decoder_decode.restype = C.c_bool
        L.ref_decoder_decode.argtypes = [vp, C.POINTER(C.c_int32), sz, C.POINTER(C.c_float), sz, C.c_int, C.c_int,
                                         C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                         cp, sz]
        L.ref_istft.restype = C.POINTER(C.c_float)
        L.ref_istft.argtypes = [C.POINTER(C.c_float), C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(sz)]
        L.ref_free.argtypes = [vp]
        if hasattr(L, "ref_wavlm_open"):
            fpp, ip = C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int)
            L.ref_wavlm_open.restype = vp
            L.ref_wavlm_open.argtypes = [cp, cp, sz]
            L.ref_wavlm_close.argtypes = [vp]
            L.ref_wavlm_extract.restype = C.c_bool
            L.ref_wavlm_extract.argtypes = [vp, cp, C.c_int, C.c_float, C.c_int, C.c_int, fpp, ip, ip, cp, sz]
            L.ref_decoder_encode_global.restype = C.c_bool
            L.ref_decoder_encode_global.argtypes = [vp, C.POINTER(C.c_float), C.c_int, C.c_int, C.c_int, fpp, ip, cp, sz]
            L.mio_tts_reference_to_embedding.restype = C.c_bool
            L.mio_tts_reference_to_embedding.argtypes = [vp, cp, MioTtsParams, fpp, C.POINTER(sz), cp, sz]
            L.mio_tts_embedding_free.argtypes = [C.POINTER(C.c_float)]
        L.ref_hw_threads.restype = C.c_int
        L.mio_tts_init_from_file.restype = vp
        L.mio_tts_init_from_file.argtypes = [cp, cp, cp, sz]
        L.mio_tts_free.argtypes = [vp]
        L.mio_tts_context_set_backend_device.restype = C.c_bool
        L.mio_tts_context_set_backend_device.argtypes = [vp, cp, cp, sz]
        L.mio_tts_synthesize.restype = C.c_bool
        L.mio_tts_synthesize.argtypes = [vp, C.POINTER(C.c_int32), sz, C.POINTER(C.c_float), sz, MioTtsParams,
                                         C.POINTER(C.POINTER(C.c_float)), C.POINTER(sz), C.POINTER(C.c_int32),
                                         cp, sz]
        L.mio_tts_audio_free.argtypes = [C.POINTER(C.c_float)]

    def hw_threads(self):
        return int(self.lib.ref_hw_threads())

    # --- decode_tokens (src/miocodec-decoder.cpp:1079) -> spec [n_frames, n_bins]
    def open_decoder(self, gguf_path):
        err = C.create_string_buffer(512)
        h = self.lib.ref_decoder_open(gguf_path.encode(), err, 512)
        if not h:
            raise RuntimeError("ref_decoder_open: " + err.value.decode())
        return h

    def close_decoder(self, h):
        self.lib.ref_decoder_close(h)

    def decode_spec(self, h, codes, emb, n_threads=1, flash=-1):
        codes = np.ascontiguousarray(codes, dtype=np.int32)
        emb = np.ascontiguousarray(emb if emb is not None else [], dtype=np.float32)   # None: static model
        out = C.POINTER(C.c_float)()
        nf, nb = C.c_int(0), C.c_int(0)
        err = C.create_string_buffer(512)
        ok = self.lib.ref_decoder_decode(h, codes.ctypes.data_as(C.POINTER(C.c_int32)), codes.size,
                                         emb.ctypes.data_as(C.POINTER(C.c_float)) if emb.size else None, emb.size,
                                         n_threads, flash,
                                         C.byref(out), C.byref(nf), C.byref(nb), err, 512)
        if not ok:
            raise RuntimeError("ref decode_tokens: " + err.value.decode())
        a = np.ctypeslib.as_array(out, shape=(nf.value, nb.value)).copy()
        self.lib.ref_free(out)
        return a

    # --- wavlm_extractor::extract_ssl_features (src/wavlm-extractor.cpp:543) -> ssl [n_frames, embed_dim]
    def open_wavlm(self, gguf_path):
        err = C.create_string_buffer(512)
        h = self.lib.ref_wavlm_open(gguf_path.encode(), err, 512)
        if not h:
            raise RuntimeError("ref_wavlm_open: " + err.value.decode())
        return h

    def close_wavlm(self, h):
        self.lib.ref_wavlm_close(h)

    def wavlm_extract(self, h, audio_path, source_sample_rate, max_seconds=20.0, n_threads=1, flash=0):
        out = C.POINTER(C.c_float)()
        nf, dim = C.c_int(0), C.c_int(0)
        err = C.create_string_buffer(512)
        ok = self.lib.ref_wavlm_extract(h, audio_path.encode(), source_sample_rate, max_seconds, n_threads, flash,
                                        C.byref(out), C.byref(nf), C.byref(dim), err, 512)
        if not ok:
            raise RuntimeError("ref extract_ssl_features: " + err.value.decode())
        if err.value:
            raise RuntimeError("ref extract_ssl_features used a fallback: " + err.value.decode())
        a = np.ctypeslib.as_array(out, shape=(nf.value, dim.value)).copy()
        self.lib.ref_free(out)
        return a

    # --- miocodec_decoder::encode_global_embedding_from_ssl (src/miocodec-decoder.cpp:824)
    def encode_global(self, h_decoder, ssl, n_threads=1):
        ssl = np.ascontiguousarray(ssl, dtype=np.float32)
        out = C.POINTER(C.c_float)()
        n = C.c_int(0)
        err = C.create_string_buffer(512)
        ok = self.lib.ref_decoder_encode_global(h_decoder, ssl.ctypes.data_as(C.POINTER(C.c_float)), ssl.shape[0],
                                                ssl.shape[1], n_threads, C.byref(out), C.byref(n), err, 512)
        if not ok:
            raise RuntimeError("ref encode_global_embedding_from_ssl: " + err.value.decode())
        if err.value:
            raise RuntimeError("ref encode_global_embedding_from_ssl fallback: " + err.value.decode())
        a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
        self.lib.ref_free(out)
        return a

    # --- mio_tts_reference_to_embedding (src/mio-tts-lib.cpp:1048)
    def reference_to_embedding(self, ctx, audio_path, n_threads=0):
        prm = MioTtsParams(n_threads, 20.0, 0, -1)
        out = C.POINTER(C.c_float)()
        n = C.c_size_t(0)
        err = C.create_string_buffer(512)
        ok = self.lib.mio_tts_reference_to_embedding(ctx, audio_path.encode(), prm, C.byref(out), C.byref(n), err, 512)
        if not ok:
            raise RuntimeError("mio_tts_reference_to_embedding: " + err.value.decode())
        a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
        self.lib.mio_tts_embedding_free(out)
        return a

    # --- embd_to_audio (src/mio-tts-lib.cpp:625)
    def istft(self, spec, n_fft, hop, n_threads=1):
        spec = np.ascontiguousarray(spec, dtype=np.float32)
        n = C.c_size_t(0)
        p = self.lib.ref_istft(spec.ctypes.data_as(C.POINTER(C.c_float)), spec.shape[0], n_fft, hop, n_threads,
                               C.byref(n))
        a = np.ctypeslib.as_array(p, shape=(max(1, n.value),)).copy()[: n.value]
        self.lib.ref_free(p)
        return a

    # --- mio_tts_synthesize (src/mio-tts-lib.cpp:1182)
    def open_context(self, gguf_path, wavlm_path=None):
        err = C.create_string_buffer(512)
        ctx = self.lib.mio_tts_init_from_file(gguf_path.encode(), wavlm_path.encode() if wavlm_path else None, err, 512)
        if not ctx:
            raise RuntimeError("mio_tts_init_from_file: " + err.value.decode())
        if not self.lib.mio_tts_context_set_backend_device(ctx, b"CPU", err, 512):
            raise RuntimeError(err.value.decode())
        return ctx

    def close_context(self, ctx):
        self.lib.mio_tts_free(ctx)

    def synthesize(self, ctx, codes, emb, n_threads=0, flash=-1):
        codes = np.ascontiguousarray(codes, dtype=np.int32)
        emb = np.ascontiguousarray(emb if emb is not None else [], dtype=np.float32)   # None: static model
        prm = MioTtsParams(n_threads, 20.0, -1, flash)
        out = C.POINTER(C.c_float)()
        n, sr = C.c_size_t(0), C.c_int32(0)
        err = C.create_string_buffer(512)
        ok = self.lib.mio_tts_synthesize(ctx, codes.ctypes.data_as(C.POINTER(C.c_int32)), codes.size,
                                         emb.ctypes.data_as(C.POINTER(C.c_float)) if emb.size else None, emb.size, prm,
                                         C.byref(out), C.byref(n), C.byref(sr), err, 512)
        if not ok:
            raise RuntimeError("mio_tts_synthesize: " + err.value.decode())
        a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
        self.lib.mio_tts_audio_free(out)
        return a, sr.value
