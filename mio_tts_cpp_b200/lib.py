"""ctypes binding to libmio_tts_b200.so (include/mio_tts_b200.h).

Mirrors the reference's public C API (/root/reference/src/mio-tts-lib.h): same function names, argument meaning,
ownership (every out-buffer is released through the matching ``*_free``) and error strings, surfaced as
``MioTtsError``.  There is no fallback of any kind: a missing library raises at import-use time.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ERR_SIZE = 512


class MioTtsError(RuntimeError):
    pass


class MioTtsParams(C.Structure):
    # /root/reference/src/mio-tts-lib.h:20-25
    _fields_ = [("n_threads", C.c_int32), ("max_reference_seconds", C.c_float),
                ("wavlm_flash_attn_type", C.c_int), ("miocodec_flash_attn_type", C.c_int)]


class Profile(C.Structure):
    _fields_ = [("prenet_ms", C.c_float), ("upsample_prior_ms", C.c_float), ("decoder_ms", C.c_float),
                ("post_head_ms", C.c_float), ("istft_ms", C.c_float), ("total_ms", C.c_float),
                ("gemm_ms", C.c_double), ("gemm_flops", C.c_double), ("gemm_launches", C.c_int32),
                ("attn_ms", C.c_double), ("attn_launches", C.c_int32), ("gemm_bytes", C.c_double)]


def lib_path():
    return os.environ.get("MIO_TTS_B200_LIB", os.path.join(_HERE, "libmio_tts_b200.so"))


_LIB = None


def load_library():
    """Loads the CUDA library or raises: the product path never degrades to a CPU implementation."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise MioTtsError(f"{path} is missing: build it with `make -C mio_tts_cpp_b200/csrc` "
                          "(or __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(path)
    vp, cp, sz, i32, u64 = C.c_void_p, C.c_char_p, C.c_size_t, C.c_int32, C.c_uint64
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    fpp, ipp, szp = C.POINTER(fp), C.POINTER(ip), C.POINTER(sz)
    sig = {
        "mio_tts_default_params": (MioTtsParams, []),
        "mio_tts_init_from_file": (vp, [cp, cp, cp, sz]),
        "mio_tts_context_clone": (vp, [vp, cp, sz]),
        "mio_tts_free": (None, [vp]),
        "mio_tts_context_is_dynamic_global": (C.c_bool, [vp]),
        "mio_tts_context_sample_rate": (i32, [vp]),
        "mio_tts_context_samples_per_token": (i32, [vp]),
        "mio_tts_context_set_backend_device": (C.c_bool, [vp, cp, cp, sz]),
        "mio_tts_vocab_map_init": (vp, [vp, cp, sz]),
        "mio_tts_vocab_map_free": (None, [vp]),
        "mio_tts_tokens_to_codes": (C.c_bool, [vp, ip, sz, ipp, szp, cp, sz]),
        "mio_tts_token_to_code": (C.c_bool, [vp, i32, ip]),
        "mio_tts_codes_load": (C.c_bool, [cp, ipp, szp, cp, sz]),
        "mio_tts_codes_save": (C.c_bool, [cp, ip, sz, cp, sz]),
        "mio_tts_codes_free": (None, [ip]),
        "mio_tts_reference_to_embedding": (C.c_bool, [vp, cp, MioTtsParams, fpp, szp, cp, sz]),
        "mio_tts_embedding_load_gguf": (C.c_bool, [cp, fpp, szp, cp, sz]),
        "mio_tts_embedding_save_gguf": (C.c_bool, [cp, fp, sz, cp, sz]),
        "mio_tts_embedding_free": (None, [fp]),
        "mio_tts_synthesize": (C.c_bool, [vp, ip, sz, fp, sz, MioTtsParams, fpp, szp, ip, cp, sz]),
        "mio_tts_reserve_workspace": (C.c_bool, [vp, i32, cp, sz]),
        "mio_tts_estimate_workspace_bytes": (C.c_bool, [vp, i32, C.POINTER(u64), cp, sz]),
        "mio_tts_reserve_reference_workspace": (C.c_bool, [vp, C.c_float, cp, sz]),
        "mio_tts_estimate_reference_workspace_bytes": (C.c_bool, [vp, C.c_float, C.POINTER(u64), cp, sz]),
        "mio_tts_audio_save_wav16": (C.c_bool, [cp, fp, sz, i32, cp, sz]),
        "mio_tts_audio_free": (None, [fp]),
        "mio_tts_b200_synthesize_batch": (C.c_bool, [vp, C.POINTER(ip), szp, C.POINTER(fp), sz, sz, MioTtsParams,
                                                     fpp, szp, ip, cp, sz]),
        "mio_tts_b200_synthesize_wav16_batch": (C.c_bool, [vp, C.POINTER(ip), szp, C.POINTER(fp), sz, sz, MioTtsParams,
                                                           C.POINTER(C.POINTER(C.c_uint8)), szp, ip, cp, sz]),
        "mio_tts_b200_wav_free": (None, [C.POINTER(C.c_uint8)]),
        "mio_tts_b200_context_enable_batching": (C.c_bool, [vp, i32, i32, i32, cp, sz]),
        "mio_tts_b200_batching_stats": (C.c_bool, [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]),
        "mio_tts_b200_audio_to_pcm16": (C.c_bool, [vp, C.POINTER(fp), szp, sz, C.POINTER(C.POINTER(C.c_int16)), cp, sz]),
        "mio_tts_b200_pcm16_static": (C.c_bool, [i32, vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64), sz, vp, vp, cp, sz]),
        "mio_tts_b200_decode_spec": (C.c_bool, [vp, ip, sz, fp, sz, fpp, ip, ip, cp, sz]),
        "mio_tts_b200_debug_stage": (C.c_bool, [vp, ip, sz, fp, sz, i32, fpp, ip, ip, cp, sz]),
        "mio_tts_b200_gather": (C.c_bool, [vp, ip, sz, fp, cp, sz]),
        "mio_tts_b200_prenet_dim": (i32, [vp]),
        "mio_tts_b200_istft": (C.c_bool, [vp, fp, i32, i32, i32, fpp, szp, cp, sz]),
        "mio_tts_b200_free": (None, [vp]),
        "mio_tts_b200_set_generic_attention": (None, [vp, i32]),
        "mio_tts_b200_prepare": (C.c_bool, [vp, C.POINTER(ip), szp, C.POINTER(fp), sz, sz, cp, sz]),
        "mio_tts_b200_run_resident": (C.c_float, [vp, i32, cp, sz]),
        "mio_tts_b200_profile_resident": (C.c_bool, [vp, C.POINTER(Profile), cp, sz]),
        "mio_tts_b200_resident_samples": (u64, [vp]),
        "mio_tts_b200_gather_bench": (C.c_float, [vp, i32, cp, sz]),
        "mio_tts_b200_istft_bench": (C.c_float, [vp, i32, i32, i32, i32, i32, u64, C.POINTER(C.c_double), cp, sz]),
        "mio_tts_b200_reference_pcm_to_embedding_batch": (C.c_bool, [vp, C.POINTER(fp), szp, i32, sz, MioTtsParams, fpp, szp,
                                                                     cp, sz]),
        "mio_tts_b200_extract_ssl": (C.c_bool, [vp, fp, sz, i32, C.c_float, fpp, ip, ip, cp, sz]),
        "mio_tts_b200_ssl_to_embedding": (C.c_bool, [vp, fp, i32, i32, fpp, szp, cp, sz]),
        "mio_tts_b200_reference_bench": (C.c_float, [vp, fp, sz, i32, i32, i32, cp, sz]),
        "mio_tts_b200_launch_count": (u64, []),
        "mio_tts_b200_workspace_bytes": (u64, [vp]),
        "mio_tts_b200_weight_bytes": (u64, [vp]),
        "mio_tts_b200_last_h2d_bytes": (u64, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)   # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    L._signatures = sig
    _LIB = L
    return L


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class MioTTS:
    """One ``mio_tts_context``.  Method names follow the C API with the ``mio_tts_`` prefix dropped."""

    def __init__(self, model_vocoder, wavlm_model=None, _handle=None):
        self.L = load_library()
        self._err = C.create_string_buffer(ERR_SIZE)
        if _handle is not None:
            self.ctx = _handle
            return
        mv = model_vocoder.encode() if model_vocoder is not None else None
        wm = wavlm_model.encode() if wavlm_model else None
        self.ctx = self.L.mio_tts_init_from_file(mv, wm, self._err, ERR_SIZE)
        if not self.ctx:
            raise MioTtsError(self._err.value.decode())

    # ---- lifetime -------------------------------------------------------------------------------------
    def clone(self):
        h = self.L.mio_tts_context_clone(self.ctx, self._err, ERR_SIZE)
        if not h:
            raise MioTtsError(self._err.value.decode())
        return MioTTS(None, _handle=h)

    def free(self):
        if getattr(self, "ctx", None):
            self.L.mio_tts_free(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def _check(self, ok):
        if not ok:
            raise MioTtsError(self._err.value.decode())

    # ---- getters --------------------------------------------------------------------------------------
    def is_dynamic_global(self):
        return bool(self.L.mio_tts_context_is_dynamic_global(self.ctx))

    def sample_rate(self):
        return int(self.L.mio_tts_context_sample_rate(self.ctx))

    def samples_per_token(self):
        return int(self.L.mio_tts_context_samples_per_token(self.ctx))

    def prenet_dim(self):
        return int(self.L.mio_tts_b200_prenet_dim(self.ctx))

    def set_backend_device(self, name):
        self._check(self.L.mio_tts_context_set_backend_device(self.ctx, name.encode() if name is not None else None,
                                                              self._err, ERR_SIZE))

    def default_params(self):
        return self.L.mio_tts_default_params()

    # ---- the hot path ---------------------------------------------------------------------------------
    def synthesize(self, codes, embedding, params=None):
        codes = _i32(codes)
        emb = _f32(embedding) if embedding is not None else None
        out, n, sr = C.POINTER(C.c_float)(), C.c_size_t(0), C.c_int32(0)
        ok = self.L.mio_tts_synthesize(self.ctx, _ptr(codes, C.c_int32) if codes.size else None, codes.size,
                                       _ptr(emb, C.c_float) if emb is not None else None,
                                       emb.size if emb is not None else 0, params or self.default_params(),
                                       C.byref(out), C.byref(n), C.byref(sr), self._err, ERR_SIZE)
        self._check(ok)
        a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
        self.L.mio_tts_audio_free(out)
        return a, sr.value

    def _batch_args(self, codes_list, emb_list):
        B = len(codes_list)
        codes = [_i32(c) for c in codes_list]
        cp = (C.POINTER(C.c_int32) * B)(*[_ptr(c, C.c_int32) for c in codes])
        cn = (C.c_size_t * B)(*[c.size for c in codes])
        if emb_list is not None:
            embs = [_f32(e) for e in emb_list]
            ep = (C.POINTER(C.c_float) * B)(*[_ptr(e, C.c_float) for e in embs])
            n_emb = embs[0].size
        else:
            embs, ep, n_emb = [], None, 0
        return B, codes, cp, cn, embs, ep, n_emb

    def synthesize_batch(self, codes_list, emb_list, copy=True):
        """mio_tts_b200_synthesize_batch.  Returns ([audio...], sample_rate)."""
        B, codes, cp, cn, embs, ep, n_emb = self._batch_args(codes_list, emb_list)
        outs = (C.POINTER(C.c_float) * B)()
        ns = (C.c_size_t * B)()
        sr = C.c_int32(0)
        ok = self.L.mio_tts_b200_synthesize_batch(self.ctx, cp, cn, ep, n_emb, B, self.default_params(), outs, ns,
                                                  C.byref(sr), self._err, ERR_SIZE)
        self._check(ok)
        res = []
        for i in range(B):
            a = np.ctypeslib.as_array(outs[i], shape=(ns[i],))
            res.append(a.copy() if copy else int(ns[i]))
            self.L.mio_tts_audio_free(outs[i])
        return res, sr.value

    def synthesize_wav16_batch(self, codes_list, emb_list, copy=True):
        """mio_tts_b200_synthesize_wav16_batch.  Returns a list of `bytes` (44-byte RIFF/WAVE header + int16 PCM)."""
        B, codes, cp, cn, embs, ep, n_emb = self._batch_args(codes_list, emb_list)
        outs = (C.POINTER(C.c_uint8) * B)()
        ns = (C.c_size_t * B)()
        ok = self.L.mio_tts_b200_synthesize_wav16_batch(self.ctx, cp, cn, ep, n_emb, B, self.default_params(), outs, ns, None,
                                                        self._err, ERR_SIZE)
        self._check(ok)
        res = []
        for i in range(B):
            res.append(C.string_at(outs[i], ns[i]) if copy else int(ns[i]))
            self.L.mio_tts_b200_wav_free(outs[i])
        return res

    def pcm16_from_audio(self, audio_list):
        arrs = [_f32(a) for a in audio_list]
        b = len(arrs)
        ptrs = (C.POINTER(C.c_float) * b)(*[_ptr(a, C.c_float) for a in arrs])
        ns = (C.c_size_t * b)(*[a.size for a in arrs])
        outs = (C.POINTER(C.c_int16) * b)()
        self._check(self.L.mio_tts_b200_audio_to_pcm16(self.ctx, ptrs, ns, b, outs, self._err, ERR_SIZE))
        res = []
        for i in range(b):
            res.append(np.ctypeslib.as_array(outs[i], shape=(arrs[i].size,)).copy())
            self.L.mio_tts_b200_free(outs[i])
        return res

    def enable_batching(self, max_batch=64, max_wait_us=300, workers=2):
        """Micro-batcher for concurrent mio_tts_synthesize calls on this context and its later clones (max_batch <= 0: off)."""
        self._check(self.L.mio_tts_b200_context_enable_batching(self.ctx, max_batch, max_wait_us, workers, self._err, ERR_SIZE))

    def batching_stats(self):
        v = [C.c_uint64(0) for _ in range(4)]
        if not self.L.mio_tts_b200_batching_stats(self.ctx, *[C.byref(x) for x in v]):
            return None
        return dict(zip(("requests", "batches", "largest", "retried_single"), [int(x.value) for x in v]))

    # ---- stage hooks ----------------------------------------------------------------------------------
    def decode_spec(self, codes, embedding):
        codes = _i32(codes)
        emb = _f32(embedding) if embedding is not None else None
        out, nf, nb = C.POINTER(C.c_float)(), C.c_int32(0), C.c_int32(0)
        ok = self.L.mio_tts_b200_decode_spec(self.ctx, _ptr(codes, C.c_int32), codes.size,
                                             _ptr(emb, C.c_float) if emb is not None else None,
                                             emb.size if emb is not None else 0, C.byref(out), C.byref(nf),
                                             C.byref(nb), self._err, ERR_SIZE)
        self._check(ok)
        a = np.ctypeslib.as_array(out, shape=(nf.value, nb.value)).copy()
        self.L.mio_tts_b200_free(out)
        return a

    def debug_stage(self, codes, embedding, stage):
        codes = _i32(codes)
        emb = _f32(embedding) if embedding is not None else None
        out, r, c = C.POINTER(C.c_float)(), C.c_int32(0), C.c_int32(0)
        ok = self.L.mio_tts_b200_debug_stage(self.ctx, _ptr(codes, C.c_int32), codes.size,
                                             _ptr(emb, C.c_float) if emb is not None else None,
                                             emb.size if emb is not None else 0, stage, C.byref(out), C.byref(r),
                                             C.byref(c), self._err, ERR_SIZE)
        self._check(ok)
        a = np.ctypeslib.as_array(out, shape=(r.value, c.value)).copy()
        self.L.mio_tts_b200_free(out)
        return a

    def gather(self, codes):
        codes = _i32(codes)
        out = np.empty((codes.size, self.prenet_dim()), np.float32)
        self._check(self.L.mio_tts_b200_gather(self.ctx, _ptr(codes, C.c_int32) if codes.size else None, codes.size,
                                               _ptr(out, C.c_float), self._err, ERR_SIZE))
        return out

    def istft(self, spec, n_fft, hop):
        spec = _f32(spec)
        out, n = C.POINTER(C.c_float)(), C.c_size_t(0)
        self._check(self.L.mio_tts_b200_istft(self.ctx, _ptr(spec, C.c_float), spec.shape[0], n_fft, hop, C.byref(out),
                                              C.byref(n), self._err, ERR_SIZE))
        a = np.ctypeslib.as_array(out, shape=(max(1, n.value),)).copy()[: n.value]
        self.L.mio_tts_b200_free(out)
        return a

    def set_generic_attention(self, enable):
        self.L.mio_tts_b200_set_generic_attention(self.ctx, 1 if enable else 0)

    # ---- workspace ------------------------------------------------------------------------------------
    def reserve_workspace(self, n_codes):
        self._check(self.L.mio_tts_reserve_workspace(self.ctx, n_codes, self._err, ERR_SIZE))

    def estimate_workspace_bytes(self, n_codes):
        v = C.c_uint64(0)
        self._check(self.L.mio_tts_estimate_workspace_bytes(self.ctx, n_codes, C.byref(v), self._err, ERR_SIZE))
        return int(v.value)

    def workspace_bytes(self):
        return int(self.L.mio_tts_b200_workspace_bytes(self.ctx))

    def weight_bytes(self):
        return int(self.L.mio_tts_b200_weight_bytes(self.ctx))

    # ---- timing hooks (bench.py) ------------------------------------------------------------------------
    def prepare(self, codes_list, emb_list):
        B, codes, cp, cn, embs, ep, n_emb = self._batch_args(codes_list, emb_list)
        self._check(self.L.mio_tts_b200_prepare(self.ctx, cp, cn, ep, n_emb, B, self._err, ERR_SIZE))

    def run_resident(self, iters=1):
        ms = self.L.mio_tts_b200_run_resident(self.ctx, iters, self._err, ERR_SIZE)
        if ms < 0:
            raise MioTtsError(self._err.value.decode())
        return float(ms)

    def profile_resident(self):
        p = Profile()
        self._check(self.L.mio_tts_b200_profile_resident(self.ctx, C.byref(p), self._err, ERR_SIZE))
        return {k: getattr(p, k) for k, _ in Profile._fields_}

    def resident_samples(self):
        return int(self.L.mio_tts_b200_resident_samples(self.ctx))

    def last_h2d_bytes(self):
        return int(self.L.mio_tts_b200_last_h2d_bytes(self.ctx))

    def gather_bench(self, iters=20):
        ms = self.L.mio_tts_b200_gather_bench(self.ctx, iters, self._err, ERR_SIZE)
        if ms < 0:
            raise MioTtsError(self._err.value.decode())
        return float(ms)

    def istft_bench(self, n_utt, frames, n_fft, hop, iters, seed=7):
        cs = C.c_double(0)
        ms = self.L.mio_tts_b200_istft_bench(self.ctx, n_utt, frames, n_fft, hop, iters, seed, C.byref(cs), self._err,
                                             ERR_SIZE)
        if ms < 0:
            raise MioTtsError(self._err.value.decode())
        return float(ms), float(cs.value)

    # ---- reference audio -> speaker embedding (WavLM + global encoder) ------------------------------------
    def reference_to_embedding(self, wav_path, params=None):
        p = params or self.default_params()
        out, n = C.POINTER(C.c_float)(), C.c_size_t(0)
        self._check(self.L.mio_tts_reference_to_embedding(self.ctx, wav_path.encode(), p, C.byref(out), C.byref(n), self._err,
                                                          ERR_SIZE))
        a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
        self.L.mio_tts_embedding_free(out)
        return a

    def reference_pcm_to_embedding_batch(self, pcm_list, sample_rate, params=None):
        p = params or self.default_params()
        arrs = [_f32(x) for x in pcm_list]
        b = len(arrs)
        ptrs = (C.POINTER(C.c_float) * b)(*[_ptr(a, C.c_float) for a in arrs])
        ns = (C.c_size_t * b)(*[a.size for a in arrs])
        outs = (C.POINTER(C.c_float) * b)()
        nout = (C.c_size_t * b)()
        self._check(self.L.mio_tts_b200_reference_pcm_to_embedding_batch(self.ctx, ptrs, ns, sample_rate, b, p, outs, nout,
                                                                         self._err, ERR_SIZE))
        res = []
        for i in range(b):
            res.append(np.ctypeslib.as_array(outs[i], shape=(nout[i],)).copy())
            self.L.mio_tts_embedding_free(outs[i])
        return res

    def extract_ssl(self, pcm, sample_rate, max_seconds=20.0):
        x = _f32(pcm)
        out = C.POINTER(C.c_float)()
        nf, dim = C.c_int32(0), C.c_int32(0)
        self._check(self.L.mio_tts_b200_extract_ssl(self.ctx, _ptr(x, C.c_float), x.size, sample_rate, max_seconds, C.byref(out),
                                                    C.byref(nf), C.byref(dim), self._err, ERR_SIZE))
        a = np.ctypeslib.as_array(out, shape=(nf.value, dim.value)).copy()
        self.L.mio_tts_b200_free(out)
        return a

    def ssl_to_embedding(self, ssl):
        x = _f32(ssl)
        out, n = C.POINTER(C.c_float)(), C.c_size_t(0)
        self._check(self.L.mio_tts_b200_ssl_to_embedding(self.ctx, _ptr(x, C.c_float), x.shape[0], x.shape[1], C.byref(out),
                                                         C.byref(n), self._err, ERR_SIZE))
        a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
        self.L.mio_tts_embedding_free(out)
        return a

    def reference_bench(self, pcm, sample_rate, batch, iters):
        x = _f32(pcm)
        ms = self.L.mio_tts_b200_reference_bench(self.ctx, _ptr(x, C.c_float), x.size, sample_rate, batch, iters, self._err,
                                                 ERR_SIZE)
        if ms < 0:
            raise MioTtsError(self._err.value.decode())
        return float(ms)

    def reserve_reference_workspace(self, max_seconds):
        self._check(self.L.mio_tts_reserve_reference_workspace(self.ctx, max_seconds, self._err, ERR_SIZE))

    def estimate_reference_workspace_bytes(self, max_seconds):
        v = C.c_uint64(0)
        self._check(self.L.mio_tts_estimate_reference_workspace_bytes(self.ctx, max_seconds, C.byref(v), self._err, ERR_SIZE))
        return int(v.value)

    def launch_count(self):
        return int(self.L.mio_tts_b200_launch_count())


# ---- context-free helpers (same names as the C API) ---------------------------------------------------------
def _err_buf():
    return C.create_string_buffer(ERR_SIZE)


def codes_load(path):
    L, e = load_library(), _err_buf()
    out, n = C.POINTER(C.c_int32)(), C.c_size_t(0)
    if not L.mio_tts_codes_load(path.encode() if path is not None else None, C.byref(out), C.byref(n), e, ERR_SIZE):
        raise MioTtsError(e.value.decode())
    a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
    L.mio_tts_codes_free(out)
    return a


def codes_save(path, codes):
    L, e = load_library(), _err_buf()
    codes = _i32(codes)
    if not L.mio_tts_codes_save(path.encode() if path is not None else None,
                                _ptr(codes, C.c_int32) if codes.size else None, codes.size, e, ERR_SIZE):
        raise MioTtsError(e.value.decode())


def embedding_load_gguf(path):
    L, e = load_library(), _err_buf()
    out, n = C.POINTER(C.c_float)(), C.c_size_t(0)
    if not L.mio_tts_embedding_load_gguf(path.encode() if path is not None else None, C.byref(out), C.byref(n), e,
                                         ERR_SIZE):
        raise MioTtsError(e.value.decode())
    a = np.ctypeslib.as_array(out, shape=(n.value,)).copy()
    L.mio_tts_embedding_free(out)
    return a


def embedding_save_gguf(path, emb):
    L, e = load_library(), _err_buf()
    emb = _f32(emb)
    if not L.mio_tts_embedding_save_gguf(path.encode() if path is not None else None,
                                         _ptr(emb, C.c_float) if emb.size else None, emb.size, e, ERR_SIZE):
        raise MioTtsError(e.value.decode())


def audio_save_wav16(path, audio, sample_rate):
    L, e = load_library(), _err_buf()
    audio = _f32(audio)
    if not L.mio_tts_audio_save_wav16(path.encode() if path is not None else None,
                                      _ptr(audio, C.c_float) if audio.size else None, audio.size, sample_rate, e,
                                      ERR_SIZE):
        raise MioTtsError(e.value.decode())


def pcm16_static(device, d_audio, offsets, lengths, d_pcm, stream=None):
    """mio_tts_b200_pcm16_static: WAV16 conversion over DEVICE buffers (raw addresses) with the library's static scratch;
    allocates nothing.  offsets / lengths are host sequences in samples."""
    L, e = load_library(), _err_buf()
    off = np.ascontiguousarray(offsets, dtype=np.int64)
    ln = np.ascontiguousarray(lengths, dtype=np.int64)
    if not L.mio_tts_b200_pcm16_static(int(device), C.c_void_p(d_audio), _ptr(off, C.c_int64), _ptr(ln, C.c_int64), off.size,
                                       C.c_void_p(d_pcm), C.c_void_p(stream), e, ERR_SIZE):
        raise MioTtsError(e.value.decode())
