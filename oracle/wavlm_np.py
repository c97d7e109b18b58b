"""CPU restatement (numpy) of the reference-audio -> speaker-embedding path -- TEST INFRASTRUCTURE ONLY.

  reference audio --(decode, peak-normalise, linear resample)--> 16 kHz waveform
                  --WavLM SSL extractor (conv stack, pos-conv, 2 gated-relative-bias layers)--> ssl [frames, 768]
                  --global encoder (ConvNeXt backbone + attentive statistics pooling)--> embedding [128]

Restates `wavlm_extractor::extract_ssl_features` (/root/reference/src/wavlm-extractor.cpp:543-1081) and
`miocodec_decoder::encode_global_embedding_from_ssl` (src/miocodec-decoder.cpp:824-1066), i.e. the body of
`mio_tts_reference_to_embedding` (src/mio-tts-lib.cpp:1048-1125).  SURVEY.md section 8f rows N1/N2.

Only tests/ (and tools/gen_golden.py) import this.  Parity pinning: like oracle/miocodec_np.py the reference has no
golden vectors for this path, so the restatement is pinned against the compiled reference (oracle/_ref:
ref_wavlm_extract / ref_decoder_encode_global / mio_tts_reference_to_embedding) live and through the committed
outputs in tests/golden/refemb_*.npz.

Conventions the reference fixes (and a from-scratch kernel gets wrong first):
* every ggml_conv_1d rounds BOTH operands to fp16 (im2col F16 x F16 kernel, fp32 accumulate, ggml.c:4409-4426);
  linears are pure fp32.
* ggml_gelu on CPU is a table over fp16 inputs: gelu(x) = fp16(gelu_tanh(fp16(x))) for |x| < 10
  (ggml-cpu/vec.h ggml_vec_gelu_f32 with GGML_GELU_FP16).
* conv0 GroupNorm has one group per channel: statistics over TIME (wavlm-extractor.cpp:669-679).
* the relative-position gate is computed for columns ordered (head fastest, time slowest) and then REINTERPRETED
  as [time, head] with time fastest (wavlm-extractor.cpp:769-790: reshape_4d(g, 1, seq, heads, 1) of a vector laid
  out head-major) -- so gate[q, h] = g_flat[q + seq*h] where g_flat[j] belongs to (head j % H, time j // H).
  The restatement (and the CUDA path) follow the reference, not torchaudio.
"""
import struct

import numpy as np

F32 = np.float32


def fp16_round(x):
    return np.asarray(x, dtype=F32).astype(np.float16).astype(F32)


# --------------------------------------------------------------------------------------------
# audio front end
# --------------------------------------------------------------------------------------------
def read_wav_mono(path):
    """PCM WAV -> float32 mono in [-1, 1) and its sample rate (what miniaudio's ma_decoder hands the reference,
    wavlm-extractor.cpp:153-203, for files already at the requested rate).  Stereo is averaged."""
    with open(path, "rb") as f:
        data = f.read()
    if data[:4] != b"RIFF" or data[8:12] != b"WAVE":
        raise ValueError("not a RIFF/WAVE file")
    pos, fmt, pcm = 12, None, None
    while pos + 8 <= len(data):
        cid, size = data[pos:pos + 4], struct.unpack("<I", data[pos + 4:pos + 8])[0]
        body = data[pos + 8:pos + 8 + size]
        if cid == b"fmt ":
            fmt = struct.unpack("<HHIIHH", body[:16])
        elif cid == b"data":
            pcm = body
        pos += 8 + size + (size & 1)
    if fmt is None or pcm is None:
        raise ValueError("WAV without fmt/data chunk")
    tag, ch, sr, _, _, bits = fmt
    if tag == 1 and bits == 16:
        x = np.frombuffer(pcm[: len(pcm) // 2 * 2], dtype="<i2").astype(F32) / F32(32768.0)
    elif tag == 1 and bits == 32:
        x = np.frombuffer(pcm[: len(pcm) // 4 * 4], dtype="<i4").astype(F32) / F32(2147483648.0)
    elif tag == 3 and bits == 32:
        x = np.frombuffer(pcm[: len(pcm) // 4 * 4], dtype="<f4").astype(F32)
    else:
        raise ValueError(f"unsupported WAV format tag={tag} bits={bits}")
    if ch > 1:
        x = x[: x.size // ch * ch].reshape(-1, ch).mean(axis=1, dtype=F32)
    return x.astype(F32), int(sr)


def normalize_audio(x):
    """wavlm-extractor.cpp:205-216: x / (max|x| + 1e-8)."""
    m = F32(np.abs(x).max(initial=0.0)) + F32(1e-8)
    return (x / m).astype(F32)


def resample_linear(x, sr_in, sr_out):
    """wavlm-extractor.cpp:218-240 (double-precision positions, float32 result)."""
    if sr_in == sr_out or x.size == 0:
        return x.astype(F32)
    ratio = float(sr_out) / float(sr_in)
    n_out = max(1, int(np.floor(x.size * ratio + 0.5)))      # llround
    pos = np.arange(n_out, dtype=np.float64) / ratio
    i0 = np.floor(pos).astype(np.int64)
    i1 = np.minimum(x.size - 1, i0 + 1)
    t = pos - i0
    xd = x.astype(np.float64)
    return ((1.0 - t) * xd[np.maximum(0, i0)] + t * xd[i1]).astype(F32)


def prepare_waveform(pcm, sr_in, sr_wavlm=16000, max_seconds=20.0):
    """extract_ssl_features up to the graph input (wavlm-extractor.cpp:562-600)."""
    if max_seconds > 0:
        pcm = pcm[: max(1, int(np.floor(sr_in * float(max_seconds) + 0.5)))]
    return resample_linear(normalize_audio(pcm), sr_in, sr_wavlm)


# --------------------------------------------------------------------------------------------
# shared op restatements
# --------------------------------------------------------------------------------------------
def gelu_ggml(x):
    """ggml_gelu on the CPU backend: tanh form evaluated through the fp16 lookup table."""
    x = np.asarray(x, dtype=F32)
    xh = x.astype(np.float16).astype(F32)
    y = F32(0.5) * xh * (F32(1.0) + np.tanh(F32(0.79788456080286535587989211986876) * xh *
                                            (F32(1.0) + F32(0.044715) * xh * xh)))
    y = y.astype(np.float16).astype(F32)
    y = np.where(x <= -10.0, F32(0.0), y)
    y = np.where(x >= 10.0, x, y)
    return y.astype(F32)


def layer_norm(x, w, b, eps):
    """ggml_norm over the last axis (ggml-cpu/ops.cpp:3642-3687) + affine (layer_norm(), miocodec-decoder.cpp:136-150)."""
    mean = x.mean(axis=-1, keepdims=True, dtype=F32)
    xc = x - mean
    var = (xc * xc).mean(axis=-1, keepdims=True, dtype=F32)
    y = xc * (F32(1.0) / np.sqrt(var + F32(eps)))
    return (y * w + b).astype(F32)


def conv1d_f16(x, w, stride, pad):
    """ggml_conv_1d: x [T, Cin], w torch layout [Cout, Cin, K]; fp16 operands, fp32 accumulate -> [T_out, Cout]."""
    T, cin = x.shape
    cout, cin_w, K = w.shape
    assert cin_w == cin
    xp = np.zeros((T + 2 * pad, cin), dtype=F32)
    xp[pad:pad + T] = fp16_round(x)
    t_out = (T + 2 * pad - K) // stride + 1
    wk = fp16_round(w)
    y = np.zeros((t_out, cout), dtype=F32)
    for j in range(K):
        y += xp[j:j + (t_out - 1) * stride + 1:stride] @ wk[:, :, j].T
    return y


def dwconv1d_f16(x, w, pad):
    """ggml_conv_1d_dw_ph: depthwise, w [C, 1, K]; fp16 operands -> [T, C]."""
    T, C = x.shape
    K = w.shape[2]
    xp = np.zeros((T + 2 * pad, C), dtype=F32)
    xp[pad:pad + T] = fp16_round(x)
    wk = fp16_round(w[:, 0, :])
    y = np.zeros((T + 2 * pad - K + 1, C), dtype=F32)
    for j in range(K):
        y += xp[j:j + y.shape[0]] * wk[:, j]
    return y


# --------------------------------------------------------------------------------------------
# GGUF loading
# --------------------------------------------------------------------------------------------
def _read_gguf(path):
    import gguf

    r = gguf.GGUFReader(path)
    t = {}
    for x in r.tensors:
        shape = tuple(int(v) for v in reversed(x.shape.tolist()))
        t[x.name] = np.array(x.data).reshape(shape)

    def kv(name, default):
        f = r.fields.get(name)
        if f is None:
            return default
        return type(default)(f.parts[f.data[0]][0])

    return t, kv


class WavLM:
    """wavlm_extractor::load (wavlm-extractor.cpp:454-540); struct defaults wavlm-extractor.h:13-25."""

    def __init__(self, path):
        self.t, kv = _read_gguf(path)
        self.hp = dict(sample_rate=16000, n_layers=2, n_heads=12, head_dim=64, embed_dim=768, num_buckets=320,
                       max_distance=800)
        for k in list(self.hp):
            self.hp[k] = kv("wavlm." + k, self.hp[k])
        self.hp["norm_eps"] = kv("wavlm.layer_norm_eps", 1e-5)
        ks, st = [10, 3, 3, 3, 3, 2, 2], [5, 2, 2, 2, 2, 2, 2]
        self.conv_kernel = [kv(f"wavlm.feat.conv{i}.kernel", ks[i]) for i in range(7)]
        self.conv_stride = [kv(f"wavlm.feat.conv{i}.stride", st[i]) for i in range(7)]


def relative_position_bucket(rel, num_buckets, max_distance):
    """wavlm-extractor.cpp:257-279 (vectorised; float32 logs like the reference)."""
    rel = np.asarray(rel, dtype=np.int64)
    half = num_buckets // 2
    out = np.where(rel > 0, half, 0).astype(np.int64)
    n = np.abs(rel)
    max_exact = half // 2
    nf = np.maximum(n, 1).astype(F32)
    denom = F32(np.log(F32(max_distance) / F32(max_exact)))
    val = max_exact + np.floor(np.log(nf / F32(max_exact)).astype(F32) / denom * F32(half - max_exact)).astype(np.int64)
    val = np.minimum(val, half - 1)
    return out + np.where(n < max_exact, n, val)


def wavlm_features(m, wav16k, stages=None):
    """The ggml graph of extract_ssl_features (wavlm-extractor.cpp:664-865): waveform -> ssl [frames, embed_dim]."""
    t, hp = m.t, m.hp
    eps = F32(hp["norm_eps"])
    x = np.asarray(wav16k, dtype=F32).reshape(-1, 1)
    for i in range(7):                                             # :664-681
        x = conv1d_f16(x, t[f"wavlm.feat.conv{i}.weight"], m.conv_stride[i], 0)
        if i == 0:   # GroupNorm with one group per channel: statistics over time (double accumulators)
            xd = x.astype(np.float64)
            mean = xd.mean(axis=0, keepdims=True)
            var = ((xd - mean) ** 2).mean(axis=0, keepdims=True)
            x = ((xd - mean) / np.sqrt(var + float(eps))).astype(F32)
            x = x * t["wavlm.feat.conv0.norm.weight"] + t["wavlm.feat.conv0.norm.bias"]
        x = gelu_ggml(x)
        if stages is not None:
            stages[f"conv{i}"] = x
    x = layer_norm(x, t["wavlm.proj.norm.weight"], t["wavlm.proj.norm.bias"], eps)          # :686
    x = (x @ t["wavlm.proj.weight"].T + t["wavlm.proj.bias"]).astype(F32)                    # :687
    if stages is not None:
        stages["proj"] = x
    seq, dim = x.shape
    # grouped positional conv, kernel 128, padding 64, 16 groups, output cropped to seq (:689-735)
    groups = 16
    pw = t["wavlm.pos_conv.weight"]                      # [dim, dim/groups, K]
    ipg, opg = dim // groups, pw.shape[0] // groups
    pos = np.zeros((seq, dim), dtype=F32)
    for g in range(groups):
        yg = conv1d_f16(x[:, g * ipg:(g + 1) * ipg], pw[g * opg:(g + 1) * opg], 1, 64)
        pos[:, g * opg:(g + 1) * opg] = yg[:seq]
    pos = gelu_ggml(pos + t["wavlm.pos_conv.bias"])
    x = layer_norm(x + pos, t["wavlm.transformer.norm.weight"], t["wavlm.transformer.norm.bias"], eps)   # :737-739
    if stages is not None:
        stages["pre"] = x
    H, hd = hp["n_heads"], hp["head_dim"]
    rel = t["wavlm.layer.0.attn.rel_embed.weight"]      # [num_buckets, heads]
    q_idx = np.arange(seq)
    bucket = relative_position_bucket(q_idx[None, :] - q_idx[:, None], hp["num_buckets"], hp["max_distance"])  # [q, k]
    raw_bias = rel[bucket]                               # [q, k, heads]  (:893-912)
    scale = F32(1.0 / np.sqrt(hd))
    ssl_sum = np.zeros_like(x)
    for li in range(hp["n_layers"]):                     # wavlm_layer_eval :762-851
        p = f"wavlm.layer.{li}."
        # gate: linear over each (time, head) slice, column order head-fastest ...
        qh = x.reshape(seq * H, hd)                      # row j = time j // H, head j % H
        g = (qh @ t[p + "attn.gru.weight"].T + t[p + "attn.gru.bias"]).astype(F32)          # [seq*H, 8]
        g0 = F32(1.0) / (F32(1.0) + np.exp(-g[:, :4].sum(axis=1, dtype=F32)))
        g1 = F32(1.0) / (F32(1.0) + np.exp(-g[:, 4:].sum(axis=1, dtype=F32)))
        # ... reinterpreted as [heads, seq] with time fastest (reshape_4d(g, 1, seq, heads, 1))
        g0 = g0.reshape(H, seq)
        g1 = g1.reshape(H, seq)
        gc = t[p + "attn.gru_const"].reshape(H, 1)
        gate = (g0 * (g1 * gc - F32(1.0)) + F32(2.0)).astype(F32)           # [heads, q]
        qkv = (x @ t[p + "attn.in_proj.weight"].T + t[p + "attn.in_proj.bias"]).astype(F32)
        q, k, v = (qkv[:, i * dim:(i + 1) * dim].reshape(seq, H, hd).transpose(1, 0, 2) for i in range(3))
        s = np.einsum("hqd,hkd->hqk", q, k).astype(F32) * scale + raw_bias.transpose(2, 0, 1) * gate[:, :, None]
        s = s - s.max(axis=-1, keepdims=True)
        pr = np.exp(s).astype(F32)
        pr = pr / pr.sum(axis=-1, keepdims=True, dtype=F32)
        att = np.einsum("hqk,hkd->qhd", pr, v).reshape(seq, dim).astype(F32)
        h = x + (att @ t[p + "attn.out_proj.weight"].T + t[p + "attn.out_proj.bias"]).astype(F32)
        n1 = layer_norm(h, t[p + "norm1.weight"], t[p + "norm1.bias"], eps)
        ff = gelu_ggml((n1 @ t[p + "ffn.w1.weight"].T + t[p + "ffn.w1.bias"]).astype(F32))
        ff = (ff @ t[p + "ffn.w2.weight"].T + t[p + "ffn.w2.bias"]).astype(F32)
        x = layer_norm(h + ff, t[p + "norm2.weight"], t[p + "norm2.bias"], eps)
        if stages is not None:
            stages[f"layer{li}"] = x
        ssl_sum = ssl_sum + x
    return (ssl_sum * F32(1.0 / max(1, hp["n_layers"]))).astype(F32)     # :853-865


def extract_ssl_from_pcm(m, pcm, sr_in, max_seconds=20.0):
    return wavlm_features(m, prepare_waveform(pcm, sr_in, m.hp["sample_rate"], max_seconds))


# --------------------------------------------------------------------------------------------
# global encoder (tensors live in the MioCodec GGUF; pass oracle.miocodec_np.Model or anything with .t)
# --------------------------------------------------------------------------------------------
def global_encoder_layers(t):
    n = 0
    while f"global_encoder.backbone.blk.{n}.dwconv.weight" in t:
        n += 1
    return n


def global_embedding(model, ssl, stages=None):
    """encode_global_embedding_from_ssl (miocodec-decoder.cpp:862-945): ssl [frames, C] -> embedding [out]."""
    t = model.t
    g = "global_encoder."
    ew = t[g + "backbone.embed.weight"]
    x = conv1d_f16(np.asarray(ssl, dtype=F32), ew, 1, ew.shape[2] // 2) + t[g + "backbone.embed.bias"]     # :871-874
    x = layer_norm(x.astype(F32), t[g + "backbone.norm.weight"], t[g + "backbone.norm.bias"], 1e-6)
    for i in range(global_encoder_layers(t)):                                                              # :883-905
        b = g + f"backbone.blk.{i}."
        dw = t[b + "dwconv.weight"]
        y = dwconv1d_f16(x, dw, dw.shape[2] // 2) + t[b + "dwconv.bias"]
        y = layer_norm(y.astype(F32), t[b + "norm.weight"], t[b + "norm.bias"], 1e-6)
        y = gelu_ggml((y @ t[b + "pw1.weight"].T + t[b + "pw1.bias"]).astype(F32))
        y = (y @ t[b + "pw2.weight"].T + t[b + "pw2.bias"]).astype(F32) * t[b + "gamma"]
        x = (x + y).astype(F32)
        if stages is not None:
            stages[f"blk{i}"] = x
    x = layer_norm(x, t[g + "backbone.final_norm.weight"], t[g + "backbone.final_norm.bias"], 1e-6)        # :908-910
    if stages is not None:
        stages["backbone"] = x
    # attentive statistics pooling (:913-937): k=1 convs (fp16 operands), softmax over TIME per channel
    a = np.tanh(conv1d_f16(x, t[g + "pool.attn0.weight"], 1, 0) + t[g + "pool.attn0.bias"]).astype(F32)
    a = (conv1d_f16(a, t[g + "pool.attn2.weight"], 1, 0) + t[g + "pool.attn2.bias"]).astype(F32)
    a = a - a.max(axis=0, keepdims=True)
    alpha = np.exp(a).astype(F32)
    alpha = alpha / alpha.sum(axis=0, keepdims=True, dtype=F32)
    mean = (alpha * x).sum(axis=0, dtype=F32)
    m2 = (alpha * x * x).sum(axis=0, dtype=F32)
    std = np.sqrt(np.clip(m2 - mean * mean, F32(1e-4), F32(1e4))).astype(F32)
    stat = np.concatenate([mean, std]).astype(F32)
    out = (t[g + "pool.proj.weight"] @ stat + t[g + "pool.proj.bias"]).astype(F32)
    out = layer_norm(out[None, :], t[g + "pool.norm.weight"], t[g + "pool.norm.bias"], 1e-5)[0]
    if not np.isfinite(out).all():                                                                           # :1048-1061
        out = np.zeros_like(out)
    return out.astype(F32)


def reference_to_embedding(wavlm, codec_model, wav_path, max_seconds=20.0):
    """mio_tts_reference_to_embedding (mio-tts-lib.cpp:1048-1125) for a WAV at the codec sample rate."""
    pcm, sr = read_wav_mono(wav_path)
    return global_embedding(codec_model, extract_ssl_from_pcm(wavlm, pcm, sr, max_seconds))
