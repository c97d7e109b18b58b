#!/usr/bin/env python3
"""Write a synthetic-weight MioCodec decoder GGUF in the reference's weight layout.

The real `miocodec.gguf` is downloaded from HuggingFace by the reference
(/root/reference/models_download.sh:45-58) and is not available offline, so every parity and
benchmark run uses seeded random weights written with exactly the tensor names, shapes and KV
keys that `miocodec_decoder::load` reads (/root/reference/src/miocodec-decoder.cpp:458-711) and that
`scripts/convert_miocodec_to_gguf.py:351-588` writes.  A real GGUF drops in unchanged.

Presets
  default    24 kHz struct defaults (/root/reference/src/miocodec-decoder.h:12-49): 91.1 M params, 364 MB
  tiny       same topology at 128-wide dims, 2+2 layers, n_fft 120 -- used for the committed golden vectors
  upsampler  wave_upsampler_layers=2 (factors [3,3], kernels [7,7], 512->256->128), n_fft 392, hop 98,
             1764 samples/token -- exercises the optional branch (miocodec-decoder.cpp:1410-1455)
  upsampler_tiny   the same branch on the tiny trunk
"""
import argparse
import hashlib
import sys

import numpy as np

PRESETS = {
    "default": dict(sample_rate=24000, n_fft=1920, hop_length=480, samples_per_token=960,
                    prenet_layers=6, prenet_dim=768, prenet_heads=12, prenet_ff=2048, prenet_window=65,
                    decoder_layers=8, decoder_dim=512, decoder_heads=8, decoder_ff=1536, decoder_window=65,
                    decoder_adanorm_dim=128, resnet_blocks=2, resnet_groups=32,
                    up_factors=[], up_kernels=[], up_channels=[]),
    "tiny": dict(sample_rate=24000, n_fft=120, hop_length=30, samples_per_token=60,
                 prenet_layers=2, prenet_dim=128, prenet_heads=2, prenet_ff=256, prenet_window=9,
                 decoder_layers=2, decoder_dim=128, decoder_heads=2, decoder_ff=256, decoder_window=9,
                 decoder_adanorm_dim=128, resnet_blocks=1, resnet_groups=32,
                 up_factors=[], up_kernels=[], up_channels=[]),
    "upsampler": dict(sample_rate=44100, n_fft=392, hop_length=98, samples_per_token=1764,
                      prenet_layers=2, prenet_dim=768, prenet_heads=12, prenet_ff=2048, prenet_window=65,
                      decoder_layers=2, decoder_dim=512, decoder_heads=8, decoder_ff=1536, decoder_window=65,
                      decoder_adanorm_dim=128, resnet_blocks=1, resnet_groups=32,
                      up_factors=[3, 3], up_kernels=[7, 7], up_channels=[256, 128], up_out=128),
    "upsampler_tiny": dict(sample_rate=44100, n_fft=56, hop_length=14, samples_per_token=252,
                           prenet_layers=1, prenet_dim=128, prenet_heads=2, prenet_ff=256, prenet_window=9,
                           decoder_layers=1, decoder_dim=128, decoder_heads=2, decoder_ff=256, decoder_window=9,
                           decoder_adanorm_dim=128, resnet_blocks=1, resnet_groups=32,
                           up_factors=[3, 3], up_kernels=[7, 5], up_channels=[64, 32], up_out=48),
    # frames != 2*T: the lerp branch of ggml_interpolate (/root/reference/src/miocodec-decoder.cpp:1097-1103,1373),
    # three frames per token (upward) and 1.5 frames per token (downward, odd T leaves a remainder)
    "tiny_lerp3": dict(sample_rate=24000, n_fft=120, hop_length=30, samples_per_token=90,
                       prenet_layers=1, prenet_dim=128, prenet_heads=2, prenet_ff=256, prenet_window=9,
                       decoder_layers=1, decoder_dim=128, decoder_heads=2, decoder_ff=256, decoder_window=9,
                       decoder_adanorm_dim=128, resnet_blocks=1, resnet_groups=32,
                       up_factors=[], up_kernels=[], up_channels=[]),
    "tiny_lerp15": dict(sample_rate=24000, n_fft=120, hop_length=30, samples_per_token=45,
                        prenet_layers=1, prenet_dim=128, prenet_heads=2, prenet_ff=256, prenet_window=9,
                        decoder_layers=1, decoder_dim=128, decoder_heads=2, decoder_ff=256, decoder_window=9,
                        decoder_adanorm_dim=128, resnet_blocks=1, resnet_groups=32,
                        up_factors=[], up_kernels=[], up_channels=[]),
    # dynamic_global = 0: plain LayerNorm weights in the decoder blocks and tail instead of the AdaLN condition
    # linears (/root/reference/src/miocodec-decoder.cpp:662-674, 701-711, 1190-1291, 1390-1400)
    "tiny_static": dict(sample_rate=24000, n_fft=120, hop_length=30, samples_per_token=60, dynamic_global=0,
                        prenet_layers=2, prenet_dim=128, prenet_heads=2, prenet_ff=256, prenet_window=9,
                        decoder_layers=2, decoder_dim=128, decoder_heads=2, decoder_ff=256, decoder_window=9,
                        decoder_adanorm_dim=128, resnet_blocks=1, resnet_groups=32,
                        up_factors=[], up_kernels=[], up_channels=[]),
}
VOCAB = 12800  # codes 0..12799, /root/reference/src/mio-tts-lib.cpp:30-31

# Global speaker-embedding encoder (optional tensors, /root/reference/src/miocodec-decoder.cpp:713-744; KVs :489-493;
# graph :824-945).  "default" = the struct defaults (miocodec-decoder.h:44-48); "tiny" pairs with the tiny WavLM.
GE_PRESETS = {
    "default": dict(input_channels=768, output_channels=128, dim=384, intermediate_dim=1152, num_layers=4,
                    embed_kernel=7, dw_kernel=7, attn_channels=128),
    "tiny": dict(input_channels=128, output_channels=128, dim=64, intermediate_dim=128, num_layers=2,
                 embed_kernel=7, dw_kernel=7, attn_channels=64),
}

# WavLM SSL extractor (/root/reference/src/wavlm-extractor.cpp:454-540 load, :664-865 graph;
# scripts/convert_wavlm_base_plus_to_gguf.py:118-186 writer).  "base" = WavLM Base+ truncated to 2 layers.
WAVLM_PRESETS = {
    "base": dict(sample_rate=16000, n_layers=2, n_heads=12, head_dim=64, embed_dim=768, num_buckets=320,
                 max_distance=800, conv_channels=512, ffn=3072, pos_kernel=128, pos_groups=16),
    "tiny": dict(sample_rate=16000, n_layers=2, n_heads=2, head_dim=64, embed_dim=128, num_buckets=320,
                 max_distance=800, conv_channels=64, ffn=256, pos_kernel=128, pos_groups=16),
}
WAVLM_CONV_K = [10, 3, 3, 3, 3, 2, 2]
WAVLM_CONV_S = [5, 2, 2, 2, 2, 2, 2]


def synth_tensors(cfg, seed=0):
    """Yield (name, ndarray) in a fixed order; numpy shapes are the reverse of ggml `ne`."""
    rng = np.random.default_rng(seed)
    f32 = np.float32

    def lin(o, i, gain=1.0):
        return (rng.standard_normal((o, i)) * (gain / np.sqrt(i))).astype(f32)

    def norm_w(c):
        return (1.0 + 0.05 * rng.standard_normal(c)).astype(f32)

    def small(c, s=0.05):
        return (s * rng.standard_normal(c)).astype(f32)

    pd, pf = cfg["prenet_dim"], cfg["prenet_ff"]
    dd, df, ad = cfg["decoder_dim"], cfg["decoder_ff"], cfg["decoder_adanorm_dim"]
    yield "token_embd", rng.standard_normal((VOCAB, pd)).astype(f32)
    for i in range(cfg["prenet_layers"]):
        p = f"wave_prenet.blk.{i}."
        yield p + "attn_norm.weight", norm_w(pd)
        yield p + "attn_norm.bias", small(pd)
        for n in ("attn_q", "attn_k", "attn_v", "attn_output"):
            yield p + n + ".weight", lin(pd, pd)
        yield p + "ffn_norm.weight", norm_w(pd)
        yield p + "ffn_norm.bias", small(pd)
        yield p + "ffn_gate.weight", lin(pf, pd)
        yield p + "ffn_up.weight", lin(pf, pd)
        yield p + "ffn_down.weight", lin(pd, pf)
    yield "wave_prenet.norm.weight", norm_w(pd)
    yield "wave_prenet.norm.bias", small(pd)
    yield "wave_prenet.output.weight", lin(dd, pd)
    yield "wave_prenet.output.bias", small(dd)
    # torch ConvTranspose1d weight [Cin, Cout, K]
    yield "wave_upsample.weight", (rng.standard_normal((dd, dd, 2)) / np.sqrt(dd)).astype(f32)
    yield "wave_upsample.bias", small(dd)

    def resblk(prefix, c):
        for j in (1, 2):
            yield f"{prefix}.norm{j}.weight", norm_w(c)
            yield f"{prefix}.norm{j}.bias", small(c)
            # torch Conv1d weight [Cout, Cin, K]
            yield f"{prefix}.conv{j}.weight", (rng.standard_normal((c, c, 3)) / np.sqrt(3 * c)).astype(f32)
            yield f"{prefix}.conv{j}.bias", small(c)

    for j in range(cfg["resnet_blocks"]):
        yield from resblk(f"wave_prior.{j}", dd)
    for j in range(cfg["resnet_blocks"]):
        yield from resblk(f"wave_post.{j}", dd)
    for i in range(cfg["decoder_layers"]):
        p = f"wave_decoder.blk.{i}."
        if cfg.get("dynamic_global", 1):
            for n in ("attn_cond", "ffn_cond"):
                yield p + n + ".weight", lin(3 * dd, ad, 0.5)
                b = small(3 * dd)
                b[2 * dd:] += 0.5  # gate ~ 0.5 so every block contributes
                yield p + n + ".bias", b
        else:
            for n in ("attn_norm", "ffn_norm"):
                yield p + n + ".weight", norm_w(dd)
                yield p + n + ".bias", small(dd)
        for n in ("attn_q", "attn_k", "attn_v", "attn_output"):
            yield p + n + ".weight", lin(dd, dd)
        yield p + "ffn_gate.weight", lin(df, dd)
        yield p + "ffn_up.weight", lin(df, dd)
        yield p + "ffn_down.weight", lin(dd, df)
    if cfg.get("dynamic_global", 1):
        yield "wave_decoder.norm_cond.weight", lin(2 * dd, ad, 0.5)
        yield "wave_decoder.norm_cond.bias", small(2 * dd)
    else:
        yield "wave_decoder.norm.weight", norm_w(dd)
        yield "wave_decoder.norm.bias", small(dd)
    c_in = dd
    for i, (f, k, c_out) in enumerate(zip(cfg["up_factors"], cfg["up_kernels"], cfg["up_channels"])):
        yield f"wave_upsampler.up.{i}.weight", (rng.standard_normal((c_in, c_out, k)) / np.sqrt(c_in * k / f)).astype(f32)
        yield f"wave_upsampler.up.{i}.bias", small(c_out)
        yield f"wave_upsampler.snake.{i}.alpha", small(c_out, 0.2)
        yield f"wave_upsampler.snake.{i}.beta", small(c_out, 0.2)
        yield from resblk(f"wave_upsampler.resblk.{i}", c_out)
        c_in = c_out
    if cfg["up_factors"]:
        yield "wave_upsampler.out_proj.weight", lin(cfg["up_out"], c_in)
        yield "wave_upsampler.out_proj.bias", small(cfg["up_out"])
        yield "wave_upsampler.out_snake.alpha", small(cfg["up_out"], 0.2)
        yield "wave_upsampler.out_snake.beta", small(cfg["up_out"], 0.2)
        c_in = cfg["up_out"]
    n_bins = cfg["n_fft"] + 2
    yield "istft_head.out.weight", lin(n_bins, c_in)
    hb = small(n_bins, 0.1)
    hb[: n_bins // 2] -= 2.0  # log-magnitude around e^-2
    yield "istft_head.out.bias", hb


def synth_global_encoder_tensors(ge, seed=0):
    """Global-encoder tensors in torch shapes (Conv1d [Cout, Cin/groups, K], Linear [out, in])."""
    rng = np.random.default_rng(seed + 7001)
    f32 = np.float32
    cin, dim, ff, out, ac = ge["input_channels"], ge["dim"], ge["intermediate_dim"], ge["output_channels"], ge["attn_channels"]

    def conv(o, i, k, gain=1.0):
        return (rng.standard_normal((o, i, k)) * (gain / np.sqrt(i * k))).astype(f32)

    def lin(o, i, gain=1.0):
        return (rng.standard_normal((o, i)) * (gain / np.sqrt(i))).astype(f32)

    def norm_w(c):
        return (1.0 + 0.05 * rng.standard_normal(c)).astype(f32)

    def small(c, sd=0.05):
        return (sd * rng.standard_normal(c)).astype(f32)

    g = "global_encoder."
    yield g + "backbone.embed.weight", conv(dim, cin, ge["embed_kernel"])
    yield g + "backbone.embed.bias", small(dim)
    yield g + "backbone.norm.weight", norm_w(dim)
    yield g + "backbone.norm.bias", small(dim)
    for i in range(ge["num_layers"]):
        b = g + f"backbone.blk.{i}."
        yield b + "dwconv.weight", conv(dim, 1, ge["dw_kernel"])
        yield b + "dwconv.bias", small(dim)
        yield b + "norm.weight", norm_w(dim)
        yield b + "norm.bias", small(dim)
        yield b + "pw1.weight", lin(ff, dim)
        yield b + "pw1.bias", small(ff)
        yield b + "pw2.weight", lin(dim, ff)
        yield b + "pw2.bias", small(dim)
        yield b + "gamma", (0.5 + 0.1 * rng.standard_normal(dim)).astype(f32)
    yield g + "backbone.final_norm.weight", norm_w(dim)
    yield g + "backbone.final_norm.bias", small(dim)
    yield g + "pool.attn0.weight", conv(ac, dim, 1)
    yield g + "pool.attn0.bias", small(ac)
    yield g + "pool.attn2.weight", conv(dim, ac, 1, 2.0)
    yield g + "pool.attn2.bias", small(dim)
    yield g + "pool.proj.weight", lin(out, 2 * dim)
    yield g + "pool.proj.bias", small(out)
    yield g + "pool.norm.weight", norm_w(out)
    yield g + "pool.norm.bias", small(out)


def synth_wavlm_tensors(cfg, seed=0):
    rng = np.random.default_rng(seed + 9001)
    f32 = np.float32
    cc, ed, hd, nh, ff = cfg["conv_channels"], cfg["embed_dim"], cfg["head_dim"], cfg["n_heads"], cfg["ffn"]

    def lin(o, i, gain=1.0):
        return (rng.standard_normal((o, i)) * (gain / np.sqrt(i))).astype(f32)

    def norm_w(c):
        return (1.0 + 0.05 * rng.standard_normal(c)).astype(f32)

    def small(c, sd=0.05):
        return (sd * rng.standard_normal(c)).astype(f32)

    yield "wavlm.feat.conv0.norm.weight", norm_w(cc)
    yield "wavlm.feat.conv0.norm.bias", small(cc, 0.2)
    yield "wavlm.feat.conv0.weight", (rng.standard_normal((cc, 1, WAVLM_CONV_K[0])) / np.sqrt(WAVLM_CONV_K[0])).astype(f32)
    for i in range(1, 7):
        k = WAVLM_CONV_K[i]
        # gain > 1: GELU halves the variance layer after layer
        yield f"wavlm.feat.conv{i}.weight", (rng.standard_normal((cc, cc, k)) * (1.6 / np.sqrt(cc * k))).astype(f32)
    yield "wavlm.proj.norm.weight", norm_w(cc)
    yield "wavlm.proj.norm.bias", small(cc)
    yield "wavlm.proj.weight", lin(ed, cc)
    yield "wavlm.proj.bias", small(ed)
    yield "wavlm.pos_conv.weight", (rng.standard_normal((ed, ed // cfg["pos_groups"], cfg["pos_kernel"]))
                                    / np.sqrt(ed // cfg["pos_groups"] * cfg["pos_kernel"])).astype(f32)
    yield "wavlm.pos_conv.bias", small(ed)
    yield "wavlm.transformer.norm.weight", norm_w(ed)
    yield "wavlm.transformer.norm.bias", small(ed)
    yield "wavlm.layer.0.attn.rel_embed.weight", (0.5 * rng.standard_normal((cfg["num_buckets"], nh))).astype(f32)
    for i in range(cfg["n_layers"]):
        p = f"wavlm.layer.{i}."
        yield p + "attn.in_proj.weight", lin(3 * ed, ed)
        yield p + "attn.in_proj.bias", small(3 * ed)
        yield p + "attn.out_proj.weight", lin(ed, ed)
        yield p + "attn.out_proj.bias", small(ed)
        yield p + "attn.gru.weight", lin(8, hd)
        yield p + "attn.gru.bias", small(8, 0.2)
        yield p + "attn.gru_const", (1.0 + 0.1 * rng.standard_normal((1, nh, 1, 1))).astype(f32)
        yield p + "norm1.weight", norm_w(ed)
        yield p + "norm1.bias", small(ed)
        yield p + "ffn.w1.weight", lin(ff, ed)
        yield p + "ffn.w1.bias", small(ff)
        yield p + "ffn.w2.weight", lin(ed, ff)
        yield p + "ffn.w2.bias", small(ed)
        yield p + "norm2.weight", norm_w(ed)
        yield p + "norm2.bias", small(ed)


def write_wavlm_gguf(path, preset="base", seed=0):
    """WavLM SSL extractor GGUF with the names/KVs of scripts/convert_wavlm_base_plus_to_gguf.py:118-186."""
    import gguf

    cfg = WAVLM_PRESETS[preset]
    w = gguf.GGUFWriter(path, "wavlm-ssl")
    w.add_name(f"WavLM SSL extractor synthetic ({preset}, seed {seed})")
    for k in ("sample_rate", "n_layers", "n_heads", "head_dim", "embed_dim", "num_buckets", "max_distance"):
        w.add_uint32("wavlm." + k, int(cfg[k]))
    w.add_float32("wavlm.layer_norm_eps", 1e-5)
    for i, (k, st) in enumerate(zip(WAVLM_CONV_K, WAVLM_CONV_S)):
        w.add_uint32(f"wavlm.feat.conv{i}.kernel", k)
        w.add_uint32(f"wavlm.feat.conv{i}.stride", st)
    n_params = 0
    for name, arr in synth_wavlm_tensors(cfg, seed):
        arr = np.ascontiguousarray(arr)
        n_params += arr.size
        w.add_tensor(name, arr)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return n_params


def write_gguf(path, preset="default", seed=0, global_encoder=None):
    """global_encoder: None (decoder only, as the committed fixtures were made) or a GE_PRESETS key."""
    import gguf

    cfg = PRESETS[preset]
    w = gguf.GGUFWriter(path, "miocodec-dec")
    w.add_name(f"MioCodec decoder synthetic ({preset}, seed {seed})")
    u32 = w.add_uint32
    u32("miocodec.model_type", 0)
    u32("miocodec.dynamic_global", int(cfg.get("dynamic_global", 1)))
    for k in ("sample_rate", "n_fft", "hop_length", "samples_per_token",
              "prenet_layers", "prenet_dim", "prenet_heads", "prenet_ff", "prenet_window",
              "decoder_layers", "decoder_dim", "decoder_heads", "decoder_ff", "decoder_window",
              "decoder_adanorm_dim", "resnet_blocks", "resnet_groups"):
        u32("miocodec." + k, int(cfg[k]))
    u32("miocodec.n_mels", 0)
    u32("miocodec.wave_upsampler_layers", len(cfg["up_factors"]))
    u32("miocodec.has_vocoder", 0)
    w.add_float32("miocodec.rope_theta", 10000.0)
    w.add_float32("miocodec.norm_eps", 1e-5)
    w.add_float32("miocodec.group_norm_eps", 1e-6)
    if cfg["up_factors"]:
        w.add_tensor("miocodec.wave_upsampler.factors", np.asarray(cfg["up_factors"], dtype=np.int32))
        w.add_tensor("miocodec.wave_upsampler.kernel_sizes", np.asarray(cfg["up_kernels"], dtype=np.int32))
    extra = []
    if global_encoder is not None:
        ge = GE_PRESETS[global_encoder]
        for k in ("input_channels", "output_channels", "dim", "intermediate_dim", "num_layers"):
            u32("miocodec.global_encoder." + k, int(ge[k]))
        extra = list(synth_global_encoder_tensors(ge, seed))
    h = hashlib.sha256()
    n_params = 0
    for name, arr in list(synth_tensors(cfg, seed)) + extra:
        arr = np.ascontiguousarray(arr)
        h.update(name.encode())
        h.update(arr.tobytes())
        n_params += arr.size
        w.add_tensor(name, arr)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return h.hexdigest(), n_params


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--preset", default="default", choices=sorted(PRESETS))
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    digest, n = write_gguf(a.out, a.preset, a.seed)
    print(f"{a.out}: preset={a.preset} seed={a.seed} params={n} sha256={digest}", file=sys.stderr)


if __name__ == "__main__":
    main()
