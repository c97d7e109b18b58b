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
}
VOCAB = 12800  # codes 0..12799, /root/reference/src/mio-tts-lib.cpp:30-31


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
        for n in ("attn_cond", "ffn_cond"):
            yield p + n + ".weight", lin(3 * dd, ad, 0.5)
            b = small(3 * dd)
            b[2 * dd:] += 0.5  # gate ~ 0.5 so every block contributes
            yield p + n + ".bias", b
        for n in ("attn_q", "attn_k", "attn_v", "attn_output"):
            yield p + n + ".weight", lin(dd, dd)
        yield p + "ffn_gate.weight", lin(df, dd)
        yield p + "ffn_up.weight", lin(df, dd)
        yield p + "ffn_down.weight", lin(dd, df)
    yield "wave_decoder.norm_cond.weight", lin(2 * dd, ad, 0.5)
    yield "wave_decoder.norm_cond.bias", small(2 * dd)
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


def write_gguf(path, preset="default", seed=0):
    import gguf

    cfg = PRESETS[preset]
    w = gguf.GGUFWriter(path, "miocodec-dec")
    w.add_name(f"MioCodec decoder synthetic ({preset}, seed {seed})")
    u32 = w.add_uint32
    u32("miocodec.model_type", 0)
    u32("miocodec.dynamic_global", 1)
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
    h = hashlib.sha256()
    n_params = 0
    for name, arr in synth_tensors(cfg, seed):
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
