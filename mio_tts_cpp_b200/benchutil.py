"""Helpers for bench.py: workload construction, clock sampling, measured-peak lookup.  No compute here."""
import json
import os
import statistics
import subprocess
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# algorithmic work of the decode path at the default hyper-parameters
FLOP_PER_TOKEN = 228.25e6          # window-65 attention, multiply-add = 2
ISTFT_BYTES_PER_FRAME = 9608       # (n_fft + 2) * 4 read + hop * 4 written


def measured_peaks():
    """Roofline denominators: MEASURED_PEAKS.json (driver-written) or the fallback stated in B200_PROFILING.md."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def config2_workload(rank=0, batch=64, n_codes=400):
    """BASELINE config 2: `batch` utterances x `n_codes` synthetic codes, seeds 42..."""
    return [np.random.default_rng(42 + rank * batch + i).integers(0, 12800, n_codes).astype(np.int32)
            for i in range(batch)]


def reduce_over_ranks(dist, device, step_ms, audio_s):
    """Multi-GPU aggregation of a replica run (no data-path collective): the step time is the MAX over ranks, the work
    is the SUM over ranks.  `dist` = torch.distributed (initialised) or None for a single process."""
    if dist is None:
        return float(step_ms), float(audio_s)
    import torch

    t = torch.tensor([step_ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    a = torch.tensor([audio_s], dtype=torch.float64, device=device)
    dist.all_reduce(a, op=dist.ReduceOp.SUM)
    return float(t.item()), float(a.item())


def speaker_embedding():
    return np.load(os.path.join(ROOT, "tests", "golden", "en_female_emb.npy"))


def synth_gguf(preset="default", tag=""):
    """Synthetic weights in the reference's GGUF layout (the real miocodec.gguf cannot be downloaded here)."""
    import sys

    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from make_synth_gguf import write_gguf

    path = f"/tmp/mio_bench_{preset}{tag}.gguf"
    if not os.path.exists(path):
        tmp = path + f".{os.getpid()}.tmp"
        write_gguf(tmp, preset, 0)
        os.replace(tmp, path)
    return path


def reference_clip(seconds, sample_rate=24000, seed=3):
    """Synthetic reference-audio clip (BASELINE config 4 stand-in: the reference's resources/ja_female.wav cannot travel):
    a dozen amplitude-modulated partials plus noise, quantised to int16 steps so a 16-bit WAV holds it exactly."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * sample_rate))
    t = np.arange(n) / sample_rate
    x = np.zeros(n)
    for _ in range(12):
        f, a, ph = rng.uniform(80, 3000), rng.uniform(0.05, 0.3), rng.uniform(0, 6.28)
        env = 0.5 * (1 + np.sin(2 * np.pi * rng.uniform(0.3, 3) * t + rng.uniform(0, 6.28)))
        x += a * env * np.sin(2 * np.pi * f * t + ph)
    x += 0.02 * rng.standard_normal(n)
    x = x / np.abs(x).max() * 0.8
    return (np.round(x * 32767) / 32768.0).astype(np.float32)


def write_wav16(path, pcm, sample_rate):
    """Mono PCM16 WAV of a clip whose samples are int16 steps / 32768 (reference_clip)."""
    import wave

    with wave.open(path, "wb") as w:
        w.setnchannels(1)
        w.setsampwidth(2)
        w.setframerate(sample_rate)
        w.writeframes(np.round(np.asarray(pcm, np.float64) * 32768.0).astype("<i2").tobytes())


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while a timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0, period_s=0.1):
        self.index, self.period, self.rows, self._stop, self._t = index, period_s, [], threading.Event(), None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().splitlines()[0].split(",")]
                self.rows.append(parts)
            except Exception:   # noqa: BLE001 -- nvidia-smi absent or busy: clocks stay unknown
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=10)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower() == "active" for r in self.rows)]
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows), "power_w_max": max(pw) if pw else None}
