import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def snr_db(got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    err = got - ref
    den = float((ref * ref).sum())
    num = float((err * err).sum())
    if num == 0.0:
        return 300.0
    return 10.0 * np.log10(den / num) if den > 0 else -300.0


@pytest.fixture(scope="session")
def emb():
    return np.load(os.path.join(GOLD, "en_female_emb.npy"))


@pytest.fixture(scope="session")
def gguf_paths(tmp_path_factory):
    """Synthetic-weight GGUFs in the reference layout (tools/make_synth_gguf.py, seed 0), written on demand."""
    from make_synth_gguf import write_gguf

    d = tmp_path_factory.mktemp("gguf")
    cache = {}

    def get(preset):
        if preset not in cache:
            p = str(d / f"{preset}.gguf")
            sha, _ = write_gguf(p, preset, 0)
            cache[preset] = (p, sha)
        return cache[preset]

    return get


def load_golden(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


GOLDEN_CASES = ["tiny_T50", "tiny_T1", "tiny_T133", "upsampler_tiny_T20", "default_T24", "upsampler_T12",
                "tiny_lerp3_T37", "tiny_lerp15_T41", "tiny_static_T50"]
# BASELINE configs 1 / 2 at full size: utterances 0 / 17 / 63 of the bench batch, reduced (tools/gen_golden.py FULL_CASES)
FULL_CASES = ["default_T400", "default_T400_u17", "default_T400_u63"]


def case_embedding(z, emb):
    """Static (dynamic_global = 0) fixtures take no embedding (lib.cpp:1222-1225)."""
    return None if "static" in str(z["preset"]) else emb


def full_case_audio_windows(z, audio):
    """The samples of `audio` that a reduced full-size fixture kept (concatenated windows)."""
    w = int(z["audio_window"])
    return np.concatenate([audio[int(s):int(s) + w] for s in z["audio_starts"]])
ISTFT_CASES = ["istft_120_30", "istft_392_98", "istft_1280_320", "istft_1920_480", "istft_2048_512"]
