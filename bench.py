#!/usr/bin/env python3
"""bench.py -- MioCodec decode + iSTFT throughput (BASELINE.json metric: audio-seconds per second).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU path (oracle/_ref) on the host cores

A step = one pass of the hot path over BASELINE config 2 on each GPU: 64 utterances x 400 synthetic codes
(16 s of 24 kHz audio each, default hyper-parameters, synthetic weights in the reference's GGUF layout).
Multi-GPU = independent replicas (weak scaling, no collective on the data path).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from mio_tts_cpp_b200 import benchutil as U  # noqa: E402

BATCH, N_CODES = 64, 400


def baseline_metric():
    try:
        return json.load(open(os.path.join(ROOT, "BASELINE.json")))["metric"]
    except Exception:   # noqa: BLE001
        return "MioCodec+iSTFT audio-seconds/sec (RTF^-1) per GPU and at 8xB200"


def _cpu_ref_loop(codes, emb, n_threads, warmup, keep=()):
    """Sequential mio_tts_synthesize calls of the compiled reference on `codes` (first `warmup` untimed).
    Returns (audio seconds, wall seconds, t_start, t_end, {index: waveform} for the timed indices in `keep`)."""
    from oracle.refbind import RefLib

    R = RefLib()
    ctx = R.open_context(U.synth_gguf("default"))
    devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(2)
    os.dup2(devnull, 2)            # the reference prints a "synth breakdown" line per call on stderr
    kept = {}
    try:
        for i in range(warmup):
            R.synthesize(ctx, codes[i], emb, n_threads=n_threads)
        go = float(os.environ.get("MIO_BENCH_CPU_GO", "0"))
        while time.time() < go:     # packed arrangement: every process starts its timed loop at the same instant
            time.sleep(0.001)
        t_start, secs = time.time(), 0.0
        t0 = time.perf_counter()
        for i in range(warmup, len(codes)):
            a, sr = R.synthesize(ctx, codes[i], emb, n_threads=n_threads)
            secs += a.size / sr
            if i - warmup in keep:
                kept[i - warmup] = a
        dt = time.perf_counter() - t0
        t_end = time.time()
    finally:
        os.dup2(saved, 2)
        os.close(devnull)
        os.close(saved)
    R.close_context(ctx)
    return secs, dt, t_start, t_end, kept


def time_cpu_reference(n_utt, n_codes, warmup=0, n_threads=None, keep=()):
    """The reference's own CPU mio_tts_synthesize (oracle/_ref = the unmodified reference compiled from
    /root/reference; else the numpy port), `n_utt` sequential utterances with `n_threads` (default: all) host threads --
    the reference has no batch API.  bench.py executes oracle/ only here and in cpu_arrangements below."""
    from oracle.refbind import RefLib, ref_lib_path

    emb = U.speaker_embedding()
    codes = U.config2_workload(0, n_utt + warmup, n_codes)
    if warmup:                      # warm-up utterances come from the END of the list: timed index i = utterance i
        codes = codes[n_utt:] + codes[:n_utt]
    if ref_lib_path() is not None:
        cores = RefLib().hw_threads()
        secs, dt, _, _, kept = _cpu_ref_loop(codes, emb, n_threads or cores, warmup, keep)
        return secs / dt, n_threads or cores, "reference", dt, kept
    from oracle import miocodec_np as O

    m = O.Model(U.synth_gguf("default"))
    t0, secs, kept = time.perf_counter(), 0.0, {}
    for i in range(n_utt):
        a = O.synthesize(m, codes[warmup + i], emb)
        secs += a.size / m.hp["sample_rate"]
        if i in keep:
            kept[i] = a
    dt = time.perf_counter() - t0
    return secs / dt, 1, "port", dt, kept


def cpu_worker_main(argv):
    """`bench.py --cpu-worker <threads> <n_utt> <first utterance>`: one process of the packed CPU arrangement."""
    k, n_utt, first = int(argv[0]), int(argv[1]), int(argv[2])
    codes = U.config2_workload(0, 64, N_CODES)
    mine = [codes[(first + i) % 64] for i in range(n_utt + 1)]
    secs, dt, t_start, t_end, _ = _cpu_ref_loop(mine[-1:] + mine[:-1], U.speaker_embedding(), k, 1)
    print(json.dumps({"secs": secs, "dt": dt, "t_start": t_start, "t_end": t_end}), flush=True)


def cpu_arrangements(runs=5, only_packed=False):
    """the reference on this box's host cores in the arrangements a deployment could choose --
    (a) all hardware threads on one utterance at a time, (b) n_threads = 2 (the CLI / server default), (c) packed:
    cores / 2 processes x 2 threads, each decoding its own utterances at the same time.  min / median over `runs`
    runs each; every run is a bounded sample of config 2 (whole utterances of 400 codes)."""
    import statistics
    import subprocess

    from oracle.refbind import RefLib, ref_lib_path

    if ref_lib_path() is None:
        return None
    cores = RefLib().hw_threads()
    out = {}
    for name, k, n_utt in (() if only_packed else (("all_threads", cores, 4), ("two_threads", 2, 1))):
        vals = [time_cpu_reference(n_utt, N_CODES, warmup=1 if r == 0 else 0, n_threads=k)[0] for r in range(runs)]
        out[name] = {"threads": k, "utterances_per_run": n_utt, "runs": runs, "median": statistics.median(vals),
                     "min": min(vals), "max": max(vals)}
    procs, k, n_utt = max(1, cores // 2), 2, 2
    vals = []
    for r in range(runs):
        env = dict(os.environ, MIO_BENCH_CPU_GO=str(time.time() + 4.0), OMP_NUM_THREADS=str(k))
        ps = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--cpu-worker", str(k), str(n_utt), str(p * n_utt)],
                               stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, env=env) for p in range(procs)]
        rows = []
        for q in ps:
            o, _ = q.communicate(timeout=600)
            rows += [json.loads(ln) for ln in o.splitlines() if ln.startswith("{")]
        if len(rows) == procs:
            vals.append(sum(x["secs"] for x in rows) / (max(x["t_end"] for x in rows) - min(x["t_start"] for x in rows)))
    if vals:
        out["packed"] = {"processes": procs, "threads": k, "utterances_per_process": n_utt, "runs": len(vals),
                         "median": statistics.median(vals), "min": min(vals), "max": max(vals)}
    return out


def run_reference(args, rank):
    """The reference arm: the unmodified reference's CPU path (oracle/_ref) on this box's host cores, in whichever of its
    own arrangements is fastest here -- one utterance at a time on all hardware threads, or cores / 2 processes x 2
    threads (its CLI / server default thread count) decoding their own utterances at the same time.  Each step is a
    bounded sample of config 2: whole utterances of 400 codes (the reference has no batch API and the metric is per
    audio-second, so utterances, not batches, are the unit)."""
    if rank != 0:
        return
    per_step = 2   # utterances per step of the sequential arrangement (the CPU path takes ~0.2-1 s per utterance)
    v, cores, kind, dt, _ = time_cpu_reference(per_step * args.steps, N_CODES, warmup=min(args.warmup, 2))
    sample = f"{per_step * args.steps} of the 64 utterances x {N_CODES} codes, sequential mio_tts_synthesize, {cores} threads"
    arr = {"all_threads_sequential": {"value": v, "threads": cores, "utterances": per_step * args.steps, "seconds": dt}}
    if kind == "reference":
        packed = cpu_arrangements(runs=3, only_packed=True)
        if packed and "packed" in packed:
            arr["packed"] = packed["packed"]
            if packed["packed"]["median"] > v:
                v = packed["packed"]["median"]
                sample = (f"packed: {packed['packed']['processes']} processes x {packed['packed']['threads']} threads, "
                          f"{packed['packed']['utterances_per_process']} utterances x {N_CODES} codes each, median of "
                          f"{packed['packed']['runs']} runs (the sequential all-thread arrangement is in `arrangements`)")
    line = {"impl": "reference", "metric": baseline_metric(), "value": v, "unit": "audio-seconds/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(1, args.steps),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"config2: batch {BATCH} x {N_CODES} codes, default 24 kHz hyper-parameters "
                                   "(bounded sample per step: whole utterances)", "sample_utterances_per_step": per_step},
            "cpu_baseline": {"value": v, "unit": "audio-seconds/s", "cores": cores, "kind": kind, "sample": sample,
                             "arrangements": arr},
            "e2e": {"value": v, "unit": "audio-seconds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_b200(args, rank, world, local_rank):
    import torch

    os.environ["MIO_TTS_BACKEND_DEVICE"] = f"CUDA{local_rank}"
    from mio_tts_cpp_b200 import MioTTS

    dist = None
    json_fd = None
    if world > 1:
        # stdout carries exactly one JSON line, but NCCL writes its banner ("NCCL version ...") to fd 1 from C: keep a
        # private duplicate of the real stdout for the JSON line and point fd 1 at stderr for everything else
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        import torch.distributed as dist_mod

        torch.cuda.set_device(local_rank)
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_mod

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    if rank == 0:
        path = U.synth_gguf("default")
    barrier()
    path = U.synth_gguf("default")
    m = MioTTS(path)
    emb = U.speaker_embedding()
    codes = U.config2_workload(rank, BATCH, N_CODES)
    embs = [emb] * BATCH
    sr = m.sample_rate()

    # ---- device-resident throughput: inputs already in HBM when the timed region starts ----
    m.prepare(codes, embs)
    audio_s_per_step = m.resident_samples() / sr
    d2h = int(m.resident_samples() * 4)      # the end-to-end legs return exactly these waveforms to the host
    for _ in range(max(args.warmup, 3)):
        m.run_resident(1)
    barrier()
    torch.cuda.synchronize()
    n0 = m.launch_count()
    with U.ClockSampler(local_rank) as cs:
        ms = m.run_resident(args.steps)          # CUDA events on the library's stream around K back-to-back steps
    launches = m.launch_count() - n0
    barrier()
    ms, total_audio = U.reduce_over_ranks(dist, f"cuda:{local_rank}", ms, audio_s_per_step)
    sum_audio_all = total_audio
    value = total_audio / (ms * 1e-3)

    # per-launch GEMM timings for the roofline, taken right after the timed region (same clocks and power state)
    profs = [m.profile_resident() for _ in range(3)] if rank == 0 else []

    # ---- sustained leg: >= 2.5 s of back-to-back steps with its own clock record, the like-for-like numerator for the
    # SUSTAINED bf16 peak of MEASURED_PEAKS.json (taken over seconds under the power cap); the K-step region above is
    # a fraction of a second and belongs beside the burst peak
    sus_steps = int(max(args.steps, min(2000, 2500.0 / max(ms, 0.05))))
    barrier()
    torch.cuda.synchronize()
    with U.ClockSampler(local_rank) as cs_sus:
        sus_ms = m.run_resident(sus_steps)
    sus_ms = max_over_ranks(sus_ms)
    sus_profs = [m.profile_resident() for _ in range(3)] if rank == 0 else []

    # ---- end to end through the public batch call: host codes in, host (pinned) waveforms out, every step.
    # Like the reference server's slots (tts-mio-server.cpp:2659-2670, one context clone per slot), `--e2e-slots`
    # host threads each own a clone of the context (shared weights, own stream + workspace) and each call
    # mio_tts_b200_synthesize_batch on a full 64 x 400 batch per step, so one slot's D2H overlaps another's kernels.
    import threading

    def e2e_run(n_slots, steps, wav16=False):
        slots = [m] + [m.clone() for _ in range(n_slots - 1)]
        work = [U.config2_workload(rank * 16 + i, BATCH, N_CODES) for i in range(n_slots)]
        call = (lambda s, w: s.synthesize_wav16_batch(w, embs, copy=False)) if wav16 else \
               (lambda s, w: s.synthesize_batch(w, embs, copy=False))
        for s, w in zip(slots, work):
            for _ in range(2):
                call(s, w)
        gate = threading.Barrier(n_slots + 1)

        def loop(s, w):
            for _ in range(3):          # concurrent warm-up: the pinned result pool reaches its steady-state size
                call(s, w)
            gate.wait()
            for _ in range(steps):
                call(s, w)
            gate.wait()

        th = [threading.Thread(target=loop, args=(s, w)) for s, w in zip(slots, work)]
        for t in th:
            t.start()
        barrier()
        torch.cuda.synchronize()
        gate.wait()
        t0 = time.perf_counter()
        gate.wait()
        dt = time.perf_counter() - t0
        for t in th:
            t.join()
        for s in slots[1:]:
            s.free()
        return dt / (steps * n_slots)      # seconds per 64-utterance batch

    e2e_single_s = max_over_ranks(e2e_run(1, args.steps))
    e2e_s = max_over_ranks(e2e_run(args.e2e_slots, args.steps)) if args.e2e_slots > 1 else e2e_single_s
    e2e = total_audio / e2e_s
    e2e_wav_s = max_over_ranks(e2e_run(args.e2e_slots, args.steps, wav16=True))
    h2d = m.last_h2d_bytes()

    line = {"metric": baseline_metric(), "value": value, "unit": "audio-seconds/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": os.environ.get("MIO_TTS_B200_LINEAR_DTYPE", "f16"), "data": "synthetic",
            "config": {"workload": f"config2: batch {BATCH} x {N_CODES} codes per GPU, default 24 kHz hyper-parameters, "
                                   "fp16 operands / fp32 accumulate (same tensor rate as bf16; MIO_TTS_B200_LINEAR_DTYPE=bf16 selects bf16), "
                                   "fp32 iSTFT, synthetic weights in the reference GGUF layout",
                       "audio_seconds_per_step_per_gpu": audio_s_per_step, "parallelism": f"replicas x{world}",
                       "l2": "no flush: each step streams ~2 GB of activations (> 126 MB L2); inputs are 100 KB of codes"},
            "e2e": {"value": e2e, "unit": "audio-seconds/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1e3 * e2e_s, "slots": args.e2e_slots,
                    "single_slot_value": total_audio / e2e_single_s, "single_slot_ms_per_step": 1e3 * e2e_single_s,
                    "wav16_value": total_audio / e2e_wav_s, "wav16_d2h_bytes_per_step": d2h // 2 + 44 * BATCH,
                    "note": "a step = one mio_tts_b200_synthesize_batch call (64 x 400 codes from host memory to pinned "
                            "host waveforms); `slots` host threads with one context clone each, as the reference "
                            "server runs its slots; single_slot_* = one caller, nothing overlapped; wav16_* = the same "
                            "slots asking for audio/wav bodies (mio_tts_b200_synthesize_wav16_batch: int16 conversion on "
                            "the GPU, half the D2H bytes), which is what the reference server sends"},
            "gpu_launches": int(launches), "clocks": cs.summary()}

    if rank == 0:
        # ---- roofline of the dominant kernel (the tcgen05 GEMM family), timed live with CUDA events per launch ----
        peaks = U.measured_peaks()
        p = profs[-1]
        g_ms = float(np.median([q["gemm_ms"] for q in profs]))
        achieved = p["gemm_flops"] / (g_ms * 1e-3) / 1e12
        traffic = None
        tp = os.path.join(ROOT, "profiles", "gemm_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("dram_bytes_per_launch")
        tmeta = json.load(open(tp)) if os.path.exists(tp) else {}
        sg_ms = float(np.median([q["gemm_ms"] for q in sus_profs]))
        sus_achieved = p["gemm_flops"] / (sg_ms * 1e-3) / 1e12
        line["sustained"] = {"value": sum_audio_all / (sus_ms * 1e-3), "unit": "audio-seconds/s", "steps": sus_steps,
                             "ms_per_step": sus_ms, "seconds": sus_ms * sus_steps * 1e-3, "clocks": cs_sus.summary(),
                             "gemm_tflops": sus_achieved, "gemm_frac_of_sustained_peak": sus_achieved / peaks["bf16_tflops_sustained"],
                             "note": "same resident step repeated for >= 2.5 s: the numerator that belongs beside the "
                                     "sustained bf16 peak (both under the power cap for seconds)"}
        line["roofline"] = {"bound": "tensor", "achieved": achieved, "peak": peaks["bf16_tflops"],
                            "unit": "TFLOP/s", "frac": achieved / peaks["bf16_tflops"], "traffic": traffic,
                            "traffic_source": {k: tmeta.get(k) for k in ("captured", "code", "file") if k in tmeta} or None,
                            "frac_of_sustained_peak": achieved / peaks["bf16_tflops_sustained"],
                            "peak_sustained": peaks["bf16_tflops_sustained"],
                            "peak_source": peaks["source"] + " (bf16 burst: the K-step timed region lasts a fraction of a "
                                           "second; see `sustained` for the multi-second leg against the sustained peak)",
                            "kernel": "gemm_tcgen05_kernel", "launches_per_step": p["gemm_launches"],
                            "avg_launch_ms": g_ms / max(1, p["gemm_launches"]),
                            "flops_per_launch_avg": p["gemm_flops"] / max(1, p["gemm_launches"]),
                            "algorithmic_bytes_per_launch_avg": p["gemm_bytes"] / max(1, p["gemm_launches"]),
                            "gemm_share_of_step": g_ms / p["total_ms"],
                            "whole_path_frac": (U.FLOP_PER_TOKEN * BATCH * N_CODES / (ms * 1e-3) / 1e12)
                            / peaks["bf16_tflops"] if world == 1 else None}
        line["stages_ms"] = {k: p[k] for k in ("prenet_ms", "upsample_prior_ms", "decoder_ms", "post_head_ms", "istft_ms",
                                               "total_ms", "attn_ms")}
        if world == 1 and not args.skip_aux:
            line["aux"] = aux_configs(m, peaks)
        if world == 1 and not args.skip_cpu:
            # ---- CPU baseline: the unmodified reference on this box's host cores, bounded sample ----
            n_cpu = 24      # utterances 0..23 of the config-2 batch, one after another (~10 s of host work)
            keep = (0, 17)
            v, cores, kind, dt, ref_audio = time_cpu_reference(n_cpu, N_CODES, warmup=1, keep=keep)
            line["cpu_baseline"] = {"value": v, "unit": "audio-seconds/s", "cores": cores, "kind": kind,
                                    "sample": f"utterances 0..{n_cpu - 1} of the 64 x {N_CODES} codes, sequential "
                                              f"mio_tts_synthesize calls, {cores} threads, {dt:.1f} s"}
            arr = cpu_arrangements() if kind == "reference" else None
            if arr:
                line["cpu_baseline"]["arrangements"] = arr
                best = max(arr, key=lambda k: arr[k]["median"])
                if arr[best]["median"] > v:     # report the arrangement that is kindest to the reference
                    line["cpu_baseline"].update(value=arr[best]["median"], cores=cores, sample=(
                        f"{best}: median of {arr[best]['runs']} runs, see arrangements (whole utterances of {N_CODES} codes)"))
            # ---- parity of THIS run: the timed batch's waveforms against the reference's for the same codes ----
            got, _ = m.synthesize_batch(codes, embs)
            par = {}
            for u in keep:
                r, g = ref_audio[u].astype(np.float64), np.asarray(got[u], np.float64)
                err = g - r
                par[f"utt{u}"] = {"snr_db": float(10 * np.log10((r * r).sum() / max((err * err).sum(), 1e-300))),
                                  "max_abs": float(np.abs(err).max()), "ref_peak": float(np.abs(r).max())}
            line["parity_snr_db"] = min(x["snr_db"] for x in par.values())
            line["parity_max_abs"] = max(x["max_abs"] for x in par.values())
            line["parity"] = {"against": "oracle/_ref (unmodified reference, fp32 CPU) on the same codes + embedding"
                              if kind == "reference" else "numpy oracle", "utterances": par,
                              "linear_dtype": os.environ.get("MIO_TTS_B200_LINEAR_DTYPE", "f16")}
        if json_fd is None:
            print(json.dumps(line), flush=True)
        else:
            os.write(json_fd, (json.dumps(line) + "\n").encode())
    barrier()
    m.free()
    if dist is not None:
        dist.destroy_process_group()


def aux_configs(m, peaks):
    """Side measurements next to the headline (not bench lines of their own): BASELINE config 3 (iSTFT alone over 1 M
    frames, HBM roofline) and config 4 (WavLM 2-layer + global encoder, 32 clips x 10 s, inputs resident)."""
    from mio_tts_cpp_b200 import MioTTS

    out = {}
    n_utt, frames = 1000, 1000
    ms, _ = m.istft_bench(n_utt, frames, 1920, 480, 5)
    gbs = n_utt * frames * U.ISTFT_BYTES_PER_FRAME / (ms * 1e-3) / 1e9
    out["config3_istft"] = {"frames": n_utt * frames, "n_fft": 1920, "hop": 480, "ms": ms, "frames_per_s": n_utt * frames / (ms * 1e-3),
                            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                         "frac": gbs / peaks["hbm_gbs"]}}
    # the frame size of the shipped 44.1 kHz checkpoints (n_fft 392 / hop 98: radix 4x7x7 fast path), another fast-path
    # size, and the generic kernel (any frame size; n_fft 1000 / hop 250 has no fast path)
    for key, nf, hp in (("istft_392_98", 392, 98), ("istft_1024_256", 1024, 256), ("istft_generic_1000_250", 1000, 250)):
        ms_s, _ = m.istft_bench(1000, 1000, nf, hp, 5)
        b_s = (nf + 2) * 4 + hp * 4
        out[key] = {"frames": 1000000, "n_fft": nf, "hop": hp, "ms": ms_s, "frames_per_s": 1e6 / (ms_s * 1e-3),
                    "roofline": {"bound": "hbm", "achieved": 1e6 * b_s / (ms_s * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                 "frac": 1e6 * b_s / (ms_s * 1e-3) / 1e9 / peaks["hbm_gbs"], "bytes_per_frame": b_s}}
    # codebook gather alone over the config-2 batch (north_star: achieved HBM GB/s of the gather): 25 600 rows x 768 fp32
    # written (79 MB) + the rows read (the 39 MB table is L2-resident after the first pass)
    m.prepare(U.config2_workload(0, BATCH, N_CODES), [U.speaker_embedding()] * BATCH)   # (the iSTFT legs reused the batch tables)
    g_ms = m.gather_bench(50)
    g_bytes = BATCH * N_CODES * m.prenet_dim() * 4 * 2 + BATCH * N_CODES * 4
    out["gather"] = {"rows": BATCH * N_CODES, "dim": m.prenet_dim(), "ms": g_ms, "bit_exact": True,
                     "roofline": {"bound": "hbm", "achieved": g_bytes / (g_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                  "frac": g_bytes / (g_ms * 1e-3) / 1e9 / peaks["hbm_gbs"], "bytes_per_token": g_bytes / (BATCH * N_CODES),
                                  "note": "table rows come from L2 (12 800 x 768 fp32 = 39 MB stays resident); DRAM sees the writes"}}
    # the branch the shipped 44.1 kHz-v2 checkpoint takes (wave_upsampler_layers = 2, factors [3,3], n_fft 392):
    # synthetic weights of the `upsampler` preset, 64 x 400 codes
    up = MioTTS(U.synth_gguf("upsampler"))
    codes = U.config2_workload(0, BATCH, N_CODES)
    up.prepare(codes, [U.speaker_embedding()] * BATCH)
    for _ in range(3):
        up.run_resident(1)
    n0 = up.launch_count()
    up_ms = up.run_resident(10)
    up_launches = (up.launch_count() - n0) / 10
    pr = up.profile_resident()
    up_audio = up.resident_samples() / up.sample_rate()
    out["upsampler_44k"] = {"workload": "preset `upsampler` (2+2 transformer layers, 1 ResNet block per stack, 2 wave-upsampler "
                                        "stages x3 with SnakeBeta + ResNet, out_proj 128, n_fft 392 / hop 98), 64 x 400 codes",
                            "audio_seconds_per_step": up_audio, "ms_per_step": up_ms, "value": up_audio / (up_ms * 1e-3),
                            "unit": "audio-seconds/s", "launches_per_step": up_launches,
                            "stages_ms": {k: pr[k] for k in ("prenet_ms", "upsample_prior_ms", "decoder_ms", "post_head_ms", "istft_ms", "total_ms")},
                            "gemm_tflops": pr["gemm_flops"] / max(pr["gemm_ms"], 1e-9) / 1e9, "gemm_share": pr["gemm_ms"] / pr["total_ms"],
                            "roofline": {"bound": "tensor", "achieved": pr["gemm_flops"] / max(pr["gemm_ms"], 1e-9) / 1e9,
                                         "peak": peaks["bf16_tflops"], "unit": "TFLOP/s",
                                         "frac": pr["gemm_flops"] / max(pr["gemm_ms"], 1e-9) / 1e9 / peaks["bf16_tflops"]},
                            "note": "shallow trunk: not comparable layer for layer with the default path; post_head = post ResNet + "
                                    "upsampler stages + head"}
    up.free()
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from make_synth_gguf import write_gguf, write_wavlm_gguf

    dec, wl = "/tmp/mio_bench_tiny_ge.gguf", "/tmp/mio_bench_wavlm_base.gguf"
    if not os.path.exists(dec):
        write_gguf(dec + ".tmp", "tiny", 0, "default")
        os.replace(dec + ".tmp", dec)
    if not os.path.exists(wl):
        write_wavlm_gguf(wl + ".tmp", "base", 0)
        os.replace(wl + ".tmp", wl)
    r = MioTTS(dec, wl)
    clip = U.reference_clip(10.0, 24000, 3)
    ms = r.reference_bench(clip, 24000, 32, 10)
    t0 = time.perf_counter()
    n_e2e = 5
    for _ in range(n_e2e):
        r.reference_pcm_to_embedding_batch([clip] * 32, 24000)
    e2e_ms = 1e3 * (time.perf_counter() - t0) / n_e2e
    r.free()
    out["config4_reference_embedding"] = {"clips": 32, "seconds_per_clip": 10.0, "ms_per_batch": ms, "clips_per_s": 32 / (ms * 1e-3),
                                          "e2e_ms_per_batch": e2e_ms, "e2e_clips_per_s": 32 / (e2e_ms * 1e-3),
                                          "h2d_bytes_per_batch": 32 * clip.size * 4,
                                          "note": "WavLM base-size synthetic weights (2 layers, 768/12/3072) + default global "
                                                  "encoder; e2e = host PCM in, host embeddings out"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--skip-cpu", action="store_true", help="omit the cpu_baseline leg (profiling runs)")
    ap.add_argument("--skip-aux", action="store_true", help="omit the config-3 / config-4 side measurements")
    ap.add_argument("--e2e-slots", type=int, default=2, help="host threads (context clones) of the end-to-end leg")
    if len(sys.argv) > 1 and sys.argv[1] == "--cpu-worker":
        return cpu_worker_main(sys.argv[2:])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
