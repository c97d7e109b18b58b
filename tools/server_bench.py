#!/usr/bin/env python3
"""BASELINE config 5: the reference's own mio-tts-server (unmodified source, tools/build_ref_apps.sh) on top of
libmio_tts_b200.so, `--parallel S` slots sharded over G GPUs (`--mio-backend-devices CUDA0,...`), driven over HTTP with
inline "codes" requests (/root/reference/src/tts-mio-server.cpp:2105-2110, 2153-2453): POST /mio/tts/stream with
stream_tokens off answers with the binary audio/wav body (:3901-4006, 768 KB per 400-code request); POST /mio/tts
(:3630-3703) writes the WAV to the server's output directory and answers with JSON.

    python tools/server_bench.py --gpus 8 --parallel 64 --clients 8,64,256 --seconds 6 --out profiles/r02_server_8gpu.json

Everything above mio_tts_synthesize (cpp-httplib, JSON parsing, the float -> int16 loop, chunked audio/wav responses)
is the reference's code and runs on the host; the numbers therefore say what the swapped library delivers THROUGH that
layer, next to tools/slots_bench.cpp which drives the same slots without HTTP.  Clients are processes x threads of
http.client with keep-alive connections, each sending 400-code requests back to back (closed loop).
"""
import argparse
import http.client
import json
import multiprocessing as mp
import os
import socket
import subprocess
import sys
import time
import urllib.request

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BIN = os.path.join(ROOT, "tools", "ref_apps", "_build", "bin")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def start_server(model, emb_path, parallel, devices, log_path, batching):
    port = free_port()
    ld = os.pathsep.join([os.path.join(ROOT, "mio_tts_cpp_b200"), BIN, os.environ.get("LD_LIBRARY_PATH", "")])
    env = dict(os.environ, LD_LIBRARY_PATH=ld)
    if batching:
        env["MIO_TTS_B200_BATCHING"] = batching
    cmd = [os.path.join(BIN, "mio-tts-server"), "-mv", model, "--host", "127.0.0.1", "--port", str(port), "--parallel",
           str(parallel), "--mio-backend-devices", devices, "--reference-file-json", json.dumps({"key": "spk", "path": emb_path})]
    log = open(log_path, "w")
    p = subprocess.Popen(cmd, env=env, stdout=log, stderr=subprocess.STDOUT)
    t0 = time.time()
    while time.time() - t0 < 300:
        if p.poll() is not None:
            raise RuntimeError("server exited: " + open(log_path).read()[-3000:])
        try:
            if json.load(urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=2)).get("status") == "ok":
                return p, port, log
        except Exception:   # noqa: BLE001
            time.sleep(0.3)
    p.kill()
    raise RuntimeError("server did not come up: " + open(log_path).read()[-3000:])


def client_proc(port, n_threads, seconds, n_codes, seed, q, go, route):
    import threading

    rng = np.random.default_rng(seed)
    bodies = [json.dumps({"codes": [int(c) for c in rng.integers(0, 12800, n_codes)], "reference_key": "spk"}).encode()
              for _ in range(8)]
    lat, errs, nbytes = [], [0], [0]
    lock = threading.Lock()

    def loop(tid):
        c = http.client.HTTPConnection("127.0.0.1", port, timeout=120)
        mine, nb, i = [], 0, tid
        while time.time() < go.value:
            time.sleep(0.001)
        t_end = go.value + seconds
        while time.time() < t_end:
            t0 = time.perf_counter()
            try:
                c.request("POST", route, body=bodies[i % 8], headers={"Content-Type": "application/json"})
                r = c.getresponse()
                data = r.read()
                if r.status != 200:
                    raise RuntimeError(data[:200])
                nb += len(data)
                mine.append(time.perf_counter() - t0)
            except Exception:   # noqa: BLE001
                with lock:
                    errs[0] += 1
                c.close()
                c = http.client.HTTPConnection("127.0.0.1", port, timeout=120)
            i += 1
        with lock:
            lat.extend(mine)
            nbytes[0] += nb

    th = [threading.Thread(target=loop, args=(t,)) for t in range(n_threads)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    q.put((lat, errs[0], nbytes[0]))


def drive(port, clients, seconds, n_codes, route="/mio/tts/stream"):
    procs = min(clients, max(1, min(32, (os.cpu_count() or 8))))
    per = [clients // procs + (1 if i < clients % procs else 0) for i in range(procs)]
    q = mp.Queue()
    go = mp.Value("d", time.time() + 3.0)
    ps = [mp.Process(target=client_proc, args=(port, per[i], seconds, n_codes, 1000 + i, q, go, route)) for i in range(procs) if per[i]]
    for p in ps:
        p.start()
    lat, errs, nbytes = [], 0, 0
    for _ in ps:
        a, e, b = q.get(timeout=seconds + 300)
        lat += a
        errs += e
        nbytes += b
    for p in ps:
        p.join()
    lat = np.sort(np.asarray(lat)) if lat else np.zeros(1)
    n = int(lat.size)
    return {"route": route, "clients": clients, "client_processes": len(ps), "seconds": seconds, "requests": n, "errors": errs,
            "requests_per_s": n / seconds, "audio_seconds_per_s": n * n_codes * 960 / 24000 / seconds,
            "response_MB_per_s": nbytes / seconds / 1e6, "latency_ms": {"p50": 1e3 * float(lat[n // 2]), "p90": 1e3 * float(lat[int(n * 0.9)]),
                                                                       "p99": 1e3 * float(lat[min(n - 1, int(n * 0.99))]), "max": 1e3 * float(lat[-1])}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--parallel", type=int, default=8, help="server slots (BASELINE config 5: 8 slots over 8 GPUs)")
    ap.add_argument("--clients", default="8,64,256")
    ap.add_argument("--seconds", type=float, default=6.0)
    ap.add_argument("--codes", type=int, default=400)
    ap.add_argument("--batching", default="64,300,2", help="MIO_TTS_B200_BATCHING for the server ('' = clones only)")
    ap.add_argument("--routes", default="/mio/tts/stream,/mio/tts")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    from mio_tts_cpp_b200 import benchutil as U
    from mio_tts_cpp_b200 import lib as L

    model = U.synth_gguf("default")
    emb_path = "/tmp/mio_server_bench_spk.emb.gguf"
    L.embedding_save_gguf(emb_path, U.speaker_embedding())
    devices = ",".join(f"CUDA{i}" for i in range(a.gpus))
    out = {"config": f"mio-tts-server --parallel {a.parallel} --mio-backend-devices {devices}, default 24 kHz synthetic weights, "
                     f"{a.codes}-code inline requests over HTTP (keep-alive), MIO_TTS_B200_BATCHING={a.batching or 'off'}",
           "host_cpus": os.cpu_count(), "runs": []}
    p, port, log = start_server(model, emb_path, a.parallel, devices, "/tmp/mio_server_bench.log", a.batching)
    try:
        drive(port, min(8, a.parallel), 2.0, a.codes)          # warm-up: workspaces, pinned pool, graphs
        for c in [int(x) for x in a.clients.split(",") if x]:
            for route in a.routes.split(","):
                r = drive(port, c, a.seconds, a.codes, route)
                print(json.dumps(r), flush=True)
                out["runs"].append(r)
    finally:
        p.terminate()
        try:
            p.wait(timeout=30)
        except subprocess.TimeoutExpired:
            p.kill()
        log.close()
    tail = open("/tmp/mio_server_bench.log").read().splitlines()
    out["server_log_head"] = [ln for ln in tail if "workspace" in ln or "backend" in ln][:12]
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        json.dump(out, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
