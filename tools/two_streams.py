import sys, threading, time
sys.path.insert(0, ".")
from mio_tts_cpp_b200 import MioTTS, benchutil as U
m = MioTTS(U.synth_gguf("default"))
emb = U.speaker_embedding()
sr = m.sample_rate()
def bench(nclones, batch, iters=30):
    ctxs = [m] + [m.clone() for _ in range(nclones - 1)]
    for i, c in enumerate(ctxs):
        codes = U.config2_workload(i, batch, 400)
        c.prepare(codes, [emb] * batch)
        for _ in range(3): c.run_resident(1)
    secs = sum(c.resident_samples() for c in ctxs) / sr
    res = [0.0] * nclones
    def run(i): res[i] = ctxs[i].run_resident(iters)
    t0 = time.perf_counter()
    th = [threading.Thread(target=run, args=(i,)) for i in range(nclones)]
    [t.start() for t in th]; [t.join() for t in th]
    dt = time.perf_counter() - t0
    print(f"clones {nclones} x batch {batch}: wall {dt/iters*1e3:.3f} ms per round, {secs*iters/dt:.0f} audio-s/s (per-clone ms {['%.2f' % r for r in res]})", flush=True)
    for c in ctxs[1:]: c.free()
for _ in range(2):
    bench(1, 64); bench(2, 32); bench(2, 64); bench(4, 16); bench(4, 32)
