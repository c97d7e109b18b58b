"""N>1 host logic on CPU, world_size 2 over gloo: what bench.py does under torchrun besides launching kernels -- each
rank builds ITS shard of the synthetic workload (benchutil.config2_workload(rank, ...): disjoint seeds, weak scaling),
and the run is aggregated as time = MAX over ranks, work = SUM over ranks (benchutil.reduce_over_ranks).  There is no
data-path collective to test: utterances are independent."""
import os
import socket

import numpy as np

from mio_tts_cpp_b200 import benchutil as U


def test_rank_shards_are_disjoint_and_reproducible():
    a0, a1 = U.config2_workload(0, 4, 50), U.config2_workload(1, 4, 50)
    assert all(x.dtype == np.int32 and x.size == 50 and 0 <= x.min() and x.max() < 12800 for x in a0 + a1)
    assert not any(np.array_equal(x, y) for x in a0 for y in a1)                 # disjoint seeds per rank
    assert all(np.array_equal(x, y) for x, y in zip(a0, U.config2_workload(0, 4, 50)))
    # rank 0 of the multi-GPU run decodes the very batch the single-GPU run decodes (BASELINE config 2, seeds 42..105)
    assert np.array_equal(U.config2_workload(0, 64, 400)[17], np.random.default_rng(59).integers(0, 12800, 400).astype(np.int32))


def test_single_process_reduce_is_identity():
    assert U.reduce_over_ranks(None, "cpu", 8.5, 1024.0) == (8.5, 1024.0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    codes = U.config2_workload(rank, 3 + rank, 100)                 # ragged on purpose: rank 1 has one utterance more
    audio_s = sum(c.size * 960 / 24000.0 for c in codes)
    ms, total = U.reduce_over_ranks(dist, "cpu", 10.0 + rank, audio_s)      # pretend rank 1 is slower
    if rank == 0:
        np.save(os.path.join(out_dir, "res.npy"), np.array([ms, total]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_replicas_aggregate(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    ms, total = np.load(tmp_path / "res.npy")
    assert ms == 11.0                                               # time = max over ranks
    assert np.isclose(total, (3 + 4) * 100 * 960 / 24000.0)         # whole-job audio seconds = sum over ranks
