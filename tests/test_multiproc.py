"""N>1 host logic on CPU: utterance -> rank partitioning (replicas, no data-path collective) and the bench's
max-over-ranks / sum-over-ranks aggregation, world_size 2 over gloo."""
import os
import socket

import numpy as np
import pytest

from mio_tts_cpp_b200.shard import shard_indices, shard_list, unshard


def test_shard_rule_matches_reference_round_robin():
    # slot i -> backend_pool[i % n_devices], /root/reference/src/tts-mio-server.cpp:2659-2670
    assert shard_indices(10, 0, 4) == [0, 4, 8] and shard_indices(10, 3, 4) == [3, 7]
    assert shard_indices(2, 3, 4) == [] and shard_indices(0, 0, 1) == []
    items = list("abcdefghij")
    parts = [shard_list(items, r, 3) for r in range(3)]
    assert sorted(sum(parts, [])) == items and unshard(parts, len(items)) == items
    with pytest.raises(ValueError):
        shard_indices(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_items, out_dir):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lens = [100 + 37 * i for i in range(n_items)]                 # codes per utterance
    mine = shard_list(lens, rank, world)
    audio_s = sum(n * 960 / 24000.0 for n in mine)                # what this replica would synthesise
    step_ms = 10.0 + rank                                          # pretend rank 1 is slower
    t = torch.tensor([audio_s], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    m = torch.tensor([step_ms], dtype=torch.float64)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    gathered = [None] * world
    dist.all_gather_object(gathered, [n * 960 for n in mine])     # per-utterance sample counts
    if rank == 0:
        np.save(os.path.join(out_dir, "res.npy"), np.array([t.item(), m.item()] + unshard(gathered, n_items)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_replicas_aggregate(tmp_path):
    import torch.multiprocessing as mp

    n_items, world = 7, 2
    mp.spawn(_worker, args=(world, _free_port(), n_items, str(tmp_path)), nprocs=world, join=True)
    res = np.load(tmp_path / "res.npy")
    lens = [100 + 37 * i for i in range(n_items)]
    assert np.isclose(res[0], sum(n * 960 / 24000.0 for n in lens))      # whole-job audio seconds
    assert res[1] == 11.0                                               # time = max over ranks
    assert res[2:].tolist() == [n * 960 for n in lens]                  # original utterance order restored
