"""Utterance -> GPU partitioning for multi-GPU runs (replicas, no collective).

Same rule as the reference server's slot -> device map, slot i -> backend_pool[i % n_devices]
(/root/reference/src/tts-mio-server.cpp:2659-2670): utterance i goes to rank i mod world_size.  Every rank holds a
full copy of the weights; nothing crosses ranks on the data path.
"""


def shard_indices(n_items, rank, world_size):
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank/world_size")
    return list(range(rank, n_items, world_size))


def shard_list(items, rank, world_size):
    return [items[i] for i in shard_indices(len(items), rank, world_size)]


def unshard(per_rank_lists, n_items):
    """Inverse of shard_list over all ranks: restores the original utterance order."""
    world = len(per_rank_lists)
    out = [None] * n_items
    for r, lst in enumerate(per_rank_lists):
        for j, v in zip(shard_indices(n_items, r, world), lst):
            out[j] = v
    return out
