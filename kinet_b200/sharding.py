"""Multi-GPU partitioning of the path: replicas only.  MSDeformAttn has no exchange step -- every
(batch item, query, head) output depends only on that batch item's value tensor -- so work is split
by independent units (video sequences for tracking inference, batches for op sweeps) with no
collective on the data path.  Mirrors the reference's round-robin assignment of MOT sequences to
ranks during evaluation (src/trackformer/engine.py:290-296: sequence i -> rank i mod world)."""


def shard_round_robin(items, rank, world):
    """Items owned by `rank` out of `world` ranks: items[rank], items[rank + world], ..."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("need 0 <= rank < world")
    return list(items)[rank::world]


def merge_round_robin(shards):
    """Inverse of shard_round_robin for a list of per-rank result lists (rank order)."""
    world = len(shards)
    total = sum(len(s) for s in shards)
    out = [None] * total
    for r, s in enumerate(shards):
        for k, v in enumerate(s):
            out[r + k * world] = v
    return out
