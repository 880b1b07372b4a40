"""Multi-GPU layout of a hub batch: contiguous shards, no per-step communication.

Hubs share nothing (each reference env is an isolated process; only read-only tables are common:
CHS.hpp:23, Aggregator_Simple.py:12-15, renewable.py:12-18), so rank g of G steps the global hub
range [lo, hi) on its own GPU.  Per-hub random streams are a function of the GLOBAL hub id
(include/chub_rng.h), which makes results independent of G.  The only exchange is one gather of
the episode-statistics buffer per rollout (torch.distributed: NCCL on GPUs, gloo in CPU tests).
"""
import numpy as np


def shard_range(n_total, rank, world):
    """Contiguous [lo, hi) of rank; the first n_total % world ranks get one extra hub."""
    if not (0 <= rank < world) or n_total < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def order_by_class(hub_class):
    """Permutation that lists the hubs class by class (stable).  Hubs are independent, so a caller is free to CREATE its
    batch in this order: a warp of the step kernel then holds 32 hubs of one class -- the same pile counts (no lane waits
    for a neighbour's longer station), one parameter record, one electrolyser table -- instead of 32 different ones.
    Returns (order, inverse): hub_class[order] is sorted; results[inverse] puts per-hub results back in the old order."""
    hc = np.asarray(hub_class)
    order = np.argsort(hc, kind="stable")
    inverse = np.empty_like(order)
    inverse[order] = np.arange(len(order))
    return order, inverse


def order_by_size(classes, hub_class):
    """Like order_by_class, but the classes themselves are listed by size (piles of both stations, then of the first):
    neighbouring warps -- the 16 to 28 tiles a block of the step kernel steps together -- then hold hubs of similar
    cost, so no warp waits at the unit's end for a neighbour with four times its piles, the warps of a block stay in the
    same part of the loop body, and the cost-ordered unit lists (chub_step_kernel) see units of one cost each.
    `classes` are the constructor kwargs (station_list).  Returns (order, inverse) like order_by_class."""
    hc = np.asarray(hub_class)
    size = np.array([[sum(c["station_list"]), c["station_list"][0]] for c in classes], dtype=np.int64)
    key = (size[hc, 0] * 4096 + size[hc, 1]) * (len(classes) + 1) + hc
    order = np.argsort(key, kind="stable")
    inverse = np.empty_like(order)
    inverse[order] = np.arange(len(order))
    return order, inverse


def hub_streams(base_seed, hub_ids):
    """(key0, key1, minstd state) of each global hub id -- Python mirror of chub_hub_streams()."""
    M = (1 << 64) - 1

    def mix(z):
        z = (z + 0x9E3779B97F4A7C15) & M
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    k0, k1, e0 = [], [], []
    for h in hub_ids:
        s = mix((int(base_seed) ^ ((int(h) * 0xD1342543DE82EF95) & M)) & M)
        r = mix(s)
        k0.append(s & 0xFFFFFFFF)
        k1.append(s >> 32)
        e0.append(1 + (r % 2147483646))
    return np.array(k0, np.uint32), np.array(k1, np.uint32), np.array(e0, np.uint32)


def nccl_comm_ptr(device=None):
    """ncclComm_t of torch's default NCCL process group as an int (what include/chub.h's collectives take), or None
    when the group is not NCCL.  The communicator is created lazily by torch: one collective must have run on it."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_backend() != "nccl":
        return None
    dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
    backend = dist.distributed_c10d._get_default_group()._get_backend(dev)
    try:
        return int(backend._comm_ptr())
    except Exception:
        return None


def gather_engine_stats(engine, n_total=None, dst=0, out=None, wait=True):
    """The path's one collective through the C ABI (chub_gather_stats): every rank's last-episode statistics on `dst`,
    [n_total, 4] in global hub order.  NCCL groups use the engine's own communication stream (the next steps overlap the
    transfer; wait=False leaves the join to engine.gather_wait()); any other group falls back to gather_stats()."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return engine.tensor("last_stats")
    comm = nccl_comm_ptr(engine.device)
    world, rank = dist.get_world_size(), dist.get_rank()
    if comm is None:
        return gather_stats(engine.tensor("last_stats"), n_total if n_total is not None else engine.n_hubs * world, dst)
    per = None
    if n_total is not None:
        per = [shard_range(n_total, r, world)[1] - shard_range(n_total, r, world)[0] for r in range(world)]
    total = sum(per) if per is not None else engine.n_hubs * world
    if rank == dst and out is None:
        out = torch.empty((total, 4), dtype=torch.float32, device="cuda:%d" % engine.device)
    engine.gather_stats(comm, dst, per, out if rank == dst else None)
    if wait:
        engine.gather_wait()
    return out if rank == dst else None


def gather_stats(local_stats, n_total, dst=0):
    """Gather per-hub episode statistics [n_local, K] of every rank on `dst`; returns [n_total, K] there
    (None elsewhere).  Shards may differ in length by one hub, so rows are padded to the longest shard."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local_stats
    world, rank = dist.get_world_size(), dist.get_rank()
    longest = shard_range(n_total, 0, world)[1]
    pad = torch.zeros((longest,) + tuple(local_stats.shape[1:]), dtype=local_stats.dtype, device=local_stats.device)
    pad[: local_stats.shape[0]] = local_stats
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst)
    if rank != dst:
        return None
    parts = []
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        parts.append(bufs[r][: hi - lo])
    return torch.cat(parts, 0)


def reduce_stats(local_stats):
    """Sum of the statistics columns over all hubs of all ranks (the 32-byte variant of the gather)."""
    import torch.distributed as dist
    s = local_stats.sum(0)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(s)
    return s
