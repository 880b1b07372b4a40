"""CPU, world_size 2 over gloo: the multi-GPU plumbing (shard ranges, global-id seeding, the one
statistics gather per rollout) with the CPU restatement standing in for the per-rank stepping."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from charginghub_env_b200 import load_tables, sharding
    from oracle import oracle as O
    from tests import common as C
    tabs = load_tables()
    lo, hi = sharding.shard_range(n_total, rank, world)
    orc = O.Oracle(hi - lo, [C.DEFAULT_KW], tables=C.oracle_tables(tabs))
    orc.seed_base(99, lo)  # streams keyed by GLOBAL hub id
    rs = np.random.RandomState(0)
    pv, wd, pr = C.synth_traces(tabs, 13, rs)
    acts = rs.uniform(-1, 1, (n_total, 47)).astype(np.float32)
    orc.reset(pv[0], wd[0], pr[0])
    for t in range(12):
        orc.step(acts[lo:hi], pv[t + 1], wd[t + 1], pr[t + 1])
    stats = torch.from_numpy(np.stack([orc.get("ep_return"), orc.get("cum_income"), orc.get("cum_draw"),
                                       orc.get("st_incomplete").sum(1).astype(np.float64)], 1).astype(np.float32))
    full = sharding.gather_stats(stats, n_total, dst=0)
    tot = sharding.reduce_stats(stats)
    if rank == 0:
        np.save(os.path.join(out_dir, "gathered.npy"), full.numpy())
        np.save(os.path.join(out_dir, "summed.npy"), tot.numpy())
    else:
        assert full is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_equal_single_rank(tmp_path, tables):
    n_total = 11  # odd on purpose: shards of 6 and 5 hubs
    mp.spawn(_worker, args=(2, _free_port(), n_total, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "gathered.npy")
    from oracle import oracle as O
    from tests import common as C
    orc = O.Oracle(n_total, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    orc.seed_base(99, 0)
    rs = np.random.RandomState(0)
    pv, wd, pr = C.synth_traces(tables, 13, rs)
    acts = rs.uniform(-1, 1, (n_total, 47)).astype(np.float32)
    orc.reset(pv[0], wd[0], pr[0])
    for t in range(12):
        orc.step(acts, pv[t + 1], wd[t + 1], pr[t + 1])
    want = np.stack([orc.get("ep_return"), orc.get("cum_income"), orc.get("cum_draw"),
                     orc.get("st_incomplete").sum(1).astype(np.float64)], 1).astype(np.float32)
    assert got.shape == (n_total, 4)
    assert np.array_equal(got, want)  # sharding changes nothing, hub for hub
    assert np.allclose(np.load(tmp_path / "summed.npy"), want.sum(0), rtol=1e-5)
