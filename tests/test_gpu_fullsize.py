"""-m gpu: parity at BASELINE's full per-GPU size (the config-5 shard, 524,288 hubs; here minus 13 so that the last
warp is ragged).  The oracle cannot step half a million hubs in test time, so it steps windows of hubs spread over the
whole range -- each seeded from its GLOBAL hub ids, which is all a hub's trajectory depends on -- and the engine must
agree with it there; plus two size-independent properties: the launch shape does not change a single bit, and a
second engine holding only the upper half of the hubs (hub_id_base) reproduces that half."""
import os

import numpy as np
import pytest

from tests import common as C

pytestmark = pytest.mark.gpu

H_FULL = 524288 - 13
WINDOW = 128


def _year(tabs, L):
    t = np.arange(L, dtype=np.float64)
    pv = 210.0 * np.maximum(0.0, np.sin(np.pi * ((t % 96) - 24) / 48.0))
    wd = np.clip(46 + 40 * np.sin(2 * np.pi * t / 355.0), 0, 98)
    price = tabs["price"][(t.astype(np.int64) - 1) % 96]
    return pv, wd, price


def _engine(H, tabs, seed, base=0):
    from charginghub_env_b200 import HubEngine
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, hub_id_base=base, diagnostics=True, tables=tabs, seed=seed)
    eng.set_traces(*_year(tabs, 98))
    return eng


def test_windows_of_the_full_shard_match_the_oracle(tables):
    import torch
    from oracle import oracle as O
    seed = 0x5EED
    eng = _engine(H_FULL, tables, seed)
    pv, wd, price = _year(tables, 98)
    bases = [0, 1000, 77777, H_FULL // 2 - 50, 300001, 444444, H_FULL - 3 * WINDOW, H_FULL - WINDOW]
    orcs = []
    for b in bases:
        o = O.Oracle(WINDOW, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
        o.seed_base(seed, b)
        o.reset(pv[0], wd[0], price[0])
        orcs.append(o)
    obs = eng.reset()
    gen = torch.Generator(device="cuda").manual_seed(5)
    S = eng.action_stride

    def compare(t, obs, rew, done, deep):
        torch.cuda.synchronize()
        for b, o in zip(bases, orcs):
            sl = slice(b, b + WINDOW)
            assert C.close(obs[sl].cpu().numpy(), o.get("obs"), atol=1e-7).all(), ("obs", t, b)
            if t:
                assert C.close(rew[sl].cpu().numpy(), o.get("reward"), atol=1e-7).all(), ("reward", t, b)
                assert np.array_equal(done[sl].cpu().numpy().astype(np.int32), o.get("hub_done")), ("done", t, b)
        if not deep:
            return
        for name in C.INT_FIELDS:
            a = eng.get(name)
            for b, o in zip(bases, orcs):
                x, y = a[b:b + WINDOW], o.get(name)
                if name in ("hub_uctr", "hub_lcg"):
                    x, y = x.astype(np.uint32), y.astype(np.uint32)
                assert np.array_equal(x, y), (name, t, b)
        for name in C.PILE_FLOATS + ["store"]:
            a = eng.get(name)
            for b, o in zip(bases, orcs):
                y = o.get(name)
                if name == "pile_power":
                    y = y * o.get("pile_car")
                if name == "store":
                    assert C.close(a[b:b + WINDOW], y).all(), (name, t, b)
                else:
                    assert np.array_equal(a[b:b + WINDOW], y), (name, t, b)      # fp32 state: bit-identical

    compare(0, obs, None, None, True)
    for t in range(96):
        a = torch.rand((H_FULL, S), device="cuda", generator=gen) * 2 - 1
        obs, rew, done = eng.step(a)
        for b, o in zip(bases, orcs):
            o.step(a[b:b + WINDOW].cpu().numpy(), pv[t + 1], wd[t + 1], price[t + 1])
        if t % 8 == 7 or t == 0:
            compare(t + 1, obs, rew, done, t in (0, 47, 95))
    assert eng.counters()["error_or"] == 0


def _digest(eng):
    """Order-sensitive checksum of everything a hub's trajectory is made of (integers and float bit patterns)."""
    import torch
    parts = [eng.tensor("obs").view(torch.int32).long().sum(dim=1), eng.tensor("reward").view(torch.int32).long(),
             eng.tensor("stats").view(torch.int32).long().sum(dim=1)]
    cols, _ = eng.situation()
    for name in ("car", "charge", "power", "soc", "target_soc", "init_soc", "emergency"):
        parts.append(cols[name].view(torch.int32).long().sum(dim=1))
    return torch.stack(parts, dim=1)


def _rollout(eng, steps, seed, H, offset=0):
    import torch
    gen = torch.Generator(device="cuda").manual_seed(seed)
    eng.reset()
    for t in range(steps):
        a = torch.rand((H, eng.action_stride), device="cuda", generator=gen) * 2 - 1
        eng.step(a[offset:offset + eng.n_hubs].contiguous())
    torch.cuda.synchronize()
    return _digest(eng)


def test_launch_shape_and_sharding_do_not_change_a_bit(tables):
    import torch
    seed, steps = 99, 40      # through the morning ramp into the busy hours
    ref = _rollout(_engine(H_FULL, tables, seed), steps, 7, H_FULL)
    old = os.environ.get("CHUB_BLOCK")
    try:
        for shape in ("1610", "810"):      # 16 and 8 warps per block instead of the fitted 28
            os.environ["CHUB_BLOCK"] = shape
            got = _rollout(_engine(H_FULL, tables, seed), steps, 7, H_FULL)
            assert torch.equal(ref, got), shape
    finally:
        if old is None:
            os.environ.pop("CHUB_BLOCK", None)
        else:
            os.environ["CHUB_BLOCK"] = old
    half = H_FULL // 2 + 5                 # an unaligned split: the upper part as its own engine (another "rank")
    upper = _rollout(_engine(H_FULL - half, tables, seed, base=half), steps, 7, H_FULL, offset=half)
    assert torch.equal(ref[half:], upper)


def test_long_horizon_rollout_matches_the_oracle(tables):
    """Config 4 in small: 21 simulated days (2,016 steps) of auto-reset episodes on 20,000 hubs, the shared traces a ring of
    7 days that the absolute step index wraps around three times, actions from a fixed pool, the engine driven by
    CUDA-graph segments of 32 steps; the oracle follows two windows of hubs step by step.  Integers must agree bit for
    bit at every check, floats within RTOL -- a drift would compound over two thousand steps."""
    import torch
    from oracle import oracle as O
    from charginghub_env_b200 import HubEngine
    seed, H, DAYS, RING, SEG = 0xABCDE, 20000, 21, 7 * 96, 32
    t = np.arange(RING, dtype=np.float64)
    pv = 210.0 * np.maximum(0.0, np.sin(np.pi * ((t % 96) - 24) / 48.0)) * (1 + 0.3 * np.cos(2 * np.pi * (t // 96) / 7.0))
    wd = np.clip(46 + 40 * np.sin(2 * np.pi * t / RING * 3) + 10 * np.sin(2 * np.pi * t / 32.0), 0, 98)
    price = tables["price"][(t.astype(np.int64) - 1) % 96] * (1 + 0.1 * np.sin(2 * np.pi * (t // 96) / 7.0))
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, diagnostics=True, tables=tables, seed=seed, auto_reset=True)
    eng.set_traces(pv, wd, price)
    bases = [0, H - 64]
    orcs = []
    for b in bases:
        o = O.Oracle(64, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
        o.seed_base(seed, b)
        o.reset(pv[0], wd[0], price[0])
        orcs.append(o)
    rs = np.random.RandomState(8)
    pool = rs.uniform(-1, 1, (SEG, H, 47)).astype(np.float32)     # step j of every segment uses pool[j]
    pool_dev = torch.from_numpy(pool).cuda()
    eng.reset()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(g, stream=side):
        for j in range(SEG):
            eng.step(pool_dev[j])
    steps = 0
    returns = [np.zeros(64), np.zeros(64)]
    for seg in range(DAYS * 96 // SEG):
        g.replay()
        for j in range(SEG):
            k = steps + j + 1                                      # absolute index of the state being entered
            for w, (b, o) in enumerate(zip(bases, orcs)):
                o.step(pool[j, b:b + 64], pv[k % RING], wd[k % RING], price[k % RING])
                if o.get("hub_done").all():
                    returns[w] = o.get("ep_return").copy()
                    o.reset(pv[k % RING], wd[k % RING], price[k % RING])
        steps += SEG
        torch.cuda.synchronize()
        deep = seg % 9 == 0 or steps == DAYS * 96
        obs = eng.tensor("obs").cpu().numpy()
        for name in (C.INT_FIELDS if deep else ["st_line", "hub_uctr", "hub_lcg", "hub_t"]):
            a = eng.get(name)
            for b, o in zip(bases, orcs):
                x, y = a[b:b + 64], o.get(name)
                if name in ("hub_uctr", "hub_lcg"):
                    x, y = x.astype(np.uint32), y.astype(np.uint32)
                assert np.array_equal(x, y), (name, steps, b)
        for b, o in zip(bases, orcs):
            assert C.close(obs[b:b + 64], o.get("obs"), atol=1e-7).all(), ("obs", steps, b)
        if deep:
            for name in C.PILE_FLOATS + ["store"]:
                a = eng.get(name)
                for b, o in zip(bases, orcs):
                    y = o.get(name)
                    if name == "pile_power":
                        y = y * o.get("pile_car")
                    if name == "store":
                        assert C.close(a[b:b + 64], y).all(), (name, steps, b)
                    else:
                        assert np.array_equal(a[b:b + 64], y), (name, steps, b)  # fp32 state: bit-identical
    assert eng.counters()["step_index"] == DAYS * 96 and eng.counters()["error_or"] == 0
    last = eng.tensor("last_stats").cpu().numpy()
    for w, b in enumerate(bases):
        assert C.close(last[b:b + 64, 0], returns[w], rtol=1e-6).all()


def test_windows_of_the_full_config3_shard_match_the_oracle(tables):
    """BASELINE config 3 at full size: 262,144 heterogeneous hubs (minus 7: ragged last warp) of 4,096 parameter classes
    (bench.py's workload: 4..32 piles per station, H2 sizes, permeate up to 0.05), action rows padded to 66, fused
    auto-reset; eight oracle windows spread over the hub range follow one and a half episodes, the oracle reset by hand
    at done.  Integers bit-exact, fp32 pile state bit-identical, fp64 within RTOL."""
    import torch
    import bench
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    wl = bench.workload("config3", tables)
    H, W, seed, S = wl["H"] - 7, 96, 0xC3, wl["stride"]
    hub_class = np.ascontiguousarray(wl["hub_class"][:H])
    eng = HubEngine(H, wl["classes"], hub_class=hub_class, device=0, diagnostics=True, tables=tables, seed=seed,
                    auto_reset=True, action_stride=S)
    T = 144
    pv, wd, price = _year(tables, T + 2)
    eng.set_traces(pv, wd, price)
    bases = [0, 4097, 65536 - 40, 100003, H // 2, 200000, H - 2 * W - 1, H - W]
    orcs = []
    for b in bases:
        o = O.Oracle(W, wl["classes"], hub_class=hub_class[b:b + W], tables=C.oracle_tables(tables))
        o.seed_base(seed, b)
        o.reset(pv[0], wd[0], price[0])
        orcs.append(o)
    obs = eng.reset()
    gen = torch.Generator(device="cuda").manual_seed(6)
    err_or = 0
    for t in range(T):
        a = torch.rand((H, S), device="cuda", generator=gen) * 2 - 1
        obs, rew, done = eng.step(a)
        rows = a.cpu().numpy()
        deep = t in (0, 40, 95, 96, T - 1)
        if deep or t % 6 == 5:
            torch.cuda.synchronize()
            o_h, r_h, d_h = obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy()
            fields = {n: eng.get(n) for n in (C.INT_FIELDS + C.PILE_FLOATS + ["store"] if deep else ["hub_uctr", "hub_lcg", "st_line"])}
        for b, o in zip(bases, orcs):
            o.step(rows[b:b + W], pv[t + 1], wd[t + 1], price[t + 1])
            want_rew, want_done = o.get("reward"), o.get("hub_done")
            if want_done.all():
                err_or |= int(np.bitwise_or.reduce(o.get("hub_err")))
                o.reset(pv[t + 1], wd[t + 1], price[t + 1])
            if not (deep or t % 6 == 5):
                continue
            sl = slice(b, b + W)
            assert np.array_equal(d_h[sl].astype(np.int32), want_done), ("done", t, b)
            assert C.close(r_h[sl], want_rew, atol=1e-7).all(), ("reward", t, b)
            assert C.close(o_h[sl], o.get("obs"), atol=1e-7).all(), ("obs", t, b)
            for name, a_all in fields.items():
                x, y = a_all[sl], o.get(name)
                if name in ("hub_uctr", "hub_lcg"):
                    x, y = x.astype(np.uint32), y.astype(np.uint32)
                if name == "pile_power":
                    y = y * o.get("pile_car")
                if name == "store":
                    assert C.close(x, y).all(), (name, t, b)
                else:
                    assert np.array_equal(x, y), (name, t, b)
    for o in orcs:
        err_or |= int(np.bitwise_or.reduce(o.get("hub_err")))
    got_err = eng.tensor("error").cpu().numpy()
    assert int(np.bitwise_or.reduce(np.concatenate([got_err[b:b + W] for b in bases]))) == err_or


@pytest.mark.parametrize("extra", [[], ["--classes", "8", "--loads-every", "3"], ["--exogenous", "device"]])
def test_parity_soak_tool_small(extra):
    """tools/parity_soak.py (the large GPU-vs-checker soak whose results are in profiles/) in small, so that it stays runnable."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "tools", "parity_soak.py"), "--hubs", "2048", "--days", "2",
                          "--threads", "4"] + extra, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "soak passed: 2048 hubs x 192 steps" in out.stdout
