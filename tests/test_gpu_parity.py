"""-m gpu: the CUDA engine (through the C ABI) against the CPU oracle on identical streams and actions."""
import numpy as np
import pytest

from tests import common as C

pytestmark = pytest.mark.gpu


def _pair(H, classes, hub_class, tabs, seed, per_hub_traces, L, rs, auto_reset=False, stride=None):
    import torch  # noqa: F401
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    eng = HubEngine(H, classes, hub_class=hub_class, device=0, diagnostics=True, tables=tabs, seed=seed,
                    auto_reset=auto_reset, action_stride=stride)
    pv, wd, price = C.synth_traces(tabs, L, rs, H if per_hub_traces else None)
    eng.set_traces(pv, wd, price, per_hub=per_hub_traces)
    orc = O.Oracle(H, classes, hub_class=hub_class, tables=C.oracle_tables(tabs))
    orc.seed_base(seed, 0)
    return eng, orc, (pv, wd, price)


def _run(eng, orc, traces, T, rs, check_every=1):
    import torch
    pv, wd, price = traces
    H, S = eng.n_hubs, eng.action_stride
    mm = C.Mismatch()
    stats = {}
    obs = eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    torch.cuda.synchronize()
    C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())
    for t in range(T):
        a = rs.uniform(-1, 1, (H, S)).astype(np.float32)
        obs, rew, done = eng.step(torch.from_numpy(a).cuda())
        orc.step(a, pv[t + 1], wd[t + 1], price[t + 1])
        if (t + 1) % check_every == 0 or t == T - 1:
            torch.cuda.synchronize()
            C.compare_state(eng, orc, t + 1, mm, C.SCALE, obs=obs.cpu().numpy(), reward=rew.cpu().numpy(),
                            done=done.cpu().numpy(), float_exact_stats=stats)
    return mm, stats


def test_default_hubs_full_episode(tables):
    """Config 2 in small: identical default hubs, 96 steps, every field compared at every step."""
    rs = np.random.RandomState(1234)
    eng, orc, tr = _pair(256, [C.DEFAULT_KW], None, tables, 0x5EED, True, 97, rs)
    mm, stats = _run(eng, orc, tr, 96, rs)
    assert not mm, str(mm)
    assert eng.counters()["error_or"] == 0
    print("fp32 fields not bit-identical (of all compared):", stats)


def test_config2_full_size_bit_exact(tables):
    """BASELINE config 2 at full size: 4,096 identical default hubs, 96 steps, per-hub traces."""
    rs = np.random.RandomState(2)
    eng, orc, tr = _pair(4096, [C.DEFAULT_KW], None, tables, 0x5EED, True, 97, rs)
    mm, stats = _run(eng, orc, tr, 96, rs, check_every=8)
    assert not mm, str(mm)
    assert eng.counters()["error_or"] == 0
    # all fp32 state came out bit-identical so far; allow a handful of last-bit differences
    assert sum(stats.values()) <= 16, stats


def _hetero_classes(rs, n):
    out = []
    for _ in range(n):
        out.append(dict(station_list=[int(rs.randint(4, 33)), int(rs.randint(4, 33))],
                        station_type_list=[["fast", "slow"], ["slow", "fast"], ["fast", "fast"], ["slow", "slow"]][rs.randint(4)],
                        constant_charging=bool(rs.rand() < 0.15),
                        hydro_prod_rate=float(rs.uniform(50, 430)), hydro_store_vlt=float(rs.uniform(10, 100)),
                        init_soc=float(rs.uniform(0.2, 0.8)), fc_max_power=float(rs.choice([50, 100, 200])),
                        fcev_permeate=float(rs.choice([0.005, 0.01, 0.02, 0.05])),
                        hydro_loss=float(rs.choice([0.0, 0.02]))))
    return out


def test_heterogeneous_hubs(tables):
    """Config 3 in small: per-hub pile counts 4..32, station order/type, H2 sizes, permeate, padded actions."""
    rs = np.random.RandomState(3)
    classes = _hetero_classes(rs, 24)
    H = 768
    hub_class = rs.randint(0, len(classes), H).astype(np.int32)
    eng, orc, tr = _pair(H, classes, hub_class, tables, 31337, False, 97, rs, stride=66)
    mm, stats = _run(eng, orc, tr, 96, rs, check_every=4)
    assert not mm, str(mm)
    assert eng.counters()["error_or"] == int(np.bitwise_or.reduce(orc.get("hub_err")))


def test_auto_reset_three_episodes(tables):
    """Fused auto-reset: 3 x 96 steps; the oracle is reset by hand at every done with the same traces."""
    import torch
    rs = np.random.RandomState(4)
    H, T = 512, 3 * 96
    eng, orc, tr = _pair(H, [C.DEFAULT_KW], None, tables, 5, False, T + 1, rs, auto_reset=True)
    pv, wd, price = tr
    eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    mm = C.Mismatch()
    for t in range(T):
        a = rs.uniform(-1, 1, (H, 47)).astype(np.float32)
        obs, rew, done = eng.step(torch.from_numpy(a).cuda())
        orc.step(a, pv[t + 1], wd[t + 1], price[t + 1])
        want_rew, want_done = orc.get("reward"), orc.get("hub_done")
        ret = orc.get("ep_return")
        if want_done.all():
            orc.reset(pv[t + 1], wd[t + 1], price[t + 1])  # what a user of the reference does after done
        if (t + 1) % 16 == 0 or want_done.any():
            torch.cuda.synchronize()
            assert np.array_equal(done.cpu().numpy().astype(np.int32), want_done)
            assert C.close(rew.cpu().numpy(), want_rew, atol=C.RTOL).all()
            C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())  # t=0 form: state + obs only
            if want_done.all():
                last = eng.tensor("last_stats").cpu().numpy()
                assert np.allclose(last[:, 0], ret, rtol=1e-4, atol=1e-4)
    assert not mm, str(mm)
    assert int(eng.get("hub_t")[0]) == 0


def test_masked_reset(tables):
    import torch
    rs = np.random.RandomState(6)
    H = 256
    eng, orc, tr = _pair(H, [C.DEFAULT_KW], None, tables, 8, False, 40, rs)
    pv, wd, price = tr
    eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    for t in range(10):
        a = rs.uniform(-1, 1, (H, 47)).astype(np.float32)
        eng.step(torch.from_numpy(a).cuda())
        orc.step(a, pv[t + 1], wd[t + 1], price[t + 1])
    mask = (rs.rand(H) < 0.3).astype(np.uint8)
    eng.reset(mask=torch.from_numpy(mask).cuda(), step_index=-1)
    orc.reset(pv[10], wd[10], price[10], mask=mask)
    mm = C.Mismatch()
    torch.cuda.synchronize()
    C.compare_state(eng, orc, 0, mm, C.SCALE)
    assert not mm, str(mm)


@pytest.mark.parametrize("stride,offset,H", [(47, 0, 100), (100, 0, 96), (47, 1, 64), (49, 3, 33)])
def test_action_row_layouts(tables, stride, offset, H):
    """The warp stages its 32 action rows in shared memory when it can; a partial last warp, a matrix too wide to
    stage and a base pointer that is not 16-byte aligned take the direct path.  All must read the same actions."""
    import torch
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    rs = np.random.RandomState(11)
    T = 12
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, diagnostics=True, tables=tables, seed=77, action_stride=stride)
    pv, wd, price = C.synth_traces(tables, T + 1, rs, None)
    eng.set_traces(pv, wd, price)
    orc = O.Oracle(H, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    orc.seed_base(77, 0)
    eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    buf = torch.empty(H * stride + 8, device="cuda", dtype=torch.float32)
    mm = C.Mismatch()
    for t in range(T):
        a = rs.uniform(-1, 1, (H, stride)).astype(np.float32)
        view = buf[offset:offset + H * stride].view(H, stride)
        view.copy_(torch.from_numpy(a))
        obs, rew, done = eng.step(view)
        orc.step(a, pv[t + 1], wd[t + 1], price[t + 1])
        torch.cuda.synchronize()
        C.compare_state(eng, orc, t + 1, mm, C.SCALE, obs=obs.cpu().numpy(), reward=rew.cpu().numpy(),
                        done=done.cpu().numpy())
    assert not mm, str(mm)
    assert eng.counters()["error_or"] == 0


def test_config3_ten_days_heterogeneous_auto_reset(tables):
    """Config 3 in small: heterogeneous hubs (4..32 piles per station, all four station-type orders,
    constant-power classes, H2 sizes, permeate up to 0.05), padded action rows of 66, fused auto-reset, 10 x 96 steps with
    per-hub traces; the scalar-load step takes every fifth step.  The oracle is reset by hand at every done."""
    import torch
    rs = np.random.RandomState(33)
    classes = _hetero_classes(rs, 16)
    H, T = 384, 10 * 96
    hub_class = rs.randint(0, len(classes), H).astype(np.int32)
    eng, orc, (pv, wd, price) = _pair(H, classes, hub_class, tables, 777, True, T + 1, rs, auto_reset=True, stride=66)
    caps = np.array([[40.0 * n for n in classes[k]["station_list"]] for k in hub_class], np.float32)
    eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    mm = C.Mismatch()
    for t in range(T):
        if t % 5 == 4:
            loads = np.concatenate([rs.uniform(0, 1, (H, 2)).astype(np.float32) ** 2 * caps,
                                    rs.uniform(-1, 1, (H, 2)).astype(np.float32)], axis=1)
            obs, rew, done = eng.step_loads(torch.from_numpy(loads).cuda())
            orc.set_load_mode(True)
            orc.step(loads, pv[t + 1], wd[t + 1], price[t + 1])
        else:
            a = rs.uniform(-1, 1, (H, 66)).astype(np.float32)
            obs, rew, done = eng.step(torch.from_numpy(a).cuda())
            orc.set_load_mode(False)
            orc.step(a, pv[t + 1], wd[t + 1], price[t + 1])
        want_rew, want_done = orc.get("reward"), orc.get("hub_done")
        if want_done.all():
            orc.reset(pv[t + 1], wd[t + 1], price[t + 1])
        if (t + 1) % 24 == 0 or want_done.any():
            torch.cuda.synchronize()
            assert np.array_equal(done.cpu().numpy().astype(np.int32), want_done), t
            assert C.close(rew.cpu().numpy(), want_rew, atol=C.RTOL).all(), t
            C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())
            assert not mm, "t=%d %s" % (t, mm)
    assert eng.counters()["error_or"] == int(np.bitwise_or.reduce(orc.get("hub_err")))
