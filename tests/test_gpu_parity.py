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
    assert sum(stats.values()) == 0, stats     # fp32 state bit-identical (compare_state already insists on it)


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
            assert C.close(rew.cpu().numpy(), want_rew, atol=1e-7).all()
            C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())  # t=0 form: state + obs only
            if want_done.all():
                last = eng.tensor("last_stats").cpu().numpy()
                assert C.close(last[:, 0], ret, rtol=1e-6).all()
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
    orc.reset(pv[0], wd[0], price[0], mask=mask)   # a hub reset mid-day restarts at slot 0 of the current day (chub_set_traces)
    mm = C.Mismatch()
    torch.cuda.synchronize()
    C.compare_state(eng, orc, 0, mm, C.SCALE, obs=eng.tensor("obs").cpu().numpy())
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
            assert C.close(rew.cpu().numpy(), want_rew, atol=1e-7).all(), t
            C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())
            assert not mm, "t=%d %s" % (t, mm)
    assert eng.counters()["error_or"] == int(np.bitwise_or.reduce(orc.get("hub_err")))


def test_episode_statistics_final_soc_and_penalty_match_the_oracle(tables):
    """The payload of the path's one collective and the trainer adapters' end-of-episode penalty: ALL FOUR
    statistics columns of the last finished episode -- return, income, grid energy, incomplete departures
    (evcssp_manager.py:259-269, CHS.hpp:916-921 / 1081-1086) -- and the terminal store SOC, over three auto-reset
    episodes of heterogeneous hubs, against the oracle's accumulators (pinned on the live reference's
    acumulate_reward / cumulated_income / cumulated_draw_ele / test_penalty and on its "charge is not complete" prints,
    tests/test_oracle_vs_reference.py)."""
    import torch
    rs = np.random.RandomState(77)
    classes = _hetero_classes(rs, 12)
    H, T = 512, 3 * 96
    hub_class = rs.randint(0, len(classes), H).astype(np.int32)
    eng, orc, (pv, wd, price) = _pair(H, classes, hub_class, tables, 4321, False, T + 1, rs, auto_reset=True, stride=66)
    eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    cap_kg = np.array([0.089 * 200 * classes[k]["hydro_store_vlt"] for k in hub_class])   # hydro_sys.py:94-98, / 1000
    init = np.array([classes[k]["init_soc"] for k in hub_class])
    episodes = 0
    for t in range(T):
        a = rs.uniform(-1, 1, (H, 66)).astype(np.float32)
        obs, rew, done = eng.step(torch.from_numpy(a).cuda())
        orc.step(a, pv[t + 1], wd[t + 1], price[t + 1])
        if orc.get("hub_done").all():
            torch.cuda.synchronize()
            last = eng.tensor("last_stats").cpu().numpy()
            want = np.stack([orc.get("ep_return"), orc.get("cum_income"), orc.get("cum_draw"),
                             orc.get("st_incomplete").sum(axis=1).astype(np.float64)], axis=1)
            # fp64 accumulators on both sides, narrowed once: the fp32 payload must be the oracle's value to fp32 accuracy
            assert np.array_equal(last[:, 3], want[:, 3]), "incomplete departures"
            for col, name in enumerate(("return", "income", "grid energy")):
                assert C.close(last[:, col], want[:, col], rtol=1e-6).all(), (name, t)
            fin = eng.tensor("final_soc").cpu().numpy()
            soc = orc.get("store_soc")
            assert np.array_equal(fin, soc.astype(np.float32)) or C.close(fin, soc, rtol=1e-6).all(), "final_soc"
            # the penalty VectorHubEnv forms from final_soc is the reference's test_penalty (evcssp_manager.py:275-290)
            pen = np.abs((fin.astype(np.float64) - init) * cap_kg / 0.2)
            assert C.close(pen, orc.get("test_penalty"), rtol=1e-5, atol=1e-5 * cap_kg / 0.2 * 1e-2).all(), "test_penalty"
            assert (eng.tensor("stats").cpu().numpy() == 0).all()        # running accumulators restart
            orc.reset(pv[t + 1], wd[t + 1], price[t + 1])
            episodes += 1
        elif t % 96 == 50:   # mid-episode: the running accumulators, fp64 on both sides
            torch.cuda.synchronize()
            run = eng.tensor("stats").cpu().numpy()
            assert run.dtype == np.float64
            assert C.close(run[:, 0], orc.get("ep_return"), rtol=1e-9, atol=1e-9).all()
            assert C.close(run[:, 1], orc.get("cum_income"), rtol=1e-9, atol=1e-9).all()
            assert C.close(run[:, 2], orc.get("cum_draw"), rtol=1e-9, atol=1e-9).all()
    assert episodes == 3


def test_no_reset_rollout_wraps_time_like_the_reference(tables):
    """Config 4's regime in small (evcssp_manager.py:271-273, 299; CHS.hpp:1204; hydro_sys.py:192-193): no reset at all --
    `done` is raised every 96 steps while the time slot wraps and stations, store and queues simply carry on -- over 32
    days = 3,072 steps, shared traces on a 4-day ring that the step index wraps eight times, the engine driven by
    CUDA-graph segments; two oracle windows follow step by step, integers bit-exact at every check."""
    import torch
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    seed, H, DAYS, RING, SEG, W = 0xC0FFEE, 4096, 32, 4 * 96, 48, 48
    t = np.arange(RING, dtype=np.float64)
    pv = 210.0 * np.maximum(0.0, np.sin(np.pi * ((t % 96) - 24) / 48.0)) * (1 + 0.3 * np.cos(2 * np.pi * (t // 96) / 4.0))
    wd = np.clip(46 + 40 * np.sin(2 * np.pi * t / RING * 3) + 10 * np.sin(2 * np.pi * t / 32.0), 0, 98)
    price = tables["price"][(t.astype(np.int64) - 1) % 96] * (1 + 0.1 * np.sin(2 * np.pi * (t // 96) / 4.0))
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, diagnostics=True, tables=tables, seed=seed, auto_reset=False)
    eng.set_traces(pv, wd, price)
    bases = [0, H - W]
    orcs = []
    for b in bases:
        o = O.Oracle(W, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
        o.seed_base(seed, b)
        o.reset(pv[0], wd[0], price[0])
        orcs.append(o)
    rs = np.random.RandomState(9)
    pool = rs.uniform(-1, 1, (SEG, H, 47)).astype(np.float32)
    pool_dev = torch.from_numpy(pool).cuda()
    eng.reset()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(g, stream=side):
        for j in range(SEG):
            eng.step(pool_dev[j])
    steps, dones = 0, 0
    for seg in range(DAYS * 96 // SEG):
        g.replay()
        for j in range(SEG):
            k = steps + j + 1
            for b, o in zip(bases, orcs):
                o.step(pool[j, b:b + W], pv[k % RING], wd[k % RING], price[k % RING])
            dones += int(orcs[0].get("hub_done")[0])
        steps += SEG
        torch.cuda.synchronize()
        deep = seg % 8 == 0 or steps == DAYS * 96
        obs = eng.tensor("obs").cpu().numpy()
        done = eng.tensor("done").cpu().numpy()
        for name in (C.INT_FIELDS if deep else ["st_line", "hub_uctr", "hub_lcg", "hub_t"]):
            a = eng.get(name)
            for b, o in zip(bases, orcs):
                x, y = a[b:b + W], o.get(name)
                if name in ("hub_uctr", "hub_lcg"):
                    x, y = x.astype(np.uint32), y.astype(np.uint32)
                assert np.array_equal(x, y), (name, steps, b)
        for b, o in zip(bases, orcs):
            assert C.close(obs[b:b + W], o.get("obs"), atol=1e-7).all(), ("obs", steps, b)
            assert np.array_equal(done[b:b + W].astype(np.int32), o.get("hub_done")), ("done", steps, b)
        if deep:
            for name in C.PILE_FLOATS:
                a = eng.get(name)
                for b, o in zip(bases, orcs):
                    y = o.get(name)
                    if name == "pile_power":
                        y = y * o.get("pile_car")
                    assert np.array_equal(a[b:b + W], y), (name, steps, b)     # fp32 state: bit-identical
            a = eng.get("store")
            for b, o in zip(bases, orcs):
                assert C.close(a[b:b + W], o.get("store")).all(), ("store", steps, b)
                soc = o.get("store_soc")
                assert (soc >= 0.1 - 1e-12).all() and (soc <= 1.0 + 1e-12).all()    # the store never leaves its bounds
    assert dones == DAYS and eng.counters()["step_index"] == DAYS * 96 and eng.counters()["error_or"] == 0


def test_traces_wrap_with_the_time_of_day_intact(tables):
    """A rollout longer than the trace wraps onto slot 0 (include/chub.h chub_set_traces): with a ring of whole days the
    slot read stays congruent to the hub's time of day.  The trace holds its own slot index, so obs[11] / obs[12] and
    the price entry reveal which slot was read."""
    import torch
    from charginghub_env_b200 import HubEngine
    H, RING = 64, 2 * 96
    slot = np.arange(RING, dtype=np.float64)
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, tables=tables, seed=3, auto_reset=True)
    eng.set_traces(slot * 210.0 / RING, slot * 92.0 / RING, slot)
    eng.reset()
    a = torch.zeros((H, 47), device="cuda")
    for k in range(1, 3 * RING + 5):
        obs, _, _ = eng.step(a)
        if k % 37 == 0 or k in (RING - 1, RING, RING + 1, 2 * RING, 3 * RING):
            o = obs.cpu().numpy()
            read = np.rint(o[:, 11] * RING)                    # pv slot
            assert (read == k % RING).all(), (k, read[:4])
            assert (np.rint(o[:, 12] * RING) == k % RING).all()
            t = eng.get("hub_t")
            assert (read % 96 == t).all(), (k, t[:4])


def test_masked_reset_keeps_the_last_reward_and_the_episode_statistics(tables):
    """TimeLimit-style truncation: a masked reset must not disturb what the truncating step returned, the cut
    episode's statistics become last_stats, and a hub reset in the middle of a day reads the trace slots of ITS time of
    day from then on."""
    import torch
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    rs = np.random.RandomState(12)
    H, T0 = 128, 30
    pv, wd, price = C.synth_traces(tables, 192, rs)
    engs = [HubEngine(H, [C.DEFAULT_KW], device=0, diagnostics=True, tables=tables, seed=21) for _ in range(2)]
    orc = O.Oracle(H, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    orc.seed_base(21, 0)
    for e in engs:
        e.set_traces(pv, wd, price)
        e.reset()
    orc.reset(pv[0], wd[0], price[0])
    acts = [rs.uniform(-1, 1, (H, 47)).astype(np.float32) for _ in range(T0 + 20)]
    for t in range(T0):
        for e in engs:
            obs, rew, done = e.step(torch.from_numpy(acts[t]).cuda())
        orc.step(acts[t], pv[t + 1], wd[t + 1], price[t + 1])
    mask = (rs.rand(H) < 0.5).astype(np.uint8)
    want_rew = engs[1].tensor("reward").clone()
    want_stats = engs[1].tensor("stats").clone()
    engs[0].reset(mask=torch.from_numpy(mask).cuda(), step_index=-1)
    torch.cuda.synchronize()
    assert torch.equal(engs[0].tensor("reward"), want_rew)                 # untouched, masked or not
    m = torch.from_numpy(mask).cuda().bool()
    assert torch.equal(engs[0].tensor("last_stats")[m].double(), want_stats[m].float().double())
    assert (engs[0].tensor("stats")[m] == 0).all() and torch.equal(engs[0].tensor("stats")[~m], want_stats[~m])
    # the oracle: masked hubs restart their day -> they see slot 0 of the current day, then 1, 2, ...
    orc.reset(pv[0], wd[0], price[0], mask=mask)
    mm = C.Mismatch()
    C.compare_state(engs[0], orc, 0, mm, C.SCALE, obs=engs[0].tensor("obs").cpu().numpy())
    for j in range(20):
        t = T0 + j
        obs, rew, done = engs[0].step(torch.from_numpy(acts[t]).cuda())
        sel = lambda x: np.where(mask.astype(bool), x[j + 1], x[t + 1])
        orc.step(acts[t], sel(pv), sel(wd), sel(price))
        torch.cuda.synchronize()
        C.compare_state(engs[0], orc, t + 1, mm, C.SCALE, obs=obs.cpu().numpy(), reward=rew.cpu().numpy())
    assert not mm, str(mm)
