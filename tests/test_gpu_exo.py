"""-m gpu: the on-device exogenous generator (chub_set_exogenous_generator) against the generator oracle
(oracle/exo_oracle.c) feeding the hub oracle, over several auto-reset episodes, plus a checkpoint round trip."""
import numpy as np
import pytest

from tests import common as C

pytestmark = pytest.mark.gpu


def _setup(H, tables, seed, rf, pf):
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, diagnostics=True, tables=tables, seed=seed, auto_reset=True)
    eng.set_exogenous_generator(tables, rf, pf)
    orc = O.Oracle(H, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    orc.seed_base(seed, 0)
    exo = O.ExoOracle(H, tables, rf, pf)
    exo.seed_base(seed, 0)
    return eng, orc, exo


def test_device_generator_three_episodes(tables):
    import torch
    rs = np.random.RandomState(21)
    H, T = 200, 3 * 96
    eng, orc, exo = _setup(H, tables, 99, 0.25, 0.1)
    obs = eng.reset()
    orc.reset(*exo.reset())
    mm = C.Mismatch()
    torch.cuda.synchronize()
    C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())
    for t in range(T):
        a = rs.uniform(-1, 1, (H, 47)).astype(np.float32)
        obs, rew, done = eng.step(torch.from_numpy(a).cuda())
        orc.step(a, *exo.step(orc.get("hub_t")))
        want_rew, want_done = orc.get("reward"), orc.get("hub_done")
        if want_done.all():
            orc.reset(*exo.reset())  # what a user of the reference does after done
        if (t + 1) % 8 == 0 or want_done.any():
            torch.cuda.synchronize()
            assert np.array_equal(done.cpu().numpy().astype(np.int32), want_done)
            assert C.close(rew.cpu().numpy(), want_rew, atol=1e-7).all()
            C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())
    assert not mm, str(mm)
    # the generator consumed exactly the oracle's share of the X stream, and the noise is really there
    assert eng.counters()["error_or"] == 0
    pv_now = obs.cpu().numpy()[:, 11]
    assert np.unique(np.round(pv_now, 6)).size > 1 or np.unique(np.round(obs.cpu().numpy()[:, 12], 6)).size > H // 2


def test_generator_state_travels_with_the_checkpoint(tables):
    import torch
    rs = np.random.RandomState(5)
    H = 96
    eng, _, _ = _setup(H, tables, 7, 0.0, 0.0)
    eng.reset()
    acts = [torch.from_numpy(rs.uniform(-1, 1, (H, 47)).astype(np.float32)).cuda() for _ in range(20)]
    for a in acts[:10]:
        eng.step(a)
    blob = eng.get_state()
    outs = []
    for a in acts[10:]:
        o, r, _ = eng.step(a)
        outs.append((o.cpu().numpy().copy(), r.cpu().numpy().copy()))
    eng.set_state(blob)
    for a, (o0, r0) in zip(acts[10:], outs):
        o, r, _ = eng.step(a)
        assert np.array_equal(o.cpu().numpy(), o0) and np.array_equal(r.cpu().numpy(), r0)


def test_batched_env_with_device_generator(tables):
    """BatchedHubEnv(traces="device"): hubs see different days and noise; statistics of a day are sane."""
    import torch
    from charginghub_env_b200.env import BatchedHubEnv
    H = 512
    env = BatchedHubEnv(H, env_kwargs=dict(C.DEFAULT_KW, renew_fluctuate=0.1, price_fluctuate=0.1), seed=3,
                        traces="device")
    obs = env.reset()
    pv_cols, wd_cols = [], []
    for t in range(96):
        a = torch.rand((H, env.action_dim), device="cuda") * 2 - 1
        obs, rew, done, info = env.step(a)
        pv_cols.append(obs[:, 11].clone())
        wd_cols.append(obs[:, 12].clone())
    torch.cuda.synchronize()
    assert bool(done.all()) and info == {}
    pv = torch.stack(pv_cols).cpu().numpy() * 210.0   # back to kW (evcssp_manager.py:338)
    wd = torch.stack(wd_cols).cpu().numpy() * 92.0
    assert pv.min() >= 0.0 and wd.min() >= 0.0
    from charginghub_env_b200.exogenous import SCALE_PV, SCALE_WD
    # OU noise (theta 0.01, sigma 1 / 1.5, gain 1.1) wanders by roughly sigma * sqrt(96) over the first day
    assert pv.max() <= SCALE_PV * (tables["pv_days"].max() + 60) and wd.max() <= SCALE_WD * (tables["wd_days"].max() + 90)
    assert np.unique(np.round(wd[10], 4)).size > H // 4      # per-hub days and noise: hubs differ
    assert abs(wd.mean() - SCALE_WD * tables["wd_days"].mean()) < 6.0 * SCALE_WD   # and average to the data's mean
    env.close()
