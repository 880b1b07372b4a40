"""-m gpu: the scalar-load step (chub_step_loads: the stations' evs_step(float), CHS.hpp:1169-1186 / 1478-1495)
against the CPU oracle in load mode on identical streams, targets and traces."""
import numpy as np
import pytest

from tests import common as C

pytestmark = pytest.mark.gpu


def _loads(rs, H, caps):
    rows = np.zeros((H, 4), np.float32)
    for k in range(2):
        rows[:, k] = rs.uniform(-0.1, 1.0, H) ** 2 * caps[:, k]
    rows[:, :2][rs.uniform(size=(H, 2)) < 0.1] = 1e9      # "everything you can take"
    rows[:, :2][rs.uniform(size=(H, 2)) < 0.05] = -5.0    # "nothing": clamps to min_power
    rows[:, 2:] = rs.uniform(-1, 1, (H, 2))
    return rows


def _run(H, classes, hub_class, tabs, seed, T, check_every):
    import torch
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    rs = np.random.RandomState(seed)
    eng = HubEngine(H, classes, hub_class=hub_class, device=0, diagnostics=True, tables=tabs, seed=seed)
    pv, wd, price = C.synth_traces(tabs, T + 1, rs, H)
    eng.set_traces(pv, wd, price, per_hub=True)
    orc = O.Oracle(H, classes, hub_class=hub_class, tables=C.oracle_tables(tabs))
    orc.set_load_mode(True)
    orc.seed_base(seed, 0)
    hc = np.zeros(H, np.int64) if hub_class is None else np.asarray(hub_class)
    caps = np.array([[40.0 * n for n in classes[k]["station_list"]] for k in hc])
    mm = C.Mismatch()
    obs = eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    torch.cuda.synchronize()
    C.compare_state(eng, orc, 0, mm, C.SCALE, obs=obs.cpu().numpy())
    n_on = 0
    for t in range(T):
        rows = _loads(rs, H, caps)
        obs, rew, done = eng.step_loads(torch.from_numpy(rows).cuda())
        orc.step(rows, pv[t + 1], wd[t + 1], price[t + 1])
        if (t + 1) % check_every == 0 or t == T - 1:
            torch.cuda.synchronize()
            C.compare_state(eng, orc, t + 1, mm, C.SCALE, obs=obs.cpu().numpy(), reward=rew.cpu().numpy(),
                            done=done.cpu().numpy())
            n_on += int(eng.get("pile_charge").sum())
    assert not mm, str(mm)
    assert eng.counters()["error_or"] == 0
    assert n_on > H     # the dispatch did switch piles on


def test_default_hubs_full_episode(tables):
    _run(512, [C.DEFAULT_KW], None, tables, 31, 96, 1)


def test_mixed_classes_with_constant_charging(tables):
    classes = [C.DEFAULT_KW,
               dict(C.DEFAULT_KW, station_list=[32, 5], station_type_list=["slow", "fast"], init_soc=0.45),
               dict(C.DEFAULT_KW, constant_charging=True),
               dict(C.DEFAULT_KW, station_list=[7, 9], station_type_list=["fast", "fast"], constant_charging=True)]
    H = 1000   # ragged last warp
    hub_class = np.random.RandomState(8).randint(0, len(classes), H).astype(np.int32)
    _run(H, classes, hub_class, tables, 32, 96, 4)


def test_per_pile_step_after_load_step(tables):
    """The two step kinds can be mixed freely on one engine (the on-bits buffer is per call)."""
    import torch
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    H, seed = 256, 77
    rs = np.random.RandomState(seed)
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, diagnostics=True, tables=tables, seed=seed)
    pv, wd, price = C.synth_traces(tables, 41, rs, H)
    eng.set_traces(pv, wd, price, per_hub=True)
    orc = O.Oracle(H, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    orc.seed_base(seed, 0)
    eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    caps = np.tile(np.array([[40.0 * n for n in C.DEFAULT_KW["station_list"]]]), (H, 1))
    mm = C.Mismatch()
    for t in range(40):
        if t % 2:
            rows = _loads(rs, H, caps)
            orc.set_load_mode(True)
            obs, rew, done = eng.step_loads(torch.from_numpy(rows).cuda())
        else:
            rows = rs.uniform(-1, 1, (H, eng.action_stride)).astype(np.float32)
            orc.set_load_mode(False)
            obs, rew, done = eng.step(torch.from_numpy(rows).cuda())
        orc.step(rows, pv[t + 1], wd[t + 1], price[t + 1])
        torch.cuda.synchronize()
        C.compare_state(eng, orc, t + 1, mm, C.SCALE, obs=obs.cpu().numpy(), reward=rew.cpu().numpy(),
                        done=done.cpu().numpy())
    assert not mm, str(mm)


def test_nan_target_sets_error_bit(tables):
    import torch
    from charginghub_env_b200 import HubEngine
    eng = HubEngine(64, [C.DEFAULT_KW], device=0, tables=tables, seed=1)
    pv, wd, price = C.synth_traces(tables, 4, np.random.RandomState(0), None)
    eng.set_traces(pv, wd, price)
    eng.reset()
    rows = torch.zeros((64, 4), device="cuda")
    rows[5, 1] = float("nan")
    eng.step_loads(rows)
    torch.cuda.synchronize()
    err = eng.tensor("error").cpu().numpy().reshape(-1)
    assert err[5] == 8 and (np.delete(err, 5) == 0).all()   # CHUB_HUBERR_BAD_LOAD, sticky, only that hub
