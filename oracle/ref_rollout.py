"""TEST INFRASTRUCTURE ONLY.  Drive the UNMODIFIED reference env (oracle/ref_env.py) hub by hub
with per-hub seeded streams and a pre-drawn action tensor, recording everything the parity tests
compare: per-pile columns and members, station outputs, hydrogen state, reward, obs, and the
exogenous traces (pv*5, wd, price incl. OU noise) that the engines take as inputs.

Only works where /root/reference exists (this container).  Used by oracle/gen_golden.py to write
tests/golden/*.npz and by tests/test_oracle_vs_reference.py.
"""
import random
import sys

import numpy as np

from . import ref_env

MAXP = 32


def hub_streams(base_seed, hub_ids):
    """Python mirror of include/chub_rng.h chub_hub_streams (uint64 arithmetic)."""
    M = (1 << 64) - 1

    def sm(z):
        z = (z + 0x9E3779B97F4A7C15) & M
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    k0, k1, e0 = [], [], []
    for h in hub_ids:
        s = sm((base_seed ^ ((int(h) * 0xD1342543DE82EF95) & M)) & M)
        r = sm(s)
        k0.append(s & 0xFFFFFFFF)
        k1.append(s >> 32)
        e0.append(1 + (r % 2147483646))
    return (np.array(k0, np.uint32), np.array(k1, np.uint32), np.array(e0, np.uint32))


def _snap_station(st, out, t, h, s):
    d = st.dump()
    n = len(d["car"])
    car = d["car"] > 0.5
    out["pile_car"][t, h, s, :n] = car
    out["pile_charge"][t, h, s, :n] = d["charge"] > 0.5
    out["pile_urgent"][t, h, s, :n] = car & (d["emergency"] > 8)
    out["pile_stay"][t, h, s, :n] = np.where(car, d["stay_time"], 0)
    out["pile_already"][t, h, s, :n] = np.where(car, d["already_stay_time"], 0)
    out["pile_soc"][t, h, s, :n] = np.where(car, d["pile_current_soc"], 0)
    out["pile_target"][t, h, s, :n] = np.where(car, d["pile_target_soc"], 0)
    out["pile_power"][t, h, s, :n] = d["power"]
    out["pile_arrive"][t, h, s, :n] = np.where(car, d["pile_arrive_soc"], 0)
    out["pile_emergency"][t, h, s, :n] = d["emergency"]
    out["st_line"][t, h, s] = d["line"]
    out["st_flow_in"][t, h, s] = d["flow_in_last"]
    out["st_car_number"][t, h, s] = d["car_number"]
    out["st_min_power"][t, h, s] = d["min_power"]
    out["st_charge_power"][t, h, s] = d["charge_power"]
    out["st_max_power"][t, h, s] = d["max_power"]


class _LoadStation(object):
    """Wiring only: forwards everything to the reference station object, but answers the aggregator's
    evs_step(Vector_float) (Aggregator_Simple.py:141-142) with the station's own evs_step(float) (CHS.hpp:1169 /
    1478) on the power target set for this step.  Both sides of the call stay the unmodified reference."""

    def __init__(self, st):
        object.__setattr__(self, "_st", st)
        object.__setattr__(self, "load", 0.0)

    def evs_step(self, _pile_actions):
        self._st.evs_step(float(np.float32(self.load)))

    def __getattr__(self, name):
        return getattr(self._st, name)

    def __setattr__(self, name, value):
        if name == "load":
            object.__setattr__(self, name, value)
        else:
            setattr(self._st, name, value)


def reference_rollout(kwargs, base_seed, hub_ids, actions, py_seed=0, env=None, loads=None):
    """actions: float32 [T, H, A].  loads: None, or float32 [T, H, 2] station power targets (kW): the stations are then
    stepped through evs_step(float) and the per-pile entries of `actions` are ignored.
    Returns (dict of arrays with leading [T+1, H], env)."""
    ps = ref_env.load_shim()
    Env = ref_env.load_reference_env_class()
    T, H, A = actions.shape
    assert H == len(hub_ids)
    k0, k1, e0 = hub_streams(base_seed, hub_ids)
    random.seed(py_seed)
    np.random.seed(py_seed)
    if env is None:
        with ref_env.quiet():
            env = Env(**kwargs)
    if loads is not None:
        objs = env.env_aggregator.evcssp_evs_objects
        for k in range(len(objs)):
            if not isinstance(objs[k], _LoadStation):
                objs[k] = _LoadStation(objs[k])
    S = T + 1
    out = {}
    for name in ("pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already"):
        out[name] = np.zeros((S, H, 2, MAXP), np.int32)
    for name in ("pile_soc", "pile_target", "pile_power", "pile_arrive", "pile_emergency"):
        out[name] = np.zeros((S, H, 2, MAXP), np.float32)
    for name in ("st_line", "st_flow_in", "st_car_number"):
        out[name] = np.zeros((S, H, 2), np.int32)
    for name in ("st_min_power", "st_charge_power", "st_max_power"):
        out[name] = np.zeros((S, H, 2), np.float32)
    for name in ("hub_t", "hub_uctr", "hub_lcg", "fcev_arrive", "fcev_qn", "hub_done"):
        out[name] = np.zeros((S, H), np.int64)
    for name in ("store", "store_soc", "reward", "income", "hy_flow", "h2_power", "fc_power", "hy_use",
                 "not_meet", "hy_to_use", "demand", "h2_grid", "hy_act",
                 "pv", "wd", "price", "cum_income", "cum_draw", "ep_return", "test_penalty", "deviation"):
        out[name] = np.zeros((S, H), np.float64)
    out["incomplete"] = np.zeros((S, H), np.int64)   # departures below target since the reset (counted from stdout)
    nobs = len(env.low_state)
    out["obs"] = np.zeros((S, H, nobs), np.float64)

    def snap(t, h):
        for s, st in enumerate(env.env_aggregator.evcssp_evs_objects):
            _snap_station(st, out, t, h, s)
        out["hub_t"][t, h] = int(env.real_state[0])
        out["hub_uctr"][t, h] = ps.hook_u_counter()
        out["hub_lcg"][t, h] = ps.hook_e_state()
        out["store"][t, h] = env.hy_sys.sty.capacity
        out["store_soc"][t, h] = env.hy_sys.sty.Store_SOC
        out["fcev_qn"][t, h] = len(env.hy_sys.hvs.needed_time_list)
        out["pv"][t, h] = env.re_pv_power
        out["wd"][t, h] = env.re_wd_power
        out["price"][t, h] = env.real_state[1]
        out["obs"][t, h] = env.state

    with ref_env.capture() as cap:
        for h in range(H):
            ps.hook_seed(int(k0[h]), int(k1[h]), 0, int(e0[h]))
            env.reset()
            snap(0, h)
            cap.new_text()
            ret0 = env.acumulate_reward       # never reset by the reference (evcssp_manager.py:269, 308)
            for t in range(T):
                if loads is not None:
                    for k, st in enumerate(env.env_aggregator.evcssp_evs_objects):
                        st.load = loads[t, h, k]
                _, r, d, _ = env.step(actions[t, h])
                snap(t + 1, h)
                out["reward"][t + 1, h] = r
                out["hub_done"][t + 1, h] = d
                out["income"][t + 1, h] = env.income
                out["fcev_arrive"][t + 1, h] = env.hy_sys.hvs.arrive_number
                out["hy_flow"][t + 1, h] = env.hy_sys.hy_flow_speed
                out["h2_power"][t + 1, h] = env.hy_sys.all_power_second
                out["fc_power"][t + 1, h] = env.fc_power
                out["hy_use"][t + 1, h] = env.hy_sys.sty.hy_use
                out["not_meet"][t + 1, h] = env.hy_sys.sty.not_meet
                out["hy_to_use"][t + 1, h] = env.hfc.hy_to_use
                out["demand"][t + 1, h] = env.hy_sys.hvs.total_mass_need
                out["h2_grid"][t + 1, h] = env.re_hydrogen_power
                out["hy_act"][t + 1, h] = env.hy_act
                out["cum_income"][t + 1, h] = env.cumulated_income
                out["cum_draw"][t + 1, h] = env.cumulated_draw_ele
                out["ep_return"][t + 1, h] = env.acumulate_reward - ret0
                out["deviation"][t + 1, h] = env.deviation
                if d:
                    out["test_penalty"][t + 1, h] = env.test_penalty
                out["incomplete"][t + 1, h] = out["incomplete"][t, h] + cap.new_text().count("charge is not complete")
    ps.hook_off()
    return out, env


def reference_tables(env=None):
    """Static tables of the reference: arrival CDF (as parsed by CHS.hpp:93-175), prices, norms."""
    ps = ref_env.load_shim()
    ref_env.load_reference_env_class()
    agg = sys.modules["lion_cpp20.Aggregator_Simple"]
    price = np.array(agg.mean_for_MAD, np.float64)
    k = 2 * np.pi / 96
    return dict(flow=ps.flow_table(), price=price, price_mean=float(np.mean(price)),
                price_std=float(np.std(price)), sin_table=np.sin(k * np.arange(96, dtype=np.float64)))
