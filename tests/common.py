"""Shared helpers of the parity tests: build matching engine / oracle pairs and compare them."""
import numpy as np

DEFAULT_KW = dict(station_list=[20, 25], station_type_list=["fast", "slow"], constant_charging=False,
                  hydro_prod_rate=100, hydro_store_vlt=25.0, init_soc=0.2, fc_max_power=100,
                  fcev_permeate=0.01, hydro_loss=0.0)  # test/env_test.py:10-30

INT_FIELDS = ["pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already", "st_line", "st_flow_in",
              "st_car_number", "hub_t", "hub_uctr", "hub_lcg", "fcev_arrive", "fcev_qn"]
PILE_FLOATS = ["pile_soc", "pile_target", "pile_power", "pile_arrive"]
ST_FLOATS = ["st_min_power", "st_charge_power", "st_max_power"]
DIAG = ["reward", "income", "hy_flow", "h2_power", "fc_power", "hy_use", "not_meet", "hy_to_use", "demand",
        "ev_grid0", "ev_grid1", "h2_grid", "hy_act", "store_soc"]
RTOL = 1e-5  # BASELINE.json north_star: float state within 1e-5 relative in fp32


def oracle_tables(tabs):
    return dict(flow=tabs["arrival_cdf"], price=tabs["price"], price_mean=tabs["price_mean"],
                price_std=tabs["price_std"], sin_table=tabs["time_sin"])


def synth_traces(tabs, L, rs, H=None):
    """Exogenous traces in the data's ranges.  H given -> per-hub [L, H]."""
    shape = (L,) if H is None else (L, H)
    t = np.arange(L).reshape((L,) + (1,) * (len(shape) - 1))
    pv = np.maximum(0.0, 5.0 * 40.0 * np.sin(np.pi * ((t % 96) - 24) / 48.0)) * rs.uniform(0.5, 1.0, shape)
    wd = np.clip(46 + 40 * np.sin(2 * np.pi * t / 355.0) + rs.normal(0, 5, shape), 0, 98)
    price = tabs["price"][(t - 1) % 96] + rs.normal(0, 0.005, shape)
    return np.ascontiguousarray(pv * np.ones(shape)), np.ascontiguousarray(wd), np.ascontiguousarray(price)


ZERO_FLOOR = 1e-12  # absolute floor, used only where the oracle's value is exactly 0


def close(a, b, rtol=RTOL, atol=0.0):
    """|a - b| <= rtol |b| (+ atol where a caller states one); where b is exactly 0, |a| <= max(atol, ZERO_FLOOR)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.abs(a - b) <= np.where(b == 0.0, np.maximum(atol, ZERO_FLOOR), atol + rtol * np.abs(b))


class Mismatch(object):
    def __init__(self):
        self.items = []

    def add(self, t, name, n, worst):
        self.items.append((t, name, n, worst))

    def __bool__(self):
        return bool(self.items)

    def __str__(self):
        return "; ".join("t=%d %s: %d bad (worst %.3g)" % x for x in self.items[:12])


def compare_state(eng, orc, t, mm, scale=None, obs=None, reward=None, done=None, float_exact_stats=None):
    """Integers bit-exact; fp32 state (per-pile SOC / target / power / arrival SOC, station min / charge / max power)
    BIT-IDENTICAL -- tests/test_gpu_curves.py proves the curves equal glibc's on every reachable argument and the station
    sums run in the reference's order; fp64 quantities and the fp32 outputs within RTOL relative (north_star's 1e-5), with
    an absolute floor only where the oracle's value is exactly 0.  `scale` is unused (kept for the callers' signature)."""
    for name in INT_FIELDS:
        a, b = eng.get(name), orc.get(name)
        if name in ("hub_uctr", "hub_lcg"):
            a, b = a.astype(np.uint32), b.astype(np.uint32)
        bad = a != b
        if bad.any():
            mm.add(t, name, int(bad.sum()), float(np.abs(a.astype(np.int64) - b.astype(np.int64)).max()))
    for name in PILE_FLOATS + ST_FLOATS:
        a, b = eng.get(name), orc.get(name)
        if name == "pile_power":
            b = b * orc.get("pile_car")
        bad = a.view(np.uint32) != b.astype(np.float32).view(np.uint32)
        bad &= ~((a == 0) & (b == 0))       # +0 / -0
        if float_exact_stats is not None:
            float_exact_stats[name] = float_exact_stats.get(name, 0) + int(bad.sum())
        if bad.any():
            mm.add(t, name, int(bad.sum()), float(np.abs(a.astype(np.float64) - b).max()))
    a, b = eng.get("store"), orc.get("store")
    ok = close(a, b)
    if not ok.all():
        mm.add(t, "store", int((~ok).sum()), float(np.abs(a - b).max()))
    if t > 0:
        for name in DIAG:
            a, b = eng.get("diag_" + name), orc.get(name)
            ok = close(a, b)
            if not ok.all():
                mm.add(t, name, int((~ok).sum()), float(np.abs(a - b).max()))
    if obs is not None:
        b = orc.get("obs")
        ok = close(obs, b, atol=1e-7)      # float32 outputs of fp64 values that may cancel to ~0 (normalised powers)
        if not ok.all():
            mm.add(t, "obs", int((~ok).sum()), float(np.abs(obs - b).max()))
    if reward is not None:
        b = orc.get("reward")
        ok = close(reward, b, atol=1e-7)
        if not ok.all():
            mm.add(t, "reward_f32", int((~ok).sum()), float(np.abs(reward - b).max()))
    if done is not None:
        b = orc.get("hub_done")
        if not np.array_equal(done.astype(np.int32), b):
            mm.add(t, "done", int((done != b).sum()), 1.0)


SCALE = {"pile_soc": 1.0, "pile_target": 1.0, "pile_power": 1.0, "pile_arrive": 1.0, "st_min_power": 10.0,
         "st_charge_power": 10.0, "st_max_power": 10.0, "reward": 1.0, "income": 10.0, "hy_flow": 1.0,
         "h2_power": 10.0, "fc_power": 10.0, "hy_use": 100.0, "not_meet": 100.0, "hy_to_use": 100.0,
         "demand": 100.0, "ev_grid0": 10.0, "ev_grid1": 10.0, "h2_grid": 10.0, "hy_act": 1.0, "store_soc": 1.0}
