"""TEST INFRASTRUCTURE ONLY.  Generate tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference through oracle/ref_env.py) on seeded streams and pre-drawn actions.

The reference's own tests hold no golden vectors, so these traces -- outputs of
the reference itself, produced in this container (glibc / libstdc++ of this image, Python 3.12) --
are what pins the oracle (tests/test_oracle_golden.py) and the CUDA engine (tests/test_gpu_golden.py)
on machines where /root/reference does not exist.  Run:  python -m oracle.gen_golden
"""
import os
import sys

import numpy as np

from . import ref_env, ref_rollout

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
BASE = dict(ref_env.DEFAULT_KWARGS)

CASES = {
    # name: (kwargs, hubs, steps, base_seed)
    "default_hub": (BASE, 6, 96, 0x5EED),
    "small_fast_big_slow_perm05": (dict(BASE, station_list=[4, 32], fcev_permeate=0.05, hydro_prod_rate=430,
                                        hydro_store_vlt=10.0, init_soc=0.8, fc_max_power=200), 4, 96, 11),
    "slow_first_loss": (dict(BASE, station_list=[32, 5], station_type_list=["slow", "fast"], fcev_permeate=0.02,
                             hydro_prod_rate=50, hydro_store_vlt=100.0, init_soc=0.45, fc_max_power=50,
                             hydro_loss=0.02), 4, 96, 12),
    "constant_charging": (dict(BASE, constant_charging=True), 3, 96, 13),
    "two_fast_low_store": (dict(BASE, station_list=[7, 9], station_type_list=["fast", "fast"], fcev_permeate=0.05,
                                hydro_store_vlt=10.0, init_soc=0.1), 3, 96, 14),
    "no_fast_station": (dict(BASE, station_list=[0, 25]), 3, 96, 16),      # 9-entry observation
    "no_slow_station": (dict(BASE, station_list=[20, 0]), 3, 96, 17),
    "max_piles_both": (dict(BASE, station_list=[32, 32], init_soc=0.7), 3, 96, 18),   # the widest stations the engine takes
    "two_slow_defaults_none": (dict(BASE, station_list=[12, 12], station_type_list=["slow", "slow"],
                                    hydro_prod_rate=None, hydro_store_vlt=None, fc_max_power=None, init_soc=0.5),
                               3, 96, 15),
}

# scalar-load path (stations stepped through evs_step(float), CHS.hpp:1169-1186 / 1478-1495): the fixture also
# holds "loads" float32 [T, H, 2], the station power targets in kW
LOAD_CASES = {
    "load_default_hub": (BASE, 5, 96, 21),
    "load_constant_charging": (dict(BASE, constant_charging=True), 3, 96, 22),
    "load_slow_first_small": (dict(BASE, station_list=[9, 3], station_type_list=["slow", "fast"], init_soc=0.6), 3, 96, 23),
}

KEEP = ["pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already", "pile_soc", "pile_target",
        "pile_power", "pile_arrive", "st_line", "st_flow_in", "st_car_number", "st_min_power", "st_charge_power",
        "st_max_power", "hub_t", "hub_uctr", "hub_lcg", "fcev_arrive", "fcev_qn", "hub_done", "store", "store_soc",
        "reward", "income", "hy_flow", "h2_power", "fc_power", "hy_use", "not_meet", "hy_to_use", "demand",
        "h2_grid", "hy_act", "pv", "wd", "price", "obs",
        # what the reference accumulates and reports at `done` (evcssp_manager.py:259-297; "charge is not complete" prints
        # of CHS.hpp:916-921 / 1081-1086 counted from its stdout): the payload of the path's one collective
        "cum_income", "cum_draw", "ep_return", "test_penalty", "deviation", "incomplete"]


def main():
    os.makedirs(OUT, exist_ok=True)
    tabs = ref_rollout.reference_tables()
    only = set(sys.argv[1:])
    for name, (kw, H, T, seed) in list(CASES.items()) + list(LOAD_CASES.items()):
        if only and name not in only:
            continue
        A = sum(kw["station_list"]) + 2
        rs = np.random.RandomState(1234 + seed)
        actions = rs.uniform(-1, 1, (T, H, A)).astype(np.float32)
        loads = None
        if name in LOAD_CASES:
            cap = np.array([40.0 * n for n in kw["station_list"]])
            loads = (rs.uniform(-0.1, 1.0, (T, H, 2)) ** 2 * cap).astype(np.float32)
            loads[rs.uniform(size=loads.shape) < 0.1] = 1e9
        out, env = ref_rollout.reference_rollout(kw, seed, np.arange(H), actions, loads=loads)
        pack = {k: out[k] for k in KEEP}
        for k in ("pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already"):
            pack[k] = pack[k].astype(np.int8)
        for k in ("st_line", "st_flow_in", "st_car_number", "hub_t", "fcev_arrive", "fcev_qn", "hub_done"):
            pack[k] = pack[k].astype(np.int16)
        pack["hub_uctr"] = pack["hub_uctr"].astype(np.uint32)
        pack["hub_lcg"] = pack["hub_lcg"].astype(np.uint32)
        pack["actions"] = actions
        if loads is not None:
            pack["loads"] = loads
        pack["el_table"] = np.array(env.hy_sys.hy_power_speed_list, np.float64)
        pack["transformer"] = np.array([s.transformer_limit for s in env.env_aggregator.evcssp_evs_objects], np.float64)
        kws = dict(kw)
        pack["kwargs_repr"] = np.array(repr(kws))
        pack["base_seed"] = np.array(seed, np.uint64)
        path = os.path.join(OUT, name + ".npz")
        np.savez_compressed(path, **pack)
        print(name, H, "hubs", T, "steps ->", os.path.getsize(path), "bytes")
    np.savez_compressed(os.path.join(OUT, "tables_check.npz"), flow_row0=tabs["flow"][0], flow_row50=tabs["flow"][50],
                        price=tabs["price"], price_mean=tabs["price_mean"], price_std=tabs["price_std"],
                        sin_table=tabs["sin_table"])


if __name__ == "__main__":
    main()
