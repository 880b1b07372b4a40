"""CPU, only where /root/reference exists: the C restatement against the LIVE unmodified reference
(pybind11 shim over CHS.hpp + the reference's own Python layer), larger than the committed fixtures."""
import numpy as np
import pytest

from oracle import ref_env
from tests import common as C

pytestmark = pytest.mark.skipif(not (ref_env.reference_available() and ref_env.shim_path()),
                                reason="needs /root/reference and oracle/_ref (make -C oracle ref)")


def _compare(kw, H, T, seed):
    from oracle import oracle as O, ref_rollout
    actions = np.random.RandomState(99 + seed).uniform(-1, 1, (T, H, sum(kw["station_list"]) + 2)).astype(np.float32)
    ref, env = ref_rollout.reference_rollout(kw, seed, np.arange(H), actions)
    tabs = ref_rollout.reference_tables()
    orc = O.Oracle(H, [kw], tables=tabs, el_tables=[np.array(env.hy_sys.hy_power_speed_list)])
    orc.seed_base(seed, 0)
    orc.reset(ref["pv"][0], ref["wd"][0], ref["price"][0])
    exact = ["pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already", "pile_soc", "pile_target",
             "pile_power", "pile_arrive", "pile_emergency", "st_line", "st_flow_in", "st_car_number",
             "st_min_power", "st_charge_power", "st_max_power", "hub_uctr", "hub_lcg", "fcev_qn"]
    for t in range(T + 1):
        if t:
            orc.step(actions[t - 1], ref["pv"][t], ref["wd"][t], ref["price"][t])
        for name in exact:
            b = orc.get(name)
            if name in ("hub_uctr", "hub_lcg"):
                b = b.astype(np.uint32)
            assert np.array_equal(ref[name][t].astype(b.dtype), b), (name, t)
        for name in ("store", "store_soc") + (("reward", "income", "demand", "hy_use", "not_meet") if t else ()):
            assert np.allclose(ref[name][t], orc.get(name), rtol=1e-12, atol=1e-12), (name, t)
        assert np.allclose(ref["obs"][t], orc.get("obs")[:, :ref["obs"].shape[2]], rtol=1e-12, atol=1e-12)
        if 0 < t <= 96:
            # what the path's one collective carries (evcssp_manager.py:259-297, CHS.hpp:916-921 / 1081-1086): the
            # reference's own accumulators, and its "charge is not complete" prints counted from its stdout
            for name in ("cum_income", "cum_draw", "ep_return", "deviation"):
                assert np.allclose(ref[name][t], orc.get(name), rtol=1e-12, atol=1e-12), (name, t)
            assert np.array_equal(ref["incomplete"][t], orc.get("st_incomplete").sum(axis=1)), ("incomplete", t)
            if t == 96:
                assert ref["hub_done"][t].all()
                assert np.allclose(ref["test_penalty"][t], orc.get("test_penalty"), rtol=1e-12, atol=0), "test_penalty"
    return ref


def test_default_hub_live():
    _compare(dict(ref_env.DEFAULT_KWARGS), 24, 96, 777)


def test_five_days_live():
    """Past the first day: the reference keeps stepping after done (time wraps, evcssp_manager.py:271-299); 5 x 96 steps.
    (Run once at 160 hubs x 960 steps and, with 9 fast + 17 slow piles, permeate 0.05 and store loss, 120 x 960: equal.)"""
    _compare(dict(ref_env.DEFAULT_KWARGS), 24, 480, 4242)


def test_unclipped_table_equals_reference_list():
    from oracle import oracle as O
    Env = ref_env.load_reference_env_class()
    with ref_env.quiet():
        env = Env(**ref_env.DEFAULT_KWARGS)
    orc = O.Oracle(1, [dict(ref_env.DEFAULT_KWARGS)])
    assert np.array_equal(np.array(env.hy_sys.hy_power_speed_list), orc.class_table(0))


def test_curves_against_reference_functions():
    """Each restated curve against the compiled reference function on 200k random fp32 arguments."""
    from oracle import oracle as O
    ps = ref_env.load_shim()
    rs = np.random.RandomState(5)
    for which, lo, hi in [(0, -1, 16), (1, -1, 16), (2, -5, 105), (3, -1, 4), (4, -1, 4), (5, -5, 105)]:
        x = rs.uniform(lo, hi, 200000).astype(np.float32)
        for cp in (False, True):
            a = np.asarray(ps.curve_batch(which, x, cp))
            b = O.curve_batch(which, x, cp)
            assert np.array_equal(a, b), (which, cp)


def _compare_loads(kw, H, T, seed):
    """Scalar-load path: the reference's stations stepped through their own evs_step(float) inside the unmodified
    Python layer (oracle/ref_rollout._LoadStation is wiring only) against the oracle in load mode."""
    from oracle import oracle as O, ref_rollout
    rs = np.random.RandomState(5 + seed)
    actions = rs.uniform(-1, 1, (T, H, sum(kw["station_list"]) + 2)).astype(np.float32)
    loads = np.zeros((T, H, 2), np.float32)
    for k in range(2):   # a power target per station: anything from "nothing" to more than every pile at full power
        cap = 40.0 * kw["station_list"][k]
        loads[:, :, k] = (rs.uniform(-0.1, 1.0, (T, H)) ** 2 * cap).astype(np.float32)
    loads[rs.uniform(size=loads.shape) < 0.1] = 1e9    # and now and then "everything"
    ref, env = ref_rollout.reference_rollout(kw, seed, np.arange(H), actions, loads=loads)
    tabs = ref_rollout.reference_tables()
    orc = O.Oracle(H, [kw], tables=tabs, el_tables=[np.array(env.hy_sys.hy_power_speed_list)])
    orc.set_load_mode(True)
    orc.seed_base(seed, 0)
    orc.reset(ref["pv"][0], ref["wd"][0], ref["price"][0])
    exact = ["pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already", "pile_soc", "pile_target",
             "pile_power", "pile_arrive", "pile_emergency", "st_line", "st_flow_in", "st_car_number",
             "st_min_power", "st_charge_power", "st_max_power", "hub_uctr", "hub_lcg", "fcev_qn"]
    n_on = 0
    for t in range(T + 1):
        if t:
            rows = np.concatenate([loads[t - 1], actions[t - 1][:, -2:]], axis=1)
            orc.step(rows, ref["pv"][t], ref["wd"][t], ref["price"][t])
        for name in exact:
            b = orc.get(name)
            if name in ("hub_uctr", "hub_lcg"):
                b = b.astype(np.uint32)
            assert np.array_equal(ref[name][t].astype(b.dtype), b), (name, t)
        n_on += int(ref["pile_charge"][t].sum())
        for name in ("store", "store_soc") + (("reward", "income", "demand", "hy_use", "not_meet") if t else ()):
            assert np.allclose(ref[name][t], orc.get(name), rtol=1e-12, atol=1e-12), (name, t)
        assert np.allclose(ref["obs"][t], orc.get("obs")[:, :ref["obs"].shape[2]], rtol=1e-12, atol=1e-12)
    assert n_on > 10 * H   # the dispatch did switch piles on


def test_scalar_load_path_live():
    _compare_loads(dict(ref_env.DEFAULT_KWARGS), 16, 96, 4242)


def test_scalar_load_path_constant_charging_live():
    _compare_loads(dict(ref_env.DEFAULT_KWARGS, constant_charging=True), 8, 96, 4243)
