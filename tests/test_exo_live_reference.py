"""CPU, only where /root/reference exists: the package's host generator of PV / wind / price
(charginghub_env_b200/exogenous.py HubExogenous) against the reference RUN LIVE -- renewable.ReNew.get_pv_power /
get_wd_power / renew_reset, OU_Noise.sample (lion_cpp20/renewable.py:21-76) and EvcsspManagerEnv_v6.make_state's draw
order and held price noise (evcssp_manager.py:344-361, 304-316) -- on the same seeds of Python's `random` and numpy's
global generator.  tests/test_exo_oracle.py then ties oracle/exo_oracle.c (and through it the device generator,
tests/test_gpu_exo.py) to HubExogenous, so this is the link that pins the whole chain on the reference."""
import random

import numpy as np
import pytest

from oracle import ref_env

pytestmark = pytest.mark.skipif(not (ref_env.reference_available() and ref_env.shim_path()),
                                reason="needs /root/reference and oracle/_ref (make -C oracle ref)")


@pytest.mark.parametrize("renew_fluctuate,price_fluctuate", [(0.0, 0.0), (0.3, 0.2)])
def test_host_generator_equals_the_live_reference(tables, renew_fluctuate, price_fluctuate):
    from charginghub_env_b200.exogenous import HubExogenous
    Env = ref_env.load_reference_env_class()
    kw = dict(ref_env.DEFAULT_KWARGS, renew_fluctuate=renew_fluctuate, price_fluctuate=price_fluctuate)
    EPISODES = 4

    # ---- the reference, live: constructor (which resets once, evcssp_manager.py:120), then reset + 96 steps, 4 times
    random.seed(7)
    np.random.seed(7)
    ref = []
    with ref_env.quiet():
        env = Env(**kw)       # its reset() consumes draws; the state it made is cleared again (evcssp_manager.py:121-123)
        for ep in range(EPISODES):
            env.reset()
            ref.append((env.re_pv_power, env.re_wd_power, float(env.real_state[1]), env.renew.pv_day, env.renew.wd_day))
            for t in range(96):
                env.step(None)
                ref.append((env.re_pv_power, env.re_wd_power, float(env.real_state[1]), env.renew.pv_day, env.renew.wd_day))

    # ---- the package's generator on the same seeds, driven in the order EvcsspManagerEnv_v6 (env.py) drives it
    random.seed(7)
    np.random.seed(7)
    exo = HubExogenous(tables, renew_fluctuate, price_fluctuate)
    got = []

    def reset():
        exo.new_day()
        v = exo.state_values(0, float(tables["price"][95]))
        exo.reset_price_count()
        got.append(v + (exo.pv_day, exo.wd_day))

    reset()
    got.clear()
    for ep in range(EPISODES):
        reset()
        for t in range(96):
            v = exo.state_values((t + 1) % 96, float(tables["price"][t]))
            got.append(v + (exo.pv_day, exo.wd_day))

    assert len(got) == len(ref) == EPISODES * 97
    for k, (a, b) in enumerate(zip(got, ref)):
        assert a[3:] == b[3:], ("days", k, a, b)
        assert a[0] == b[0] and a[1] == b[1] and a[2] == b[2], ("values", k, a, b)   # bit for bit
    days = {(g[3], g[4]) for g in got}
    assert len(days) >= 3                      # several different PV / wind days were drawn


def test_price_table_and_day_tables_equal_the_reference_data(tables):
    """The tables HubExogenous reads are the reference's own data files (renewable.py:9-18, Aggregator_Simple.py:9-15)."""
    import sys
    ref_env.load_reference_env_class()
    ren = sys.modules["lion_cpp20.renewable"]
    agg = sys.modules["lion_cpp20.Aggregator_Simple"]
    assert np.array_equal(np.asarray(ren.pv_power_data, np.float64), tables["pv_days"])
    assert np.array_equal(np.asarray(ren.wd_power_data, np.float64), tables["wd_days"])
    assert np.array_equal(np.asarray(agg.mean_for_MAD, np.float64), tables["price"])
