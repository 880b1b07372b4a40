"""-m gpu: the CUDA engine against the committed golden traces of the UNMODIFIED reference."""
import glob
import os

import numpy as np
import pytest

from tests import common as C
from tests.test_oracle_golden import GOLD, load_case

pytestmark = pytest.mark.gpu

INTS = ["pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already", "st_line", "st_flow_in",
        "st_car_number", "hub_t", "hub_uctr", "hub_lcg", "fcev_qn"]
F32 = ["pile_soc", "pile_target", "pile_power", "pile_arrive", "st_min_power", "st_charge_power", "st_max_power"]
DIAG = ["reward", "income", "hy_flow", "h2_power", "fc_power", "hy_use", "not_meet", "hy_to_use", "demand",
        "h2_grid", "hy_act", "store_soc"]


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_engine_reproduces_reference_trace(path, tables):
    import torch
    from charginghub_env_b200 import HubEngine
    z, kw = load_case(path)
    T, H, A = z["actions"].shape
    eng = HubEngine(H, [dict(kw, el_table=z["el_table"])], device=0, diagnostics=True, tables=tables,
                    seed=int(z["base_seed"]))
    eng.set_traces(z["pv"], z["wd"], z["price"], per_hub=True)
    nobs = z["obs"].shape[2]
    obs = eng.reset()

    def check(t, obs, rew=None, done=None):
        torch.cuda.synchronize()
        for name in INTS + (["fcev_arrive"] if t else []):
            a, b = z[name][t], eng.get(name)
            if name in ("hub_uctr", "hub_lcg"):
                b = b.astype(np.uint32)
            if name == "pile_urgent":
                a = a * z["pile_car"][t]
            assert np.array_equal(a.astype(np.int64), b.astype(np.int64)), (name, t)
        for name in F32:
            a, b = z[name][t].astype(np.float64), eng.get(name)
            if name == "pile_power":
                a = a * z["pile_car"][t]
            assert np.array_equal(b, a.astype(np.float32)), (name, t)     # fp32 state: bit-identical to the reference's
        assert C.close(eng.get("store"), z["store"][t]).all(), ("store", t)
        assert C.close(obs.cpu().numpy()[:, :nobs], z["obs"][t], atol=1e-7).all(), ("obs", t)
        if t:
            for name in DIAG:
                assert C.close(eng.get("diag_" + name), z[name][t], atol=0.0).all(), (name, t)
            assert C.close(rew.cpu().numpy(), z["reward"][t], atol=1e-7).all()
            assert np.array_equal(done.cpu().numpy().astype(np.int64), z["hub_done"][t].astype(np.int64))

    check(0, obs)
    for t in range(T):
        if "loads" in z.files:   # scalar-load fixtures: stations stepped through evs_step(float)
            rows = np.concatenate([z["loads"][t], z["actions"][t][:, -2:]], axis=1).astype(np.float32)
            obs, rew, done = eng.step_loads(torch.from_numpy(rows).cuda())
        else:
            obs, rew, done = eng.step(torch.from_numpy(z["actions"][t]).cuda())
        check(t + 1, obs, rew, done)
    assert eng.counters()["error_or"] == 0
    assert np.allclose(eng.class_table(0), z["el_table"])
    if T == 96 and "cum_income" in z.files:
        # the statistics frozen at done and the terminal store SOC against what the REFERENCE accumulated and printed
        last = eng.tensor("last_stats").cpu().numpy()
        for col, name in enumerate(("ep_return", "cum_income", "cum_draw")):
            assert C.close(last[:, col], z[name][96], rtol=1e-6).all(), name
        assert np.array_equal(last[:, 3], z["incomplete"][96].astype(np.float32))
        fin = eng.tensor("final_soc").cpu().numpy()
        assert C.close(fin, z["store_soc"][96], rtol=1e-6).all()
        vlt = kw.get("hydro_store_vlt")
        cap_kg = 0.089 * 200 * (5000.0 if vlt is None else vlt)              # hydro_sys.py:94-98, grams / 1000
        pen = np.abs((fin.astype(np.float64) - kw["init_soc"]) * cap_kg / 0.2)   # VectorHubEnv's soc_penalty="abs"
        assert C.close(pen, z["test_penalty"][96], rtol=1e-5, atol=1e-7 * cap_kg / 0.2).all()
