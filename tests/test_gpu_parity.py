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
