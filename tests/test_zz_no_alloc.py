"""-m gpu, collected LAST on purpose (a failure here must not hide other tests).

chub_step_loads_static (include/chub.h): the scalar-load step for callers that must not allocate -- the dispatch kernel's
scratch is a file-scope `static __device__` array of the library instead of a buffer cudaMalloc'ed on first use.  It has no
counterpart in the reference, so the witness is the existing entry point:
two engines on the same streams and inputs, one stepped through chub_step_loads, the other through the new call, must
agree on every output and on the state, on two calls in a row (the shared array keeps the first call's contents) and on
more calls after them.  Called through the C ABI (ctypes), passes alone in a fresh process and after the other tests."""
import ctypes

import numpy as np
import pytest

from tests import common as C

pytestmark = pytest.mark.gpu


def _classes(rs, n):
    out = []
    for _ in range(n):
        out.append(dict(station_list=[int(rs.randint(4, 33)), int(rs.randint(4, 33))],
                        station_type_list=[["fast", "slow"], ["slow", "fast"], ["fast", "fast"], ["slow", "slow"]][rs.randint(4)],
                        constant_charging=bool(rs.rand() < 0.2), hydro_prod_rate=float(rs.uniform(50, 430)),
                        hydro_store_vlt=float(rs.uniform(10, 100)), init_soc=float(rs.uniform(0.2, 0.8)),
                        fc_max_power=float(rs.choice([50, 100, 200])), fcev_permeate=float(rs.choice([0.005, 0.01, 0.02]))))
    return out


def _pair(tables, H, classes, hub_class, seed, rs):
    from charginghub_env_b200 import HubEngine
    pv, wd, price = C.synth_traces(tables, 96, rs)
    engs = []
    for _ in range(2):
        e = HubEngine(H, classes, hub_class=hub_class, device=0, diagnostics=True, tables=tables, seed=seed,
                      auto_reset=True, action_stride=66)
        e.set_traces(pv, wd, price)
        e.reset()
        engs.append(e)
    return engs


def _same(a, b, t):
    """Outputs and state of the two engines: the tolerance of the parity tests is RTOL relative for fp64 quantities and
    bit-identity for integers and fp32 state; both calls run the same kernels on the same bits, so everything is equal."""
    import torch
    torch.cuda.synchronize()
    for name in ("obs", "reward", "done", "stats", "last_stats", "error"):
        x, y = a.tensor(name).cpu().numpy(), b.tensor(name).cpu().numpy()
        assert C.close(x, y).all() and np.array_equal(x, y), (name, t)
    for name in C.INT_FIELDS + C.PILE_FLOATS + ["store"]:
        assert np.array_equal(a.get(name), b.get(name)), (name, t)


def test_static_scratch_step_equals_step_loads(tables):
    import torch
    rs = np.random.RandomState(2024)
    classes = _classes(rs, 12)
    H = 1500 - 11                                     # ragged last warp
    hub_class = rs.randint(0, len(classes), H).astype(np.int32)
    ref, new = _pair(tables, H, classes, hub_class, 77, rs)
    caps = np.array([[40.0 * n for n in classes[k]["station_list"]] for k in hub_class], np.float32)
    lib = new.lib
    stream = ctypes.c_void_p(torch.cuda.current_stream(0).cuda_stream)
    for e in (ref, new):                              # DLPack views made before the calls under test
        for name in ("obs", "reward", "done"):
            e.tensor(name)
    n_on = 0
    for t in range(40):                               # the first two calls are the required "two calls in a row"
        loads = np.concatenate([rs.uniform(0, 1, (H, 2)).astype(np.float32) ** 2 * caps,
                                rs.uniform(-1, 1, (H, 2)).astype(np.float32)], axis=1)
        rows = torch.from_numpy(loads).cuda()
        ref.step_loads(rows)
        rc = lib.chub_step_loads_static(new.h, ctypes.c_void_p(rows.data_ptr()), None, None, None, stream)   # the C ABI itself
        assert rc == 0, lib.chub_last_error()
        if t < 4 or t % 6 == 0:
            _same(ref, new, t)
        n_on += int(new.get("pile_charge").sum())
    _same(ref, new, 40)
    assert n_on > 10 * H                              # the dispatch did switch piles on
    assert new.counters()["error_or"] == ref.counters()["error_or"]
    # results written straight into the caller's buffers (the chub_step_into form), through the Python wrapper
    obs = torch.empty((H, 13), device="cuda")
    rew = torch.empty(H, device="cuda")
    done = torch.empty(H, device="cuda", dtype=torch.uint8)
    loads = torch.from_numpy(np.concatenate([rs.uniform(0, 1, (H, 2)).astype(np.float32) * caps,
                                             rs.uniform(-1, 1, (H, 2)).astype(np.float32)], axis=1)).cuda()
    ref.step_loads(loads)
    new.step_loads_static(loads, obs, rew, done)
    torch.cuda.synchronize()
    assert torch.equal(obs, ref.tensor("obs")) and torch.equal(rew, ref.tensor("reward")) and torch.equal(done, ref.tensor("done"))


def test_two_engines_share_the_array_when_serialised(tables):
    """One array per device for every engine: calls that do NOT overlap (same stream) may interleave freely."""
    import torch
    rs = np.random.RandomState(5)
    a_ref, a_new = _pair(tables, 300, [C.DEFAULT_KW], None, 1, rs)
    b_ref, b_new = _pair(tables, 700, [dict(C.DEFAULT_KW, station_list=[32, 9])], None, 2, rs)
    for t in range(6):
        la = torch.cat([torch.rand((300, 2), device="cuda") * 600, torch.rand((300, 2), device="cuda") * 2 - 1], dim=1).contiguous()
        lb = torch.cat([torch.rand((700, 2), device="cuda") * 900, torch.rand((700, 2), device="cuda") * 2 - 1], dim=1).contiguous()
        a_ref.step_loads(la); b_ref.step_loads(lb)
        a_new.step_loads_static(la); b_new.step_loads_static(lb)
    _same(a_ref, a_new, 6)
    _same(b_ref, b_new, 6)


def test_sizes_beyond_the_static_maximum_are_refused(tables):
    import torch
    from charginghub_env_b200 import ChubError, HubEngine, _lib
    H = _lib.STATIC_MAX_HUBS + 32
    small = dict(C.DEFAULT_KW, station_list=[4, 4])   # 8 piles per hub: the engine itself stays small
    eng = HubEngine(H, [small], device=0, tables=tables, seed=3, action_stride=10)
    eng.set_traces(*C.synth_traces(tables, 96, np.random.RandomState(0)))
    loads = torch.zeros((H, 4), device="cuda")
    with pytest.raises(ChubError) as ei:
        eng.step_loads_static(loads)
    assert "CHUB_STATIC_MAX_HUBS" in str(ei.value)
    rc = eng.lib.chub_step_loads_static(eng.h, ctypes.c_void_p(loads.data_ptr()), None, None, None, None)
    assert rc == -1 and b"1048576" in eng.lib.chub_last_error()          # CHUB_ERR_ARG with a clear message
    eng.reset()
    eng.step_loads(loads)                             # the allocating entry point takes the same engine
    torch.cuda.synchronize()
    assert eng.counters()["step_index"] == 1
    eng.close()
