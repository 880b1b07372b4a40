"""-m gpu: behaviour of the C-ABI surface itself -- host path, graph rollouts, checkpoints, DLPack, errors."""
import numpy as np
import pytest

from tests import common as C

pytestmark = pytest.mark.gpu


def _engine(H, tables, seed=1, **kw):
    from charginghub_env_b200 import HubEngine
    rs = np.random.RandomState(seed)
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, tables=tables, seed=seed, **kw)
    eng.set_traces(*C.synth_traces(tables, 400, rs))
    return eng, rs


def test_unclipped_table_equals_oracle_table(tables):
    from oracle import oracle as O
    eng, _ = _engine(32, tables)
    assert np.array_equal(eng.class_table(0), O.Oracle(1, [C.DEFAULT_KW]).class_table(0))
    assert eng.algorithmic_bytes(0) == 1845  # SURVEY.md 8d, default hub


def test_host_path_equals_device_path(tables):
    import torch
    a_eng, rs = _engine(300, tables, seed=3)
    b_eng, _ = _engine(300, tables, seed=3)
    a_eng.reset()
    obs_h = b_eng.reset_host().copy()
    assert np.array_equal(a_eng.tensor("obs").cpu().numpy(), obs_h)
    for t in range(20):
        a = rs.uniform(-1, 1, (300, 47)).astype(np.float32)
        o, r, d = a_eng.step(torch.from_numpy(a).cuda())
        oh, rh, dh = b_eng.step_host(a)
        assert np.array_equal(o.cpu().numpy(), oh) and np.array_equal(r.cpu().numpy(), rh)
        assert np.array_equal(d.cpu().numpy(), dh)


def test_graph_rollout_equals_stepwise(tables):
    import torch
    a_eng, rs = _engine(1000, tables, seed=5, auto_reset=True)
    b_eng, _ = _engine(1000, tables, seed=5, auto_reset=True)
    acts = torch.from_numpy(rs.uniform(-1, 1, (1000, 47)).astype(np.float32)).cuda()
    a_eng.reset()
    b_eng.reset()
    b_eng.capture_rollout(acts, 16)
    for rep in range(7):  # 112 steps: crosses the day boundary, so the trace index must advance per replay
        for _ in range(16):
            a_eng.step(acts)
        b_eng.rollout()
    torch.cuda.synchronize()
    assert a_eng.counters()["step_index"] == b_eng.counters()["step_index"] == 112
    for name in ("pile_soc", "pile_stay", "st_line", "hub_uctr", "hub_lcg", "store", "hub_t"):
        assert np.array_equal(a_eng.get(name), b_eng.get(name)), name
    assert np.array_equal(a_eng.tensor("obs").cpu().numpy(), b_eng.tensor("obs").cpu().numpy())
    assert np.array_equal(a_eng.tensor("stats").cpu().numpy(), b_eng.tensor("stats").cpu().numpy())


def test_checkpoint_roundtrip(tables):
    import torch
    eng, rs = _engine(500, tables, seed=7)
    acts = [torch.from_numpy(rs.uniform(-1, 1, (500, 47)).astype(np.float32)).cuda() for _ in range(20)]
    eng.reset()
    for a in acts[:10]:
        eng.step(a)
    blob = eng.get_state()
    for a in acts[10:]:
        eng.step(a)
    torch.cuda.synchronize()
    want = {n: eng.get(n) for n in ("pile_soc", "pile_power", "pile_stay", "st_line", "store", "hub_lcg")}
    want_obs = eng.tensor("obs").cpu().numpy().copy()
    eng.set_state(blob)
    for a in acts[10:]:
        eng.step(a)
    torch.cuda.synchronize()
    for n, v in want.items():
        assert np.array_equal(eng.get(n), v), n
    assert np.array_equal(eng.tensor("obs").cpu().numpy(), want_obs)


def test_dlpack_views_alias_engine_memory(tables):
    import torch
    eng, rs = _engine(64, tables)
    obs = eng.reset()
    assert obs.is_cuda and obs.shape == (64, 13) and obs.dtype == torch.float32
    rew, done = eng.tensor("reward"), eng.tensor("done")
    assert rew.shape == (64,) and done.dtype == torch.uint8
    before = obs.clone()
    eng.step(torch.from_numpy(rs.uniform(-1, 1, (64, 47)).astype(np.float32)).cuda())
    torch.cuda.synchronize()
    assert not torch.equal(before, obs)  # same memory, updated in place by the kernel
    keep = eng.tensor("stats")
    eng.close()  # views keep the engine alive until they die
    assert torch.isfinite(keep).all()


def test_errors_are_reported_not_printed(tables):
    import torch
    from charginghub_env_b200 import ChubError, HubEngine
    eng = HubEngine(16, [C.DEFAULT_KW], device=0, tables=tables)
    with pytest.raises(ChubError):  # no traces yet
        eng.reset()
    with pytest.raises(ValueError):  # action width, evcssp_manager.py:148
        eng.set_traces(*C.synth_traces(tables, 10, np.random.RandomState(0)))
        eng.step(torch.zeros((16, 46), device="cuda"))
    with pytest.raises(ChubError):
        HubEngine(16, [dict(C.DEFAULT_KW, init_soc=0.05)], device=0, tables=tables)  # hydro_sys.py:137
    with pytest.raises(ChubError):
        HubEngine(16, [dict(C.DEFAULT_KW, station_list=[40, 2])], device=0, tables=tables)


def test_fcev_backlog_overflow_sets_error_bit(tables):
    """A permeate of 1.0 floods the FCEV list; the engine keeps CHUB_FCEV_QCAP entries and raises the sticky bit."""
    import torch
    from charginghub_env_b200 import HubEngine
    eng = HubEngine(64, [dict(C.DEFAULT_KW, fcev_permeate=1.0)], device=0, tables=tables, seed=2)
    eng.set_traces(*C.synth_traces(tables, 100, np.random.RandomState(0)))
    eng.reset()
    a = torch.zeros((64, 47), device="cuda")
    for _ in range(60):
        eng.step(a)
    assert eng.counters()["error_or"] & 1
    assert (eng.tensor("error").cpu().numpy() & 1).any()


def test_situation_matrices_match_the_oracle(tables):
    """chub_situation: the stations' `situation` columns (CHS.hpp:204-231) as device matrices, against the oracle's
    pile fields, on mixed hub classes (different pile counts, constant charging) in the middle of a day."""
    import torch
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    classes = [C.DEFAULT_KW, dict(C.DEFAULT_KW, station_list=[7, 30], station_type_list=["slow", "fast"]),
               dict(C.DEFAULT_KW, station_list=[12, 4], constant_charging=True)]
    H, seed = 200, 5
    rs = np.random.RandomState(seed)
    hub_class = rs.randint(0, 3, H).astype(np.int32)
    eng = HubEngine(H, classes, hub_class=hub_class, device=0, diagnostics=True, tables=tables, seed=seed)
    pv, wd, price = C.synth_traces(tables, 50, rs, H)
    eng.set_traces(pv, wd, price, per_hub=True)
    orc = O.Oracle(H, classes, hub_class=hub_class, tables=C.oracle_tables(tables))
    orc.seed_base(seed, 0)
    eng.reset()
    orc.reset(pv[0], wd[0], price[0])
    buf = torch.empty((8, H, 50), dtype=torch.float32, device="cuda")
    for t in range(45):
        a = rs.uniform(-1, 1, (H, eng.action_stride)).astype(np.float32)
        eng.step(torch.from_numpy(a).cuda())
        orc.step(a, pv[t + 1], wd[t + 1], price[t + 1])
        if t % 11 and t != 44:
            continue
        cols, (s0, s1) = eng.situation(out=buf)
        torch.cuda.synchronize()
        assert (s0, s1) == (20, 30)
        car = orc.get("pile_car")
        for name, field in (("car", "pile_car"), ("charge", "pile_charge"), ("emergency", "pile_emergency"),
                            ("power", "pile_power"), ("soc", "pile_soc"), ("init_soc", "pile_arrive"),
                            ("target_soc", "pile_target")):
            want = orc.get(field).astype(np.float32)
            if name != "car":
                want = want * car     # the oracle keeps some members of empty piles; the table reads 0 there
            got = cols[name].cpu().numpy()
            assert got.shape == (H, s0 + s1)
            assert np.array_equal(got[:, :s0], want[:, 0, :s0]), (name, t, 0)
            assert np.array_equal(got[:, s0:], want[:, 1, :s1]), (name, t, 1)
        assert not cols["assign"].any()
        assert car.sum() > H


def _hetero_engines(tables, n_eng, H=3000, seed=11, **kw):
    from charginghub_env_b200 import HubEngine
    from tests.test_gpu_parity import _hetero_classes
    rs = np.random.RandomState(seed)
    classes = _hetero_classes(rs, 12)
    hub_class = rs.randint(0, len(classes), H).astype(np.int32)
    traces = C.synth_traces(tables, 400, rs)
    engs = []
    for _ in range(n_eng):
        e = HubEngine(H, classes, hub_class=hub_class, device=0, tables=tables, seed=seed, action_stride=66, auto_reset=True, **kw)
        e.set_traces(*traces)
        e.reset()
        engs.append(e)
    return engs, rs, H


def _same_outputs(a, b):
    import torch
    torch.cuda.synchronize()
    for name in ("obs", "reward", "done", "stats"):
        assert np.array_equal(a.tensor(name).cpu().numpy(), b.tensor(name).cpu().numpy()), name
    for name in ("pile_soc", "pile_stay", "pile_charge", "st_line", "hub_uctr", "hub_lcg", "store"):
        assert np.array_equal(a.get(name), b.get(name)), name


def test_packed_step_equals_row_step(tables):
    """chub_step_packed on records packed by the host packer = chub_step on the float rows (heterogeneous pile counts,
    actions on both sides of the action_to_real threshold)."""
    import torch
    (a_eng, b_eng), rs, H = _hetero_engines(tables, 2)
    edge = np.array([0.0, -0.0, -2.0 ** -25, -2.0 ** -24, -3e-8, 1.0, -1.0], np.float32)
    for t in range(120):
        a = rs.uniform(-1, 1, (H, 66)).astype(np.float32)
        idx = rs.randint(0, a.size, a.size // 10)
        a.reshape(-1)[idx] = edge[rs.randint(0, edge.size, idx.size)]
        a_eng.step(torch.from_numpy(a).cuda())
        b_eng.step_packed(torch.from_numpy(b_eng.pack_actions(a).view(np.int32)).cuda())
        if t % 24 == 23:
            _same_outputs(a_eng, b_eng)
    assert a_eng.counters()["step_index"] == b_eng.counters()["step_index"] == 120


@pytest.mark.parametrize("threads,ranges", [(0, 0), (1, 1), (3, 5), (8, 32)])
def test_host_pipeline_settings(tables, threads, ranges):
    """chub_step_host with unpacked rows, one thread, odd range counts: the same bits as the device path."""
    import torch
    (a_eng, b_eng), rs, H = _hetero_engines(tables, 2, H=40000 if ranges == 32 else 6000)
    b_eng.set_host_threads(threads, ranges)
    for t in range(30):
        a = rs.uniform(-1, 1, (H, 66)).astype(np.float32)
        o, r, d = a_eng.step(torch.from_numpy(a).cuda())
        oh, rh, dh = b_eng.step_host(a)
        assert np.array_equal(o.cpu().numpy(), oh) and np.array_equal(r.cpu().numpy(), rh) and np.array_equal(d.cpu().numpy(), dh), t
    _same_outputs(a_eng, b_eng)


def test_policy_and_env_in_one_torch_graph(tables):
    """BASELINE configs[4] in small: a torch policy on the DLPack obs view + chub_step + rollout-buffer stores, captured with
    torch.cuda.graph (the engine launches on torch's capture stream) and replayed, equals the same loop run eagerly."""
    import torch
    H, S, T = 2000, 47, 16
    a_eng, rs = _engine(H, tables, seed=9, auto_reset=True)
    b_eng, _ = _engine(H, tables, seed=9, auto_reset=True)
    Wp = torch.from_numpy(rs.normal(0, 0.7, (13, S)).astype(np.float32)).cuda()
    bufs = []

    def segment(eng):
        obs, rew, done = eng.tensor("obs"), eng.tensor("reward"), eng.tensor("done")
        act = torch.empty((T, H, S), device="cuda")
        ob, rb, db = torch.empty((T, H, 13), device="cuda"), torch.empty((T, H), device="cuda"), torch.empty((T, H), device="cuda", dtype=torch.uint8)
        bufs.append((act, ob, rb, db))

        def run():
            for t in range(T):
                ob[t].copy_(obs)
                torch.mm(obs, Wp, out=act[t])
                act[t].tanh_()
                eng.step(act[t])
                rb[t].copy_(rew)
                db[t].copy_(done)
        return run

    run_a, run_b = segment(a_eng), segment(b_eng)
    a_eng.reset()
    b_eng.reset()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        run_b()                       # warm-up of the captured loop's libraries; consumes one segment on b ...
    torch.cuda.current_stream().wait_stream(side)
    run_a()                           # ... so a runs the same segment eagerly
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run_b()
    for rep in range(7):              # 7 x 16 more steps: crosses the day boundary and the auto-reset
        run_a()
        g.replay()
        torch.cuda.synchronize()
        for x, y in zip(bufs[0], bufs[1]):
            assert torch.equal(x, y), rep
    assert a_eng.counters()["step_index"] == b_eng.counters()["step_index"] == 8 * T
    _same_outputs(a_eng, b_eng)


def test_hub_report_equals_field_reads(tables):
    """chub_hub_report (one call) returns what the per-field read-backs return."""
    import torch
    from charginghub_env_b200 import _lib
    eng, rs = _engine(100, tables, seed=4, diagnostics=True)
    eng.reset()
    for t in range(30):
        eng.step(torch.from_numpy(rs.uniform(-1, 1, (100, 47)).astype(np.float32)).cuda())
    for h in (0, 37, 99):
        rep = eng.hub_report(h)
        for name in _lib.DIAG:
            assert rep[name] == float(eng.get("diag_" + name)[h]), name
        for name in ("st_min_power", "st_charge_power", "st_max_power", "st_line", "st_flow_in"):
            assert tuple(rep[name]) == tuple(float(x) if "power" in name else int(x) for x in eng.get(name)[h]), name
        assert rep["t"] == int(eng.get("hub_t")[h]) and rep["store"] == float(eng.get("store")[h])
        assert rep["fcev_arrive"] == int(eng.get("fcev_arrive")[h]) and rep["fcev_qn"] == int(eng.get("fcev_qn")[h])
    with pytest.raises(Exception):
        eng.hub_report(100)


def test_step_into_equals_step(tables):
    """chub_step_into: results written into the caller's tensors are what chub_step leaves in the engine's buffers."""
    import torch
    H = 1500
    a_eng, rs = _engine(H, tables, seed=12, auto_reset=True)
    b_eng, _ = _engine(H, tables, seed=12, auto_reset=True)
    a_eng.reset()
    b_eng.reset()
    T = 100
    ob = torch.empty((T, H, 13), device="cuda")
    rb = torch.empty((T, H), device="cuda")
    db = torch.empty((T, H), device="cuda", dtype=torch.uint8)
    for t in range(T):
        a = torch.from_numpy(rs.uniform(-1, 1, (H, 47)).astype(np.float32)).cuda()
        o, r, d = a_eng.step(a)
        if t % 3 == 0:
            b_eng.step_into(a, ob[t], rb[t], db[t])
        else:   # mixed: only some outputs redirected
            b_eng.step_into(a, ob[t], None, db[t])
            rb[t].copy_(b_eng.tensor("reward"))
        assert torch.equal(o, ob[t]) and torch.equal(r, rb[t]) and torch.equal(d, db[t]), t
    _same_outputs_state = ("pile_soc", "pile_stay", "st_line", "hub_uctr", "hub_lcg", "store")
    for name in _same_outputs_state:
        assert np.array_equal(a_eng.get(name), b_eng.get(name)), name
    with pytest.raises(ValueError):
        b_eng.step_into(a, ob[0][:, :12])


def test_checkpoint_carries_the_last_outputs_and_validates_its_header(tables):
    """a trainer that resumes mid-episode must see the observation it left; a blob of another engine shape
    must be refused by the C ABI itself (header: magic, layout version, hubs, pile slots, classes, generator flag)."""
    import torch
    from charginghub_env_b200 import ChubError, HubEngine
    eng, rs = _engine(200, tables, seed=9)
    eng.reset()
    for _ in range(7):
        eng.step(torch.from_numpy(rs.uniform(-1, 1, (200, 47)).astype(np.float32)).cuda())
    torch.cuda.synchronize()
    want = {n: eng.tensor(n).clone() for n in ("obs", "reward", "done", "final_soc", "stats")}
    blob = eng.get_state()
    for _ in range(5):
        eng.step(torch.from_numpy(rs.uniform(-1, 1, (200, 47)).astype(np.float32)).cuda())
    eng.set_state(blob)
    torch.cuda.synchronize()
    for n, v in want.items():
        assert torch.equal(eng.tensor(n), v), n
    other, _ = _engine(232, tables, seed=9)
    big = other.get_state()
    with pytest.raises((ChubError, ValueError)):
        eng.set_state(big)
    lib = eng.lib
    import ctypes
    bad = blob.copy()
    bad[0] ^= 0xFF                                        # magic
    assert lib.chub_set_state(eng.h, bad.ctypes.data_as(ctypes.c_void_p)) == -1
    small = HubEngine(200, [dict(C.DEFAULT_KW, station_list=[20, 24])], device=0, tables=tables, seed=9)
    sb = small.get_state()
    assert lib.chub_set_state(eng.h, sb.ctypes.data_as(ctypes.c_void_p)) == -1   # other pile layout: refused by the C ABI
    assert b"another shape" in lib.chub_last_error()


def test_wrong_dtype_or_device_is_refused(tables):
    """the kernel would silently reinterpret the bytes of a float64 / half / int64 tensor."""
    import torch
    eng, rs = _engine(64, tables)
    eng.reset()
    for bad in (torch.zeros((64, 47), device="cuda", dtype=torch.float64), torch.zeros((64, 47), device="cuda", dtype=torch.float16),
                torch.zeros((64, 47), device="cuda", dtype=torch.int64)):
        with pytest.raises(ValueError):
            eng.step(bad)
    with pytest.raises(ValueError):
        eng.step_loads(torch.zeros((64, 4), device="cuda", dtype=torch.float64))
    with pytest.raises(ValueError):
        eng.reset(mask=torch.zeros(64, device="cuda", dtype=torch.float32))
    with pytest.raises(ValueError):
        eng.reset(mask=torch.zeros(63, device="cuda", dtype=torch.uint8))
    with pytest.raises(ValueError):
        eng.step(torch.zeros((64, 47)))                   # host tensor
    eng.step(torch.zeros((64, 47), device="cuda"))       # and the right one still works


def test_setters_invalidate_a_captured_rollout(tables):
    """a captured graph bakes trace / pile pointers; replaying it after a setter freed them must fail cleanly."""
    import torch
    from charginghub_env_b200 import ChubError
    eng, rs = _engine(256, tables, seed=4, auto_reset=True)
    acts = torch.from_numpy(rs.uniform(-1, 1, (256, 47)).astype(np.float32)).cuda()
    eng.reset()
    eng.capture_rollout(acts, 4)
    eng.rollout()
    eng.set_traces(*C.synth_traces(tables, 96, rs))
    with pytest.raises(ChubError):
        eng.rollout()
    eng.capture_rollout(acts, 4)
    eng.rollout()
    torch.cuda.synchronize()
    assert eng.counters()["step_index"] == 8


def test_per_launch_timers(tables):
    import torch
    eng, rs = _engine(4096, tables, seed=4)
    assert eng.unit_hubs() % 32 == 0 and 128 <= eng.unit_hubs() <= 1024   # whole tiles, one per warp of a block
    acts = torch.from_numpy(rs.uniform(-1, 1, (4096, 47)).astype(np.float32)).cuda()
    eng.reset()
    eng.set_timing(True)
    for _ in range(10):
        eng.step(acts)
    ms = eng.step_times()
    assert ms.shape == (10,) and (ms > 0).all() and (ms < 50).all()
    eng.set_timing(False)
    eng.step(acts)
    assert eng.step_times().shape == (0,)


def test_gather_and_reduce_stats_through_the_c_abi(tables):
    """chub_gather_stats / chub_reduce_stats on a one-rank NCCL communicator (torch's): the root receives its own
    statistics, the sums equal a float64 sum of the buffer.  (The multi-rank form runs in bench.py --gpus N.)"""
    import os
    import torch
    import torch.distributed as dist
    from charginghub_env_b200 import sharding
    eng, rs = _engine(3000, tables, seed=6, auto_reset=True)
    acts = torch.from_numpy(rs.uniform(-1, 1, (3000, 47)).astype(np.float32)).cuda()
    eng.reset()
    for _ in range(97):
        eng.step(acts)
    torch.cuda.synchronize()
    last = eng.tensor("last_stats")
    assert (last[:, 0] != 0).all()
    sums = eng.reduce_stats()                              # no communicator: local sums
    torch.cuda.synchronize()
    assert torch.allclose(sums, last.double().sum(0), rtol=1e-12)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29613")
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    try:
        dist.all_reduce(torch.zeros(1, device="cuda"))     # torch creates the communicator on first use
        comm = sharding.nccl_comm_ptr(0)
        assert comm
        out = torch.zeros((3000, 4), device="cuda")
        eng.gather_stats(comm, 0, None, out)
        eng.step(acts)                                      # overlaps the transfer
        eng.gather_wait()
        torch.cuda.synchronize()
        assert torch.equal(out, last)
        sums2 = eng.reduce_stats(comm)
        torch.cuda.synchronize()
        assert torch.equal(sums2, sums)
        got = sharding.gather_engine_stats(eng)
        assert torch.equal(got, last)
    finally:
        if created:
            dist.destroy_process_group()
