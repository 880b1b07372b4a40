"""CPU: host-side logic of the package that needs no GPU -- shard arithmetic, stream derivation,
hub-class marshalling, packaged tables."""
import numpy as np

from charginghub_env_b200 import make_hub_class, sharding
from oracle import oracle as O
from tests import common as C


def test_shard_ranges_partition_the_batch():
    for n, w in [(1048576, 8), (10, 4), (7, 8), (4194304, 8), (5, 1)]:
        edges = [sharding.shard_range(n, r, w) for r in range(w)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        assert all(edges[r][1] == edges[r + 1][0] for r in range(w - 1))
        sizes = [b - a for a, b in edges]
        assert max(sizes) - min(sizes) <= 1


def test_hub_orders_are_permutations_that_group_classes_and_sizes():
    rs = np.random.RandomState(5)
    classes = [dict(station_list=[int(rs.randint(4, 33)), int(rs.randint(4, 33))]) for _ in range(40)]
    hc = rs.randint(0, 40, 5000)
    for order, inverse in (sharding.order_by_class(hc), sharding.order_by_size(classes, hc)):
        assert sorted(order.tolist()) == list(range(5000))
        assert (hc[order][inverse] == hc).all()
        runs = 1 + int((np.diff(hc[order]) != 0).sum())
        assert runs == len(set(hc.tolist()))                      # every class is one contiguous run
    size = np.array([sum(c["station_list"]) for c in classes])[hc[sharding.order_by_size(classes, hc)[0]]]
    assert (np.diff(size) >= 0).all()                             # and the runs come in order of pile count


def test_python_stream_mirror_equals_c_header():
    """sharding.hub_streams (used to seed explicit streams) == chub_hub_streams() in include/chub_rng.h."""
    ids = np.array([0, 1, 2, 4095, 4096, 1048575, 4194303], np.int64)
    k0, k1, e0 = sharding.hub_streams(0x5EED, ids)
    for j, h in enumerate(ids):
        orc = O.Oracle(1, [C.DEFAULT_KW])
        orc.seed_base(0x5EED, int(h))
        assert np.uint32(orc.get("hub_lcg")[0]) == e0[j]
        # keys are not readable back, so check them through the first U draw: step-free, via reset draws
    assert len(set(e0.tolist())) == len(ids) and (e0 >= 1).all() and (e0 <= 2147483646).all()
    assert len(set(zip(k0.tolist(), k1.tolist()))) == len(ids)


def test_streams_do_not_depend_on_sharding(tables):
    """Hub 5 of a 1-shard batch == hub 1 of the shard that starts at global id 4."""
    rs = np.random.RandomState(3)
    pv, wd, pr = C.synth_traces(tables, 9, rs)
    a = rs.uniform(-1, 1, (8, 47)).astype(np.float32)
    whole = O.Oracle(8, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    whole.seed_base(7, 0)
    part = O.Oracle(4, [C.DEFAULT_KW], tables=C.oracle_tables(tables))
    part.seed_base(7, 4)
    whole.reset(pv[0], wd[0], pr[0])
    part.reset(pv[0], wd[0], pr[0])
    for t in range(8):
        whole.step(a, pv[t + 1], wd[t + 1], pr[t + 1])
        part.step(a[4:], pv[t + 1], wd[t + 1], pr[t + 1])
    for name in ("pile_soc", "pile_stay", "st_line", "reward", "store"):
        assert np.array_equal(whole.get(name)[4:], part.get(name)), name


def test_hub_class_marshalling():
    c, keep = make_hub_class(**C.DEFAULT_KW)
    assert (c.n_piles[0], c.n_piles[1]) == (20, 25) and (c.station_type[0], c.station_type[1]) == (0, 1)
    assert c.hydro_prod_rate == 100.0 and c.hydro_store_vlt == 25.0 and c.init_soc == 0.2 and keep is None
    c, _ = make_hub_class(station_list=[3, 4], station_type_list=["slow", "slow"])
    assert c.hydro_prod_rate < 0 and c.hydro_store_vlt < 0 and c.fc_max_power < 0  # None -> engine defaults
    tab = np.arange(102, dtype=np.float64)
    c, keep = make_hub_class(el_table=tab)
    assert keep is not None and c.el_table[101] == 101.0


def test_packaged_tables(tables):
    assert tables["arrival_cdf"].shape == (96, 301) and tables["price"].shape == (96,)
    assert tables["pv_days"].shape == (100, 96) and tables["wd_days"].shape == (150, 96)
    cdf = tables["arrival_cdf"]
    assert (np.diff(cdf, axis=1) >= -1e-6).all() and cdf.max() <= 1.0 + 1e-6  # rows are CDFs
    assert abs(tables["price_mean"] - tables["price"].mean()) < 1e-15


# ---- arithmetic facts the step kernel relies on (charginghub_env_b200/csrc/chub_kernels.cu) ----

def test_action_binarisation_is_a_single_compare():
    """action_to_real on float32: on iff (a + 1) / 2 >= 0.5 (evcssp_manager.py:384-404).  The kernel tests
    a >= -2^-25 instead; the two agree on every float32, checked here around the boundary and on specials."""
    edge = np.float32(-2.0 ** -25)
    xs = [edge]
    for _ in range(64):
        xs.append(np.nextafter(xs[-1], np.float32(-1), dtype=np.float32))
    up = edge
    for _ in range(64):
        up = np.nextafter(up, np.float32(1), dtype=np.float32)
        xs.append(up)
    xs += [np.float32(v) for v in (0.0, -0.0, 1.0, -1.0, 1e-30, -1e-30, 3e38, -3e38, np.inf, -np.inf, np.nan, -1e-8, -3e-8)]
    rs = np.random.RandomState(0)
    xs = np.concatenate([np.array(xs, np.float32), rs.uniform(-1, 1, 200000).astype(np.float32),
                         (rs.uniform(-1, 1, 200000) * 1e-7).astype(np.float32)])
    with np.errstate(invalid="ignore", over="ignore"):
        ref = ((xs + np.float32(1.0)) / np.float32(2.0)) >= np.float32(0.5)
        fast = xs >= np.float32(-2.0 ** -25)
    assert np.array_equal(ref, fast)


def test_curve_breakpoints_are_not_float32_values():
    """The curves' segment tests mix '<' and '<=' against double constants (CHS.hpp:503-509, 559-578, 629-635,
    649-657).  None of the constants is a float32 value, so both comparisons pick the same side for every float32
    time and one segment index decides soc and power together (slow_seg / fast_seg in chub_device.cuh)."""
    for v in (2.33 * 4, 3.67 * 4, 28.7 / 15, 51.2 / 15):
        assert float(np.float32(v)) != v


def test_charging_power_is_never_negative():
    """The sign bit of the posted pile power carries the urgency flag to the station sums, which is only sound
    if no curve ever returns a negative power (or -0.0).  Checked on the oracle's curves over every time the
    step can produce: t = s2t(soc) + 1 in [1, t_max + 1], and t = s2t(arrival soc) for add_car."""
    for which, tmax in ((0, 14.68), (3, 51.2 / 15)):
        t = np.linspace(0.0, tmax + 1.5, 2000001).astype(np.float32)
        for cp in (False, True):
            p = O.curve_batch(which, t, constant_power=cp)
            assert not np.signbit(p).any() and np.isfinite(p).all()


def test_minstd_jump_ahead_constants():
    """chub_minstd_jump(x, A_k) == k plain steps (include/chub_rng.h), checked through a pure-Python mirror."""
    m, a = 2147483647, 16807
    assert (a * a) % m == 282475249 and pow(a, 3, m) == 1622650073 and pow(a, 4, m) == 984943658


def test_segment_thresholds_as_float32():
    """slow_seg / fast_seg compare the float time with a float constant; it must order every float exactly as the
    reference's double comparison does (CHUB_*_SEG_THR in csrc/chub_device.cuh)."""
    import os, re
    src = open(os.path.join(os.path.dirname(__file__), "..", "charginghub_env_b200", "csrc", "chub_device.cuh")).read()
    for name, dbl in (("CHUB_SLOW_SEG_THR", 2.33 * 4), ("CHUB_FAST_SEG_THR", 28.7 / 15)):
        lit = re.search(r"#define %s (\S+)f\b" % name, src).group(1)
        thr = np.float32(float.fromhex(lit))
        assert float(thr) == float.fromhex(lit)           # the literal is a float32 value
        below = np.nextafter(thr, np.float32(-np.inf))
        assert float(thr) >= dbl and float(below) < dbl     # smallest float not below the double
