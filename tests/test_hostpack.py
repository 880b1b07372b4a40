"""CPU: the host-side action packer of chub_step_host (csrc/chub_hostpack.cpp) against action_to_real
(evcssp_manager.py:384-404) restated in numpy float32 -- boundary values, NaN, narrow and ragged rows, and the thread pool."""
import ctypes

import numpy as np
import pytest


def action_to_real_bits(rows, na):
    """on iff (a + 1)/2 >= 0.5 evaluated in float32, as the reference does on its float32 action vector."""
    a = rows[:, :na].astype(np.float32)
    half = (a + np.float32(1.0)) / np.float32(2.0)
    with np.errstate(invalid="ignore"):
        return half >= np.float32(0.5)


def expected(rows, na_row):
    n = rows.shape[0]
    out = np.zeros((n, 4), np.uint32)
    for r in range(n):
        na = int(na_row[r])
        on = action_to_real_bits(rows[r:r + 1], na)[0]
        bits = 0
        for j in np.nonzero(on)[0]:
            bits |= 1 << int(j)
        out[r, 0] = bits & 0xFFFFFFFF
        out[r, 1] = bits >> 32
        out[r, 2:4] = rows[r, na:na + 2].view(np.uint32)
    return out


def special_rows(rs, n, stride):
    rows = rs.uniform(-1, 1, (n, stride)).astype(np.float32)
    edge = np.array([0.0, -0.0, -2.0 ** -25, np.nextafter(np.float32(-2.0 ** -25), np.float32(-1)),
                     np.nextafter(np.float32(-2.0 ** -25), np.float32(0)), -2.0 ** -24, np.nan, np.inf, -np.inf,
                     -1e-38, 1e-45, -1e-45, -1.0, 1.0], np.float32)
    idx = rs.randint(0, rows.size, rows.size // 5)
    rows.reshape(-1)[idx] = edge[rs.randint(0, edge.size, idx.size)]
    return rows


def lib():
    from charginghub_env_b200 import _lib
    return _lib.load()


@pytest.mark.parametrize("stride,na", [(47, 45), (66, 64), (66, 8), (10, 8), (9, 7), (3, 1), (34, 32), (40, 33)])
def test_pack_rows_uniform(stride, na):
    rs = np.random.RandomState(stride * 100 + na)
    rows = special_rows(rs, 257, stride)
    out = np.zeros((257, 4), np.uint32)
    lib().chub_pack_rows(rows.ctypes.data, stride, None, na, 257, out.ctypes.data)
    assert np.array_equal(out, expected(rows, np.full(257, na)))


def test_pack_rows_ragged_classes():
    rs = np.random.RandomState(7)
    stride = 66
    rows = special_rows(rs, 1000, stride)
    na_row = rs.randint(1, 65, 1000).astype(np.int32)
    out = np.zeros((1000, 4), np.uint32)
    lib().chub_pack_rows(rows.ctypes.data, stride, na_row.ctypes.data, 0, 1000, out.ctypes.data)
    assert np.array_equal(out, expected(rows, na_row))


def test_pack_rows_empty_and_exact_end():
    rows = np.full((1, 10), 0.5, np.float32)   # the row ends where the buffer ends: no read beyond na + 2 entries
    out = np.full((2, 4), 0xDEADBEEF, np.uint32)
    lib().chub_pack_rows(rows.ctypes.data, 10, None, 8, 0, out.ctypes.data)
    assert (out == 0xDEADBEEF).all()
    lib().chub_pack_rows(rows.ctypes.data, 10, None, 8, 1, out.ctypes.data)
    assert out[0, 0] == 0xFF and out[0, 1] == 0 and (out[1] == 0xDEADBEEF).all()


@pytest.mark.parametrize("threads", [1, 3, 8])
def test_pool_packs_every_range(threads):
    L = ctypes.CDLL(lib()._name)
    L.chub_pack_pool_create.restype = ctypes.c_void_p
    L.chub_pack_pool_create.argtypes = [ctypes.c_int32]
    L.chub_pack_pool_destroy.argtypes = [ctypes.c_void_p]
    L.chub_pack_pool_start.restype = ctypes.c_int32
    L.chub_pack_pool_start.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32,
                                       ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]
    L.chub_pack_pool_wait.argtypes = [ctypes.c_void_p, ctypes.c_int32]
    pool = L.chub_pack_pool_create(threads)
    rs = np.random.RandomState(threads)
    try:
        for n, per in [(5000, 1024), (5, 1024), (4096, 1024), (33, 1)]:
            rows = special_rows(rs, n, 47)
            out = np.zeros((n, 4), np.uint32)
            nr = L.chub_pack_pool_start(pool, rows.ctypes.data, 47, None, 45, n, per, out.ctypes.data)
            if -(-n // per) > 32:
                assert nr == -1
                continue
            assert nr == -(-n // per)
            for r in range(nr):
                L.chub_pack_pool_wait(pool, r)
                lo, hi = r * per, min(n, (r + 1) * per)
                assert np.array_equal(out[lo:hi], expected(rows[lo:hi], np.full(hi - lo, 45))), (n, per, r)
        for _ in range(200):   # back-to-back jobs: the workers never see a half-written job
            rows = special_rows(rs, 64, 47)
            out = np.zeros((64, 4), np.uint32)
            nr = L.chub_pack_pool_start(pool, rows.ctypes.data, 47, None, 45, 64, 16, out.ctypes.data)
            for r in range(nr):   # a wait covers its own range only
                L.chub_pack_pool_wait(pool, r)
            assert np.array_equal(out, expected(rows, np.full(64, 45)))
    finally:
        L.chub_pack_pool_destroy(pool)
