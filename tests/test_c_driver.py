"""The C ABI used from plain C: include/chub.h compiles as C99, a C driver (tests/c/abi_driver.c) builds against
libchub.so with gcc alone, and -- on a GPU -- reproduces what the Python binding computes, bit for bit."""
import os
import struct
import subprocess

import numpy as np
import pytest

from tests import common as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "charginghub_env_b200", "csrc")


def build_driver(tmp):
    exe = os.path.join(tmp, "abi_driver")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", "abi_driver.c"), "-o", exe, "-L", CSRC, "-lchub",
                           "-Wl,-rpath," + CSRC])
    return exe


def test_header_is_c99_and_driver_links(tmp_path):
    from charginghub_env_b200 import _lib
    _lib.load()
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c",
                           os.path.join(ROOT, "include", "chub.h")])
    exe = build_driver(str(tmp_path))
    assert subprocess.call([exe]) == 1   # no arguments: usage error, but the loader resolved every symbol it uses


@pytest.mark.gpu
def test_c_driver_equals_python_binding(tables, tmp_path):
    import torch
    from charginghub_env_b200 import HubEngine
    H, T, L, n0, n1 = 5000, 100, 200, 20, 25
    rs = np.random.RandomState(77)
    pv, wd, price = C.synth_traces(tables, L, rs)
    acts = rs.uniform(-1, 1, (T, H, n0 + n1 + 2)).astype(np.float32)
    seed = 0xC0FFEE
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("<5i", H, T, L, n0, n1))
        for a in (tables["arrival_cdf"], tables["price"], np.array([tables["price_mean"], tables["price_std"]]),
                  tables["time_sin"], pv, wd, price):
            f.write(np.ascontiguousarray(a, np.float64).tobytes())
        f.write(struct.pack("<Q", seed))
        f.write(acts.tobytes())
    exe = build_driver(str(tmp_path))
    out = subprocess.run([exe, fin, fout], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert out.stdout.startswith("steps %d " % T) and out.stdout.strip().endswith("error_bits 0")
    raw = np.fromfile(fout, np.uint8)
    eng = HubEngine(H, [C.DEFAULT_KW], device=0, tables=tables, seed=seed, auto_reset=True, diagnostics=False)
    eng.set_traces(pv, wd, price)
    obs = eng.reset()
    torch.cuda.synchronize()
    off = 0

    def take(n, dt):
        nonlocal off
        a = raw[off:off + n * np.dtype(dt).itemsize].view(dt)
        off += a.nbytes
        return a

    assert np.array_equal(take(H * 13, np.float32).reshape(H, 13), obs.cpu().numpy())
    for t in range(T):
        o, r, d = eng.step(torch.from_numpy(acts[t]).cuda())
        assert np.array_equal(take(H * 13, np.float32).reshape(H, 13), o.cpu().numpy()), t
        assert np.array_equal(take(H, np.float32), r.cpu().numpy()), t
        assert np.array_equal(take(H, np.uint8), d.cpu().numpy()), t
    rep = take(29, np.float64)
    assert rep[24] == eng.get("hub_t")[H - 1] and rep[25] == eng.get("store")[H - 1]
    assert off == raw.size
