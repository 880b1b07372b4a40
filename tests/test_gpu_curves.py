"""-m gpu: the charging curves AS THE STEP KERNEL COMPILES THEM against glibc, on EVERY reachable float32 argument.

Integer state hangs on these functions -- stay_time = ceil(s2t(target) - s2t(soc)) + late (CHS.hpp:933-937, 1098-1102),
urgency from time_left <= ceil(needed) (CHS.hpp:885, 1050) -- and the device evaluates them with fma-Horner chains and
its own log / exp / divide (csrc/chub_device.cuh, csrc/chub_fastmath.h) where the reference calls glibc pow / exp / log
(CHS.hpp:551-588, 684-725).  A sample cannot prove the fp32 results equal; the reachable domains are small enough to
enumerate, so this test enumerates them through chub_debug_curve (the kernel's own device functions) and compares every
value with oracle/hub_oracle.c's restatement, which tests/test_oracle_vs_reference.py pins on the compiled reference.

Reachable arguments: an EV arrives with SOC = 75 - 5 clamp(N(7,3), 1, 10) in [25, 70] (CHS.hpp:804-814) and charges to at
most 100, so SOC lies in [25, 100] (swept: [16, 100]); charge time is s2t(SOC) + 1 <= 15.68 + 1 slow, <= 3.42 + 1 fast
(swept: [1, 16) and [1, 4.5)), and add_car's initial power is read at s2t(arrival SOC) (swept from 0.5 up).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CHUNK = 1 << 22


def _bits(x):
    return int(np.float32(x).view(np.uint32))


def _sweep(eng, which_dev, which_orc, lo, hi, cp=False):
    """Every float32 in [lo, hi]: device vs oracle, bit patterns compared.  Returns (count, list of mismatches)."""
    import torch
    from oracle import oracle as O
    b0, b1 = _bits(lo), _bits(hi) + 1
    bad, n_all = [], 0
    out = torch.empty(CHUNK, dtype=torch.float32, device="cuda")
    for first in range(b0, b1, CHUNK):
        n = min(CHUNK, b1 - first)
        eng.debug_curve(which_dev + (8 if cp else 0), first, n, out=out)
        x = (np.arange(n, dtype=np.uint32) + np.uint32(first)).view(np.float32)
        want = O.curve_batch(which_orc, x, cp)           # overlaps the kernel
        got = out[:n].cpu().numpy()
        diff = np.nonzero(got.view(np.uint32) != want.view(np.uint32))[0]
        for i in diff[:20]:
            bad.append((float(x[i]), float(got[i]), float(want[i])))
        n_all += n
    return n_all, bad


# (device function, oracle function, lo, hi): see include/chub.h chub_debug_curve for the device numbering
SWEEPS = [
    ("slow car_step soc", 1, 1, 1.0, 16.0), ("slow car_step power", 6, 0, 1.0, 16.0),
    ("slow add_car power", 0, 0, 0.5, 16.0), ("slow soc_to_time", 2, 2, 16.0, 100.0),
    ("fast car_step soc", 4, 4, 1.0, 4.5), ("fast car_step power", 7, 3, 1.0, 4.5),
    ("fast add_car power", 3, 3, 0.5, 4.5), ("fast soc_to_time", 5, 5, 16.0, 100.0),
]


@pytest.mark.parametrize("name,dev,orc,lo,hi", SWEEPS)
def test_every_reachable_argument_matches_glibc(tables, name, dev, orc, lo, hi):
    from charginghub_env_b200 import HubEngine
    from tests import common as C
    eng = HubEngine(32, [C.DEFAULT_KW], device=0, tables=tables, seed=1)
    n, bad = _sweep(eng, dev, orc, lo, hi)
    print("%s: %d arguments in [%g, %g], %d differ" % (name, n, lo, hi, len(bad)), bad[:8])
    assert not bad, (name, bad[:8])


@pytest.mark.parametrize("name,dev,orc,lo,hi", SWEEPS)
def test_constant_power_variants(tables, name, dev, orc, lo, hi):
    """constant_charging=True (CHS.hpp:496-501, 514-521, 528-535, 622-627, 640-647, 662-669), same domains."""
    from charginghub_env_b200 import HubEngine
    from tests import common as C
    eng = HubEngine(32, [C.DEFAULT_KW], device=0, tables=tables, seed=1)
    n, bad = _sweep(eng, dev, orc, lo, hi, cp=True)
    assert not bad, (name, bad[:8])


def test_target_table_matches_glibc(tables):
    """s2t(target) for the 1000 possible targets uniform_rand(80, 100) = r / 999.f * 20 + 80 (CHS.hpp:41-42, 867): the
    device builds this table with its own curve code; compare through the sweep entry on exactly those arguments."""
    from charginghub_env_b200 import HubEngine
    from oracle import oracle as O
    from tests import common as C
    eng = HubEngine(32, [C.DEFAULT_KW], device=0, tables=tables, seed=1)
    r = np.arange(1000, dtype=np.float32)
    tr = (r / np.float32(999.0)).astype(np.float32)
    target = (tr * np.float32(20.0) + np.float32(80.0)).astype(np.float32)
    for dev, orc in ((2, 2), (5, 5)):
        for cp in (False, True):
            want = O.curve_batch(orc, target, cp)
            got = np.array([float(eng.debug_curve(dev + (8 if cp else 0), _bits(t), 1).cpu()[0]) for t in target[::37]],
                           np.float32)
            assert np.array_equal(got.view(np.uint32), want[::37].view(np.uint32)), (dev, cp)
