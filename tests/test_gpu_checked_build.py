"""-m gpu: the parity scenarios through the SELF-CHECKING build (libchub_checked.so = the same sources with -DCHUB_CHECKS).

compute-sanitizer is closed on the pool this engine is developed on (profiles/r02_compute_sanitizer_closed_on_this_pool.txt),
so the kernels carry their own bounds assertions for everything the step kernel indexes by hand -- the per-warp car_step
rings, the add_car slots, the posted-power rows, the warp's pile tile, the staged action rows, the unit queue -- counted in
a device counter (include/chub.h chub_checks).  A subprocess loads that build (CHUB_LIB), steps heterogeneous hubs (all
station orders, constant power, permeate up to 0.05, ragged batch) with per-pile and scalar-load actions through auto-reset
episodes, the graph rollout and the host pipeline, compares with the oracle, and must report zero violations."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r'''
import numpy as np, torch, sys
sys.path.insert(0, %r)
from charginghub_env_b200 import HubEngine, load_tables
from oracle import oracle as O
from tests import common as C
from tests.test_gpu_parity import _hetero_classes
tabs = load_tables()
rs = np.random.RandomState(5)
classes = _hetero_classes(rs, 20)
H, T = 1000 - 7, 2 * 96 + 8
hub_class = rs.randint(0, len(classes), H).astype(np.int32)
eng = HubEngine(H, classes, hub_class=hub_class, device=0, diagnostics=True, tables=tabs, seed=99, auto_reset=True, action_stride=66)
compiled, fails = eng.checks()
assert compiled, "not the checked build"
pv, wd, price = C.synth_traces(tabs, T + 1, rs, H)
eng.set_traces(pv, wd, price, per_hub=True)
orc = O.Oracle(H, classes, hub_class=hub_class, tables=C.oracle_tables(tabs))
orc.seed_base(99, 0)
eng.reset(); orc.reset(pv[0], wd[0], price[0])
caps = np.array([[40.0 * n for n in classes[k]["station_list"]] for k in hub_class], np.float32)
mm = C.Mismatch()
for t in range(T):
    if t %% 4 == 3:
        rows = np.concatenate([rs.uniform(0, 1, (H, 2)).astype(np.float32) ** 2 * caps, rs.uniform(-1, 1, (H, 2)).astype(np.float32)], axis=1)
        obs, rew, done = eng.step_loads(torch.from_numpy(rows).cuda()); orc.set_load_mode(True)
    elif t %% 4 == 1:
        rows = rs.uniform(-1, 1, (H, 66)).astype(np.float32)
        obs, rew, done = (torch.from_numpy(x.copy()) for x in eng.step_host(rows)); orc.set_load_mode(False)
    else:
        rows = rs.uniform(-1, 1, (H, 66)).astype(np.float32)
        obs, rew, done = eng.step(torch.from_numpy(rows).cuda()); orc.set_load_mode(False)
    orc.step(rows, pv[t + 1], wd[t + 1], price[t + 1])
    if orc.get("hub_done").all():
        orc.reset(pv[t + 1], wd[t + 1], price[t + 1])
    if t %% 24 == 23 or t == T - 1:
        torch.cuda.synchronize()
        C.compare_state(eng, orc, 0, mm, obs=obs.cpu().numpy())
assert not mm, str(mm)
acts = torch.from_numpy(rs.uniform(-1, 1, (H, 66)).astype(np.float32)).cuda()
eng.capture_rollout(acts, 8); eng.rollout(); eng.rollout()
big = HubEngine(200000, [C.DEFAULT_KW], device=0, tables=tabs, seed=3, auto_reset=True)      # the persistent queue with many units
big.set_traces(*C.synth_traces(tabs, 192, rs))
big.reset()
a = torch.rand((200000, 47), device="cuda") * 2 - 1
for t in range(100):
    big.step(a)
compiled, fails = eng.checks()
print("checked build: violations", fails)
assert fails == 0
'''


def test_checked_build_reports_no_violation():
    lib = os.path.join(ROOT, "charginghub_env_b200", "csrc", "libchub_checked.so")
    assert os.path.exists(lib), "build it: python __graft_entry__.py"
    env = dict(os.environ, CHUB_LIB=lib)
    out = subprocess.run([sys.executable, "-c", SCRIPT % ROOT], capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "checked build: violations 0" in out.stdout
