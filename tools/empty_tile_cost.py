"""What a tile costs when its stations are EMPTY: the arrival CDF is replaced by one that always draws zero cars, the cars
of reset() leave within a day, and what remains of a step is its fixed part -- action rows, hub record, the hydrogen /
reward / observation section -- plus the launch itself.  Prints the per-launch kernel time of the second day."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from charginghub_env_b200 import HubEngine, load_tables

H = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
tabs = dict(load_tables())
tabs["arrival_cdf"] = np.ones_like(tabs["arrival_cdf"])          # u <= cdf[0] always: N = 0 cars per slot
eng = HubEngine(H, [bench.DEFAULT_KW], action_stride=47, auto_reset=False, tables=tabs, seed=3)
eng.set_traces(*bench.year_traces(tabs, 35040))
acts = torch.rand((H, 47), device="cuda") * 2 - 1
eng.reset()
for _ in range(192):
    eng.step(acts)
occ = eng.get("pile_car").sum()
eng.set_timing(True)
for _ in range(48):
    eng.step(acts)
ms = eng.step_times(128)[-48:]
print("hubs %d, cars left %d: empty-station step %.4f ms (min %.4f max %.4f), error bits %d"
      % (H, int(occ), float(ms.mean()), float(ms.min()), float(ms.max()), eng.counters()["error_or"]))
