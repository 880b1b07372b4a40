"""Per-step duration of the step kernel over one simulated day (the load varies with the time of day)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
from charginghub_env_b200 import HubEngine, load_tables
tabs = load_tables()
wl = bench.workload("config5", tabs)
H = wl["H"]
eng = HubEngine(H, wl["classes"], action_stride=wl["stride"], auto_reset=True, tables=tabs, seed=1)
eng.set_traces(*bench.year_traces(tabs, 35040))
pool = [torch.rand((H, wl["stride"]), device="cuda") * 2 - 1 for _ in range(4)]
eng.reset()
for i in range(96):
    eng.step(pool[i % 4])
ev = [torch.cuda.Event(enable_timing=True) for _ in range(97)]
torch.cuda.synchronize()
ev[0].record()
for i in range(96):
    eng.step(pool[i % 4])
    ev[i + 1].record()
torch.cuda.synchronize()
ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(96)]
print("step_ms " + " ".join("%.3f" % x for x in ms))
print("mean %.4f min %.4f max %.4f" % (np.mean(ms), np.min(ms), np.max(ms)))
