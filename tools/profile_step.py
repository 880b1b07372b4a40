"""Minimal driver for ncu: N hubs, a few warm-up steps past the day-0 transient, then a few steps."""
import argparse
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from charginghub_env_b200 import HubEngine, load_tables

ap = argparse.ArgumentParser()
ap.add_argument("--hubs", type=int, default=524288)
ap.add_argument("--warm", type=int, default=40)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--workload", default="config5")
ap.add_argument("--control", default="piles", choices=["piles", "loads"])
ap.add_argument("--phase", default="lockstep", choices=["lockstep", "spread", "spread-warps", "spread-quarters"],
                help="spread: de-phase the batch first (bench.dephase) -- every launch then holds every time of day")
a = ap.parse_args()
tabs = load_tables()
wl = bench.workload(a.workload, tabs)
H = a.hubs
hc = None if wl["hub_class"] is None else np.resize(wl["hub_class"], H)
eng = HubEngine(H, wl["classes"], hub_class=hc, action_stride=wl["stride"], auto_reset=True, tables=tabs, seed=1)
eng.set_traces(*bench.year_traces(tabs, 35040))
acts = torch.rand((H, wl["stride"]), device="cuda") * 2 - 1
eng.reset()
if a.control == "loads":   # one power target per station: [load0 kW, load1 kW, a[-2], a[-1]]
    caps = torch.tensor([[40.0 * n for n in wl["classes"][0]["station_list"]]], device="cuda")
    acts = torch.cat([torch.rand((H, 2), device="cuda") ** 2 * caps, torch.rand((H, 2), device="cuda") * 2 - 1], dim=1).contiguous()
if a.phase != "lockstep":
    bench.dephase(eng, H, eng.step_loads if a.control == "loads" else eng.step, [acts], a.phase)
for _ in range(a.warm + a.steps):
    (eng.step_loads if a.control == "loads" else eng.step)(acts)
torch.cuda.synchronize()
print("ok", eng.counters())
