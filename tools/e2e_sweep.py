"""chub_step_host timing over host-thread counts and range counts (config-5 shard).  Prints hub-steps/s per setting."""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from charginghub_env_b200 import HubEngine, load_tables

ap = argparse.ArgumentParser()
ap.add_argument("--hubs", type=int, default=524288)
ap.add_argument("--threads", default="0,1,4,8,12,15")
ap.add_argument("--ranges", default="4,8,16")
ap.add_argument("--steps", type=int, default=24)
a = ap.parse_args()
tabs = load_tables()
wl = bench.workload("config5", tabs)
H = a.hubs
eng = HubEngine(H, wl["classes"], action_stride=wl["stride"], auto_reset=True, tables=tabs, seed=1)
eng.set_traces(*bench.year_traces(tabs, 35040))
eng.reset()
hb = eng.host_buffers()
bufs = [hb["actions"], eng.pinned_actions(), eng.pinned_actions()]   # rotated: 296 MB, beyond the host's caches
for k, b in enumerate(bufs):
    b[...] = np.random.RandomState(k).uniform(-1, 1, b.shape).astype(np.float32)
print("cores", os.cpu_count())
for th in [int(x) for x in a.threads.split(",")]:
    for rg in [int(x) for x in a.ranges.split(",")]:
        eng.set_host_threads(th, rg)
        for k in range(3):
            eng.step_host(bufs[k % 3], copy=False)
        t0 = time.perf_counter()
        for k in range(a.steps):
            o, r, d = eng.step_host(bufs[k % 3], copy=False)
            _ = float(r[0])
        dt = (time.perf_counter() - t0) / a.steps
        print("threads %2d ranges %2d: %.3f ms/step  %.4g hub-steps/s" % (th, rg, dt * 1e3, H / dt), flush=True)
# the packing alone
pk = np.empty((H, 4), np.uint32)
for _ in range(2):
    eng.lib.chub_pack_actions(eng.h, hb["actions"].ctypes.data, pk.ctypes.data)
t0 = time.perf_counter()
for _ in range(5):
    eng.lib.chub_pack_actions(eng.h, hb["actions"].ctypes.data, pk.ctypes.data)
print("pack on one thread: %.3f ms" % ((time.perf_counter() - t0) / 5 * 1e3))
