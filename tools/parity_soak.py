"""Parity soak: a large batch stepped for many days on the GPU and by the CPU checker (oracle/, one instance per host
thread), every integer field and the fp32 pile state compared every few steps.  Hunts the one-in-10^7 events the unit
tests are too small to meet.  usage: parity_soak.py [--hubs 16384] [--days 10] [--classes 0|N] [--seed S] [--every 8]"""
import argparse, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from charginghub_env_b200 import HubEngine, load_tables
from oracle import oracle as O
from tests import common as C
from tests.test_gpu_parity import _hetero_classes

ap = argparse.ArgumentParser()
ap.add_argument("--hubs", type=int, default=16384)
ap.add_argument("--days", type=int, default=10)
ap.add_argument("--classes", type=int, default=0, help="0: default hubs; N: N random heterogeneous classes")
ap.add_argument("--seed", type=int, default=2024)
ap.add_argument("--every", type=int, default=8)
ap.add_argument("--threads", type=int, default=os.cpu_count() or 4)
ap.add_argument("--exogenous", default="traces", choices=["traces", "device"],
                help="device: the on-device PV / wind / price generator against oracle/exo_oracle.c (default hubs only)")
ap.add_argument("--host", action="store_true", help="drive the per-pile steps through chub_step_host (host rows, pack pool)")
ap.add_argument("--loads-every", type=int, default=0, help="N: every Nth step is a scalar-load step (chub_step_loads)")
a = ap.parse_args()
tabs = load_tables()
rs = np.random.RandomState(a.seed)
H, T = a.hubs, a.days * 96
if a.classes:
    classes = _hetero_classes(rs, a.classes)
    hub_class = rs.randint(0, len(classes), H).astype(np.int32)
    S = 66
else:
    classes, hub_class, S = [C.DEFAULT_KW], None, 47
pv, wd, price = C.synth_traces(tabs, T + 1, rs)
eng = HubEngine(H, classes, hub_class=hub_class, device=0, diagnostics=False, tables=tabs, seed=a.seed, auto_reset=True,
                action_stride=S)
GEN = a.exogenous == "device"
if GEN:
    eng.set_exogenous_generator(tabs, 0.25, 0.1)
else:
    eng.set_traces(pv, wd, price)
nt = max(1, min(a.threads, H // 64))
per = -(-H // nt)
orcs = []
for k in range(nt):
    lo, hi = k * per, min(H, (k + 1) * per)
    o = O.Oracle(hi - lo, classes, hub_class=None if hub_class is None else np.ascontiguousarray(hub_class[lo:hi]),
                 tables=C.oracle_tables(tabs))
    o.seed_base(a.seed, lo)
    if GEN:
        o.exo = O.ExoOracle(hi - lo, tabs, 0.25, 0.1)
        o.exo.seed_base(a.seed, lo)
        o.reset(*o.exo.reset())
    else:
        o.reset(pv[0], wd[0], price[0])
    orcs.append((lo, hi, o))
eng.reset()
gen = torch.Generator(device="cuda").manual_seed(a.seed)
FIELDS = ["hub_lcg", "hub_uctr", "st_line", "st_flow_in", "pile_car", "pile_charge", "pile_urgent", "pile_stay",
          "pile_already", "fcev_arrive", "fcev_qn", "hub_t"]
FLOATS = ["pile_soc", "pile_power", "pile_target", "store"]


caps = torch.tensor([[40.0 * n for n in classes[0 if hub_class is None else k]["station_list"]]
                     for k in (range(1) if hub_class is None else hub_class)], device="cuda")


def is_load_step(t):
    return a.loads_every > 0 and t % a.loads_every == a.loads_every - 1


def orc_steps(o, lo, hi, acts, t0):
    for j in range(acts.shape[0]):
        t = t0 + j
        if is_load_step(t):
            o.set_load_mode(True)
            o.step(np.ascontiguousarray(acts[j, lo:hi, :4]), pv[t + 1], wd[t + 1], price[t + 1])
            o.set_load_mode(False)
        elif GEN:
            o.step(acts[j, lo:hi], *o.exo.step(o.get("hub_t")))
        else:
            o.step(acts[j, lo:hi], pv[t + 1], wd[t + 1], price[t + 1])
        if o.get("hub_done").all():
            if GEN:
                o.reset(*o.exo.reset())
            else:
                o.reset(pv[t + 1], wd[t + 1], price[t + 1])


t_start = time.time()
bad_total = 0
for t0 in range(0, T, a.every):
    n = min(a.every, T - t0)
    acts = torch.rand((n, H, S), device="cuda", generator=gen) * 2 - 1
    for j in range(n):
        if is_load_step(t0 + j):   # columns 0..1 become power targets in kW, 2..3 stay the analog actions
            acts[j, :, :2] = ((acts[j, :, :2] + 1) * 0.5) ** 2 * caps
            eng.step_loads(acts[j, :, :4].contiguous())
        elif a.host:
            eng.step_host(acts[j].cpu().numpy())
        else:
            eng.step(acts[j])
    acts_h = acts.cpu().numpy()
    th = [threading.Thread(target=orc_steps, args=(o, lo, hi, acts_h, t0)) for lo, hi, o in orcs]
    [x.start() for x in th]
    [x.join() for x in th]
    torch.cuda.synchronize()
    for name in FIELDS + FLOATS:
        x = eng.get(name)
        y = np.concatenate([o.get(name) for _, _, o in orcs])
        if name == "pile_power":
            y = y * np.concatenate([o.get("pile_car") for _, _, o in orcs])
        if name in ("hub_lcg", "hub_uctr"):
            x, y = x.astype(np.uint32), y.astype(np.uint32)
        bad = (x != y) if name in FIELDS else ~C.close(x, y, atol=C.RTOL)
        bad = bad.reshape(H, -1).any(axis=1)
        if bad.any():
            hs = np.nonzero(bad)[0]
            print("MISMATCH after step %d field %s hubs %s (of %d bad)" % (t0 + n, name, hs[:6], hs.size), flush=True)
            h = int(hs[0])
            if hub_class is not None:
                print("  class", classes[hub_class[h]])
            for n2 in ("pile_car", "pile_charge", "pile_urgent", "pile_stay", "pile_already", "pile_soc", "pile_target",
                       "st_line", "st_flow_in", "hub_uctr", "hub_lcg", "fcev_arrive", "fcev_qn", "hub_t", "store"):
                yy = np.concatenate([o.get(n2) for _, _, o in orcs])[h]
                print("  ", n2, "\n     gpu", np.array2string(np.asarray(eng.get(n2)[h]), max_line_width=250, precision=7),
                      "\n     orc", np.array2string(np.asarray(yy), max_line_width=250, precision=7))
            bad_total += 1
            break
    if bad_total:
        break
    if (t0 // a.every) % 12 == 0:
        print("step %d ok (%.0f s)" % (t0 + n, time.time() - t_start), flush=True)
print("soak %s: %d hubs x %d steps = %.3g hub-steps, error bits %d, %.0f s"
      % ("FAILED" if bad_total else "passed", H, T, float(H) * T, eng.counters()["error_or"], time.time() - t_start))
sys.exit(1 if bad_total else 0)
