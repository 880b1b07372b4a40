"""bench.py -- hub-env-steps/s of the batched step path on N B200s (one process per GPU).

A "step" is ONE launch of the fused hub step kernel over this rank's shard of hubs (its last block out advances the
trace index; only the ranged launches of the host pipeline still end with a one-thread tick kernel).  Workloads (BASELINE.json configs):
  config5 (default): 524,288 default hubs per GPU (4,194,304 at 8 GPUs, weak scaling), shared synthetic year traces,
                     auto-reset on, actions from a rotating pool of pre-drawn matrices
  config2          : 4,096 identical default hubs (the bit-exact parity configuration; far too small to fill a B200)
  config3          : 262,144 heterogeneous hubs, auto-reset
  config4          : 1,048,576 default hubs IN TOTAL (sharded over the GPUs: strong scaling), a year = 35,040 steps without
                     any reset (done every 96 steps, time wraps), CUDA-graph segments of one day
Phase: a hub's cost follows its time of day (0.18 ms per step at night, 0.48 ms when every pile is occupied), so a window of
K steps of a batch that runs in lock step measures whatever hours it happens to cover.  By default the batch is therefore
DE-PHASED during warm-up (--phase spread): 96 groups of warps are reset one step apart, so every launch holds all 96 times of
day in equal shares and any K steps cost the day's mean.  --phase lockstep keeps the reference's lock step; the line also
carries a lock-step whole-day curve (roofline.lockstep_day) measured with the engine's per-launch timers.
Metric value = hubs x steps / max-over-ranks device time, inputs already resident in HBM.
e2e        = the same through chub_step_host, the call a host caller of the reference makes: float action rows in pinned host
             memory in (three matrices in rotation, so they come from DRAM), obs/reward/done out to pinned host memory, every
             step; the rows are thresholded into 16-byte records by host threads before they cross PCIe (--host-threads 0: not).
e2e_packed = the same pipeline fed 16-byte records that already lie in host memory (chub_step_host_packed).
ppo_loop   = BASELINE configs[4]: torch policy + env + rollout buffers, one CUDA graph per 128-step segment (--loop ppo
             runs only that and prints it as the line).
--impl reference times the CPU restatement of the reference path (oracle/, "port") on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DEFAULT_KW = dict(station_list=[20, 25], station_type_list=["fast", "slow"], constant_charging=False,
                  hydro_prod_rate=100, hydro_store_vlt=25.0, init_soc=0.2, fc_max_power=100,
                  fcev_permeate=0.01, hydro_loss=0.0)  # reference test/env_test.py:10-30
METRIC = "hub_env_steps_per_sec"
UNIT = "hub-env-steps/s"
YEAR = 35040   # steps; a whole number of days: the engine wraps around the trace


def year_traces(tabs, L):
    """SURVEY.md section 8d config 4 synthetic traces (shared by all hubs)."""
    t = np.arange(L, dtype=np.float64)
    tau, d = t % 96, t // 96
    pv = 210.0 * np.maximum(0.0, np.sin(np.pi * (tau - 24) / 48.0)) * (1 + 0.3 * np.cos(2 * np.pi * d / 365.0))
    wd = np.clip(46 + 40 * np.sin(2 * np.pi * t / 355.0) + 10 * np.sin(2 * np.pi * t / 37.0), 0, 98)
    price = tabs["price"][(t.astype(np.int64) - 1) % 96] * (1 + 0.1 * np.sin(2 * np.pi * d / 7.0))
    return pv, wd, price


def workload(name, tabs):
    if name == "config5":
        return dict(H=524288, classes=[DEFAULT_KW], hub_class=None, auto_reset=True, stride=47, scaling="weak",
                    desc="config5 shard: 524,288 default hubs (20 fast + 25 slow) per GPU, auto-reset, shared year traces")
    if name == "config2":
        return dict(H=4096, classes=[DEFAULT_KW], hub_class=None, auto_reset=False, stride=47, scaling="weak",
                    desc="config2: 4,096 identical default hubs")
    if name == "config4":
        return dict(H=1048576, classes=[DEFAULT_KW], hub_class=None, auto_reset=False, stride=47, scaling="strong",
                    desc="config4: 1,048,576 default hubs in total, one year (35,040 steps) without reset -- done every 96 "
                         "steps, time wraps -- synthetic shared year traces, CUDA-graph segments of one day")
    if name == "config3":
        rs = np.random.RandomState(2)
        ncls = 4096  # distinct parameter classes; hubs draw a class at random
        classes = []
        for _ in range(ncls):
            classes.append(dict(station_list=[int(rs.randint(4, 33)), int(rs.randint(4, 33))],
                                station_type_list=["fast", "slow"], hydro_prod_rate=float(rs.uniform(50, 430)),
                                hydro_store_vlt=float(rs.uniform(10, 100)), init_soc=float(rs.uniform(0.2, 0.8)),
                                fc_max_power=float(rs.choice([50, 100, 200])),
                                fcev_permeate=float(rs.choice([0.005, 0.01, 0.02, 0.05]))))
        H = 262144
        hub_class = rs.randint(0, ncls, H).astype(np.int32)
        desc = "config3: 262,144 heterogeneous hubs (4..32 piles per station), auto-reset"
        order = os.environ.get("CHUB_BENCH_ORDER", "size")
        if order == "class":
            hub_class = np.sort(hub_class)
            desc += ", hubs created in class order"
        elif order == "size":
            # hubs are independent and their order is the caller's choice: created class by class, the classes by pile
            # count (sharding.order_by_size), a warp holds hubs of one class -- same pile counts, one parameter record --
            # and the warps of a block hubs of similar cost
            from charginghub_env_b200 import sharding
            hub_class = hub_class[sharding.order_by_size(classes, hub_class)[0]]
            desc += ", hubs created class by class in order of pile count"
        return dict(H=H, classes=classes, hub_class=hub_class, auto_reset=True, stride=66, scaling="weak", desc=desc)
    raise SystemExit("unknown workload " + name)


PHASE_TEXT = {"spread": "; de-phased: every launch holds all 96 times of day in equal shares (groups of one kernel unit of hubs "
                        "reset one step apart during warm-up)",
              "spread-warps": "; de-phased by single warps", "spread-quarters": "; de-phased by quarter units",
              "lockstep": "; lock step (all hubs at the same time of day)"}


def phase_of(args, wl):
    return args.phase or ("spread" if wl["auto_reset"] else "lockstep")


def config_dict(wl, H, world, args):
    """The `config` object, identical in both arms (the driver compares them): the workload, how its hubs are spread over
    the day, and what the GPU arm does about the L2 (the CPU arm times whole days from reset: every time of day weighs
    the same there too)."""
    phase = phase_of(args, wl)
    npool = 4
    return {"workload": wl["desc"] + (PHASE_TEXT[phase] if wl["auto_reset"] or phase != "lockstep" else ""),
            "hubs_per_gpu": H, "hubs_total": H * world, "action_stride": wl["stride"], "phase": phase,
            "l2": "GPU arm: inputs larger than L2 (state %.0f MB + %d rotating action matrices of %.0f MB)"
                  % (H * (sum(DEFAULT_KW["station_list"]) * 16 + 80) / 1e6, npool, H * wl["stride"] * 4 / 1e6)}


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.p, self.index = [], None, index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 6 and r[2 + k] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(workload, H, phase):
    """DRAM bytes of one step-kernel launch of THIS workload and phase mode from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    try:
        for d in json.load(open(p))["captures"]:
            if d["workload"] == workload and d["hubs"] == H and d["phase"] == phase:
                return d["dram_bytes"], d
    except Exception:
        pass
    return None, None


def reference_env_rate():
    """The real reference env's rate as measured in the CPU container (tools/reference_env_rate.py), labelled so."""
    p = os.path.join(ROOT, "profiles", "r02_reference_env_cpu_container.json")
    try:
        d = json.load(open(p))
        best = max((v for k, v in d.items() if k.startswith("procs_")), key=lambda v: v["env_steps_per_s"])
        return {"value": best["env_steps_per_s"], "processes": best["processes"], "per_process": best["per_process"],
                "unit": UNIT, "what": d["what"], "where": d["where"], "source": "profiles/r02_reference_env_cpu_container.json"}
    except Exception:
        return None


def cpu_port_rate(tabs, classes, hub_class, H, T, threads, stride):
    """hub-steps/s of the C restatement (oracle/) on `threads` host threads over T steps from reset; a bounded sample."""
    from oracle import oracle as O
    from tests.common import oracle_tables
    rs = np.random.RandomState(0)
    pv, wd, price = year_traces(tabs, T + 1)
    per = max(1, H // threads)
    orcs = []
    for k in range(threads):
        hc = None if hub_class is None else np.ascontiguousarray(hub_class[k * per:(k + 1) * per])
        o = O.Oracle(per, classes, hub_class=hc, tables=oracle_tables(tabs))
        o.seed_base(0x5EED, k * per)
        o.reset(pv[0], wd[0], price[0])
        orcs.append(o)
    acts = rs.uniform(-1, 1, (per, stride)).astype(np.float32)

    def work(o):
        for t in range(T):
            o.step(acts, pv[t + 1], wd[t + 1], price[t + 1])  # ctypes releases the GIL during the C call

    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(o,)) for o in orcs]
    [x.start() for x in th]
    [x.join() for x in th]
    dt = time.perf_counter() - t0
    return per * threads * T / dt, per * threads, dt


def run_reference(args, tabs, wl):
    """--impl reference: the reference path on the host cores.  The reference env itself (Python over Boost.Python C++)
    cannot run on the GPU box (it is not there), so this arm times its C restatement (oracle/hub_oracle.c, pinned bit for
    bit on the reference) on every host thread: each bench step is one WHOLE DAY (96 env steps from reset, so every time
    of day weighs the same, like the de-phased GPU batch) of a bounded sample of the workload's hubs."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    H = wl["H"] if wl["scaling"] == "weak" else wl["H"] // world
    Hs = min(H, 1024 * cores)
    vals = []
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_port_rate(tabs, wl["classes"], wl["hub_class"], Hs, 4, cores, wl["stride"])
    t_all = time.perf_counter()
    for _ in range(args.steps):
        v, hs, dt = cpu_port_rate(tabs, wl["classes"], wl["hub_class"], Hs, 96, cores, wl["stride"])
        vals.append(v)
        if time.perf_counter() - t_all > 120:
            break
    value = float(np.mean(vals))
    sample = ("%d hubs x 96 steps (one day from reset) per bench step on %d threads, C restatement of CHS.hpp + the "
              "Python layer (oracle/hub_oracle.c)" % (hs, cores))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * hs * 96 / value, "higher_is_better": True,
        "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f64/f32", "data": "synthetic",
        "config": config_dict(wl, H, world, args), "bench_steps_run": len(vals),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "reference_env_steps_per_s": reference_env_rate(),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def ppo_loop(args, eng, H, S, rank, world, local, barrier, n_segments):
    """BASELINE.json configs[4]: PPO-style rollout loop -- segments of 128 env steps; at every step a torch policy reads the
    latest observation and writes the action rows, the engine steps, and obs / reward / done land in the segment's rollout
    buffers.  Policy, step kernel and stores of a whole segment are ONE CUDA graph (torch.cuda.graph around chub_step_into:
    the engine launches on torch's capture stream).  The policy is a stand-in (one linear layer + tanh through torch); it
    is the trainer's code, not this repo's product.  Returns a dict (rank 0) or None."""
    import torch
    import torch.distributed as dist
    T = args.segment
    dev = torch.device("cuda", local)
    gen = torch.Generator(device="cuda").manual_seed(4321 + rank)
    Wp = torch.randn((13, S), device=dev, generator=gen) * 0.7
    bp = torch.randn((S,), device=dev, generator=gen) * 0.3
    act_buf = torch.empty((T, H, S), device=dev)
    obs_buf = torch.empty((T + 1, H, 13), device=dev)    # obs_buf[t] = the observation action t is chosen on
    rew_buf = torch.empty((T, H), device=dev)
    done_buf = torch.empty((T, H), device=dev, dtype=torch.uint8)

    def segment():
        obs_buf[0].copy_(obs_buf[T])                    # carry the last observation of the previous segment over
        for t in range(T):
            torch.addmm(bp, obs_buf[t], Wp, out=act_buf[t])   # policy: obs [H,13] -> action rows [H,S], written in place
            act_buf[t].tanh_()
            # the engine writes obs / reward / done of this step straight into the rollout storage (chub_step_into)
            eng.step_into(act_buf[t], obs_buf[t + 1], rew_buf[t], done_buf[t])

    obs_buf[T].copy_(eng.tensor("obs"))
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                       # warm-up outside capture (cuBLAS workspaces, launch shapes)
        segment()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        segment()
    g.replay()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(n_segments):
        g.replay()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    K = n_segments * T
    del g, act_buf, obs_buf, rew_buf, done_buf
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    bytes_loop = eng.algorithmic_bytes(0) + 13 * 4 + S * 4 + 2 * S * 4   # + policy: read obs, write rows, tanh in place
    return {"value": H * world * K / (ms_max * 1e-3), "unit": UNIT, "steps": K, "ms_per_step": ms_max / K, "segment": T,
            "graph_replays": n_segments, "algorithmic_bytes_per_hub_step": bytes_loop,
            "what": "configs[4]: %d-step segments, torch policy (linear + tanh) on DLPack views, obs/reward/done written "
                    "into the rollout buffers by chub_step_into, one CUDA graph per segment; per step 1 step kernel "
                    "(ours) + 1 GEMM + 1 tanh (torch)" % T}


def dephase(eng, H, step, pool, mode):
    """Spread the hubs over the day: group g = a scattered function of the warp (or block) index is reset one step after
    group g - 1, all hubs stepping on in between, so that after 96 steps every launch holds all 96 times of day in equal
    shares.  Traces stay aligned with each hub's own time of day (include/chub.h chub_set_traces)."""
    import torch
    unit = {"spread": eng.unit_hubs(), "spread-warps": 32, "spread-quarters": eng.unit_hubs() // 4}[mode]   # one unit of the step kernel's queue
    grp = ((torch.arange(H, device="cuda") // unit) * 37) % 96
    for j in range(96):
        eng.reset(mask=(grp == j).to(torch.uint8), step_index=-1)
        step(pool[j % len(pool)])
    torch.cuda.synchronize()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps (default 192; config4: the whole year, 35,040)")
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="config5")
    ap.add_argument("--hubs", type=int, default=0, help="override hubs per GPU")
    ap.add_argument("--e2e-steps", type=int, default=24)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ppo", action="store_true", help="skip the configs[4] PPO-loop figure of the default line")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--loop", default="step", choices=["step", "ppo"],
                    help="step: bare env steps on pre-drawn actions (default, the headline metric); ppo: configs[4], "
                         "128-step segments with a torch policy and rollout buffers, one CUDA graph per segment")
    ap.add_argument("--segment", type=int, default=128)
    ap.add_argument("--phase", default="", choices=["", "spread", "spread-warps", "spread-quarters", "lockstep"],
                    help="spread (default with auto-reset): every launch holds all 96 times of day; lockstep: as after reset()")
    ap.add_argument("--host-threads", type=int, default=-1,
                    help="threads that pack the action rows inside chub_step_host (0: float rows cross PCIe unpacked; "
                         "default: the host's cores, less one, divided between the ranks of this box, at most 16)")
    ap.add_argument("--host-ranges", type=int, default=0, help="hub ranges of the chub_step_host pipeline")
    ap.add_argument("--control", default="piles", choices=["piles", "loads"],
                    help="piles: per-pile action rows (the gym env's step, default); loads: one power target per station "
                         "(the stations' evs_step(float)), which adds the dispatch kernel to every step")
    ap.add_argument("--exogenous", default="traces", choices=["traces", "device"],
                    help="traces: shared synthetic year (default); device: the on-device PV/wind/price generator")
    args = ap.parse_args()

    from charginghub_env_b200 import load_tables
    tabs = load_tables()
    wl = workload(args.workload, tabs)
    if args.hubs:
        wl["H"] = args.hubs
        if wl["hub_class"] is not None:
            wl["hub_class"] = np.resize(wl["hub_class"], args.hubs)
    if not args.steps:
        args.steps = YEAR if args.workload == "config4" else 192
    if args.exogenous == "device":
        wl["desc"] += ", exogenous inputs generated on the device"
    if args.control == "loads":
        wl["desc"] += ", scalar-load control (dispatch kernel + step kernel)"
    if args.impl == "reference":
        return run_reference(args, tabs, wl)

    import torch
    import torch.distributed as dist
    from charginghub_env_b200 import HubEngine, sharding
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run)" % (args.gpus, world))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    S = wl["stride"]
    if wl["scaling"] == "strong":      # config 4: the batch is fixed, the GPUs share it
        lo, hi = sharding.shard_range(wl["H"], rank, world)
        H, base = hi - lo, lo
        per_rank = [sharding.shard_range(wl["H"], r, world)[1] - sharding.shard_range(wl["H"], r, world)[0] for r in range(world)]
    else:
        H, base, per_rank = wl["H"], rank * wl["H"], None
    K, W = args.steps, max(args.warmup, 3)
    hub_class = wl["hub_class"]
    eng = HubEngine(H, wl["classes"], hub_class=hub_class, device=local, hub_id_base=base,
                    action_stride=S, auto_reset=wl["auto_reset"], diagnostics=False, tables=tabs, seed=0x5EED)
    if args.exogenous == "device":
        eng.set_exogenous_generator(tabs, 0.0, 0.0)
    else:
        eng.set_traces(*year_traces(tabs, YEAR))
    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    npool = 4  # rotating action matrices; state + actions per step exceed the 126 MB L2 for config5/config3/config4
    pool = [torch.rand((H, S), device="cuda", generator=gen) * 2 - 1 for _ in range(npool)]
    step = eng.step
    if args.control == "loads":
        caps = torch.tensor([[40.0 * n for n in c["station_list"]] for c in wl["classes"]], device="cuda")
        caps = caps[torch.as_tensor(hub_class, device="cuda").long()] if hub_class is not None else caps[:1]
        pool = [torch.cat([torch.rand((H, 2), device="cuda", generator=gen) ** 2 * caps,
                           torch.rand((H, 2), device="cuda", generator=gen) * 2 - 1], dim=1).contiguous()
                for _ in range(npool)]
        step = eng.step_loads
        host_rows = torch.rand((H, S), device="cuda", generator=gen) * 2 - 1   # the e2e leg stays the per-pile call
    phase = phase_of(args, wl)
    if phase != "lockstep" and not wl["auto_reset"]:
        raise SystemExit("--phase %s needs an auto-reset workload" % phase)
    eng.reset()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    comm = None
    gathered = None
    if world > 1:
        dist.all_reduce(torch.zeros(1, device="cuda"))    # torch creates its NCCL communicator on first use
        comm = sharding.nccl_comm_ptr(local)
        total = sum(per_rank) if per_rank else H * world
        gathered = torch.empty((total, 4), device="cuda") if rank == 0 else None

    def gather_start():  # the one collective of the path, through the C ABI: snapshot now, transfer on the comm stream
        if comm:
            eng.gather_stats(comm, 0, per_rank, gathered)

    if phase != "lockstep":
        dephase(eng, H, step, pool, phase)
    if args.loop == "ppo":
        res = ppo_loop(args, eng, H, S, rank, world, local, barrier, max(1, K // args.segment))
        if rank == 0:
            peak, peak_src = measured_peak()
            ach = res["algorithmic_bytes_per_hub_step"] * H / (res["ms_per_step"] * 1e-3) / 1e9
            print(json.dumps({
                "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": res["steps"],
                "warmup": args.segment * 2, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": wl["scaling"],
                "vs_baseline": None, "dtype": "f32 state / f64 arithmetic", "data": "synthetic",
                "config": dict(config_dict(wl, H, world, args), loop=res["what"]), "e2e": None, "gpu_launches": res["steps"],
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                             "peak_source": peak_src, "kernel": "whole loop step (policy + chub_step_kernel + stores)"},
                "hub_error_bits": int(eng.counters()["error_or"])}))
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- the timed region: K steps + the rollout's one collective (statistics of the episodes finished so far) ----
    graph_steps = 96 if args.workload == "config4" else 0
    soc_lo = soc_hi = None
    if graph_steps:   # config 4: one CUDA graph per day; the action matrices rotate inside it
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for j in range(graph_steps):
                step(pool[j % npool])
        torch.cuda.current_stream().wait_stream(side)
        K = max(graph_steps, K // graph_steps * graph_steps)
        W = graph_steps
        soc_lo = torch.full((1,), 2.0, device="cuda")
        soc_hi = torch.full((1,), -1.0, device="cuda")
        g.replay()     # warm-up: day 0
    else:
        for i in range(W):
            step(pool[i % npool])
    gather_start()     # warm-up includes the collective: NCCL sets its channels up on first use
    eng.gather_wait()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = eng.counters()["launches"]
    barrier()
    ev0.record()
    gather_start()          # statistics of the episodes that ended before this rollout travel while it runs
    if graph_steps:
        obs = eng.tensor("obs")
        for d in range(K // graph_steps):
            g.replay()
            torch.minimum(soc_lo, obs[:, 10].min(), out=soc_lo)   # Store_SOC (obs[10]) must stay inside [0.1, 1]
            torch.maximum(soc_hi, obs[:, 10].max(), out=soc_hi)
    else:
        for i in range(K):
            step(pool[i % npool])
    eng.gather_wait()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.counters()["launches"] - launches0 + (K if graph_steps else 0)
    tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    hubs_all = sum(per_rank) if per_rank else H * world
    value = hubs_all * K / (ms_max * 1e-3)

    # ---- the step kernel alone: per-launch CUDA events on the launch stream (engine timers), same state, same inputs ----
    eng.set_timing(True)
    KT = min(96, max(K, 8))
    for i in range(KT):
        step(pool[i % npool])
    launch_ms = eng.step_times(128)[-KT:]
    eng.set_timing(False)
    kernel_ms = float(np.mean(launch_ms))

    ppo = None
    if not args.no_ppo and args.workload == "config5" and args.control == "piles":
        ppo = ppo_loop(args, eng, H, S, rank, world, local, barrier, 2)

    # ---- end to end through the host-buffer C ABI calls (pinned host memory both ways) ----
    e2e = e2e_packed = None
    host_chunks = 0
    if not args.no_e2e and not graph_steps:
        # three pinned action matrices in rotation (296 MB at config 5: beyond any host cache), fresh values in each
        nhost = 3
        host_pool = [eng.host_buffers()["actions"]] + [eng.pinned_actions() for _ in range(nhost - 1)]
        for j, a in enumerate(host_pool):
            a[...] = (host_rows if args.control == "loads" else pool[j % npool]).cpu().numpy()
        KE = max(2, min(args.e2e_steps, K))
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
        host_threads = args.host_threads
        if host_threads < 0:   # the ranks of a box share its cores; below two packing threads the rows go unpacked
            host_threads = min(16, max(2, ((os.cpu_count() or 2) - 1) // local_world))
        eng.set_host_threads(host_threads, args.host_ranges)

        def timed_host(fn, bufs, windows=2):
            """KE steps, max over ranks; the better of `windows` back-to-back windows (the host side of this path shares its
            cores and memory with whatever else runs on the box: one window in five came out 25 % low on a busy host)."""
            for j in range(3):
                fn(bufs[j % len(bufs)])
            best = None
            for _ in range(windows):
                barrier()
                t0 = time.perf_counter()
                for j in range(KE):
                    obs_h, rew_h, done_h = fn(bufs[j % len(bufs)])
                    _ = float(rew_h[0])
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                te = torch.tensor([dt], device="cuda", dtype=torch.float64)
                if world > 1:
                    dist.all_reduce(te, op=dist.ReduceOp.MAX)
                best = float(te.item()) if best is None else min(best, float(te.item()))
            return H * world * KE / best, best

        e2e_value, e2e_s = timed_host(lambda a: eng.step_host(a, copy=False), host_pool)
        recs = [eng.pinned_records() for _ in range(nhost)]
        for j in range(nhost):
            recs[j][...] = eng.pack_actions(host_pool[j])
        e2e_packed_value, e2e_packed_s = timed_host(eng.step_host_packed, recs)
        n_ranges = args.host_ranges or (16 if host_threads else 8)
        per = -(-(-(-H // n_ranges)) // 1024) * 1024  # chub_step_host cuts the batch into hub ranges of whole blocks
        host_chunks = -(-H // per)
        # what the host's memory system delivers to the same number of threads: memmove of the same rows, slice per thread
        import ctypes
        dst = np.empty((H, S), np.float32)
        bounds = np.linspace(0, H, max(1, host_threads) + 1).astype(np.int64)

        def cp(lo, hi):
            ctypes.memmove(dst[lo:hi].ctypes.data, host_pool[1][lo:hi].ctypes.data, int(hi - lo) * S * 4)   # releases the GIL
        best = 1e9
        for _ in range(3):
            th = [threading.Thread(target=cp, args=(bounds[i], bounds[i + 1])) for i in range(len(bounds) - 1)]
            t0 = time.perf_counter()
            [x.start() for x in th]
            [x.join() for x in th]
            best = min(best, time.perf_counter() - t0)
        host_copy_gbs = 2 * H * S * 4 / best / 1e9   # bytes read + bytes written
        d2h = int(H * (13 * 4 + 4 + 1))
        e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(H * (16 if host_threads else S * 4)),
               "d2h_bytes_per_step": d2h, "steps": KE, "windows": "the better of 2 windows of %d steps" % KE,
               "host_rows_bytes_per_step": int(H * S * 4), "host_threads": host_threads,
               "host_rows_read_gbs": H * S * 4 * KE / e2e_s / 1e9, "host_memmove_gbs_same_threads": host_copy_gbs,
               "api": "chub_step_host: float action rows in pinned host memory -> %s -> device, step, obs/reward/done -> "
                      "pinned host, pipelined over %d hub ranges"
                      % ("thresholded into 16-byte records by %d host threads" % host_threads if host_threads
                         else "copied as they are", host_chunks)}
        e2e_packed = {"value": e2e_packed_value, "unit": UNIT, "h2d_bytes_per_step": int(H * 16), "d2h_bytes_per_step": d2h,
                      "steps": KE, "api": "chub_step_host_packed: 16-byte records in pinned host memory -> device, step, "
                                          "obs/reward/done -> pinned host, same pipeline"}

    # ---- lock-step whole day: the per-step curve behind the mean (per-launch engine timers) ----
    day = None
    if wl["auto_reset"] and args.control == "piles":
        eng.reset()
        for i in range(96):          # day 0 from reset() starts with half-empty stations; time day 1
            step(pool[i % npool])
        eng.set_timing(True)
        for i in range(96):
            step(pool[i % npool])
        curve = eng.step_times(128)[-96:]
        eng.set_timing(False)
        day = {"mean_ms": float(curve.mean()), "min_ms": float(curve.min()), "max_ms": float(curve.max()),
               "busy_hours_ms": float(np.sort(curve)[48:-1].mean()), "reset_step_ms": float(curve[-1]),
               "curve_ms": [round(float(x), 4) for x in curve],
               "what": "96 consecutive launches of a batch in lock step (second day after reset), per-launch CUDA events"}
    err = eng.counters()["error_or"]

    if rank == 0:
        peak, peak_src = measured_peak()
        bytes_per_hub = eng.algorithmic_bytes(0) if len(wl["classes"]) == 1 else float(
            np.mean([2 * (16 * sum(c["station_list"]) + 80) + 4 * (sum(c["station_list"]) + 2) + 57
                     for c in (wl["classes"][k] for k in hub_class)]))
        traffic, traffic_detail = measured_traffic(args.workload, H, phase)
        achieved = bytes_per_hub * H / (kernel_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_detail": traffic_detail, "peak_source": peak_src,
                "algorithmic_bytes_per_hub_step": bytes_per_hub, "kernel": "chub_step_kernel", "kernel_ms": kernel_ms,
                "kernel_ms_min_max": [float(np.min(launch_ms)), float(np.max(launch_ms))], "launches_timed": int(KT),
                "timed_region_ms_per_step": ms / K}
        if day:
            roof["lockstep_day"] = day
            roof["day_mean_frac"] = bytes_per_hub * H / (day["mean_ms"] * 1e-3) / 1e9 / peak
            roof["busy_hours_frac"] = bytes_per_hub * H / (day["busy_hours_ms"] * 1e-3) / 1e9 / peak
        cfg = config_dict(wl, H if wl["scaling"] == "weak" else wl["H"] // world, world, args)
        measurement = {
            "window": "steps %d..%d after the warm-up of a de-phased batch: every step holds every time of day" % (0, K - 1)
                      if phase != "lockstep" else "steps %d..%d after reset()" % (W, W + K - 1),
            "dephase_group_hubs": eng.unit_hubs() if phase == "spread" else None,
            "timer": "CUDA events on the launch stream, max over ranks; kernel_ms from per-launch events (chub_step_times)",
            "collective": "chub_gather_stats (C ABI, NCCL send/recv on the engine's communication stream) started with "
                          "the rollout, joined before the closing event" if world > 1 else "none at 1 GPU"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup if not graph_steps else W,
            "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
            "dtype": "f32 state / f64 arithmetic", "data": "synthetic", "config": cfg, "clocks": clocks,
            "e2e": e2e, "e2e_packed": e2e_packed, "ppo_loop": ppo,
            "gpu_launches": int(launches) * (2 if args.control == "loads" else 1)
                            + (int(e2e["steps"]) * (host_chunks + 1) * 2 * 2 if e2e else 0),   # e2e + e2e_packed, two windows each
            "roofline": roof, "measurement": measurement, "hub_error_bits": int(err),
            "reference_env_steps_per_s": reference_env_rate(),
        }
        if graph_steps:
            line["year"] = {"days": K // graph_steps, "store_soc_min": float(soc_lo.item()), "store_soc_max": float(soc_hi.item()),
                            "store_within_bounds": bool(0.1 - 1e-6 <= float(soc_lo.item()) and float(soc_hi.item()) <= 1.0 + 1e-6),
                            "episodes_flagged_done": K // 96, "graph": "one CUDA graph of 96 step launches per day"}
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            v, hs, dt = cpu_port_rate(tabs, wl["classes"], hub_class, min(H, 1024 * cores), 96, cores, S)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "%d hubs x 96 steps (one day from reset, %.1f s) of the same hub classes on %d host "
                                              "threads, oracle/hub_oracle.c" % (hs, dt, cores)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
