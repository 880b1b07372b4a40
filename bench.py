"""bench.py -- hub-env-steps/s of the batched step path on N B200s (one process per GPU).

A "step" is ONE launch of the fused hub step kernel over this rank's shard of hubs (plus the
one-thread tick kernel that advances the trace index).  Workloads (BASELINE.json configs):
  config5 (default): 524,288 default hubs per GPU (4,194,304 at 8 GPUs, weak scaling), shared synthetic
                     year traces, auto-reset on, actions from a rotating pool of pre-drawn matrices
  config2          : 4,096 identical default hubs (the bit-exact parity configuration; far too small to
                     fill a B200 -- reported for completeness)
  config3          : 262,144 heterogeneous hubs, auto-reset
Metric value = hubs x steps / max-over-ranks device time, inputs already resident in HBM.
e2e = the same through chub_step_host, the call a host caller of the reference makes: float action rows in pinned host
      memory in (three matrices in rotation, so they come from DRAM), obs/reward/done out to pinned host memory, every
      step; the rows are thresholded into 16-byte records by host threads before they cross PCIe (--host-threads 0: not).
--loop ppo = BASELINE configs[4]: torch policy + env + rollout buffers, one CUDA graph per 128-step segment.
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
        return dict(H=524288, classes=[DEFAULT_KW], hub_class=None, auto_reset=True, stride=47,
                    desc="config5 shard: 524,288 default hubs (20 fast + 25 slow) per GPU, auto-reset, shared year traces")
    if name == "config2":
        return dict(H=4096, classes=[DEFAULT_KW], hub_class=None, auto_reset=False, stride=47,
                    desc="config2: 4,096 identical default hubs")
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
        return dict(H=H, classes=classes, hub_class=rs.randint(0, ncls, H).astype(np.int32), auto_reset=True,
                    stride=66, desc="config3: 262,144 heterogeneous hubs (4..32 piles per station), auto-reset")
    raise SystemExit("unknown workload " + name)


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


def measured_traffic(workload, H):
    """DRAM bytes of one step-kernel launch from the committed ncu capture (profiles/), or None."""
    p = os.path.join(ROOT, "profiles", "r01j_traffic.json")
    try:
        d = json.load(open(p))
        if d["workload"] == workload and d["hubs"] == H:
            return d["busy_hour_bytes"], {"busy_hour": d["busy_hour_bytes"], "quiet_hour": d["quiet_hour_bytes"],
                                          "source": "profiles/r01j_traffic.json (ncu --set full, one launch)"}
    except Exception:
        pass
    return None, None


def cpu_port_rate(tabs, classes, hub_class, H, T, threads, stride):
    """hub-steps/s of the C restatement (oracle/) on `threads` host threads; a bounded sample."""
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
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    Hs = min(wl["H"], 2048 * cores)
    vals = []
    for _ in range(args.warmup):
        cpu_port_rate(tabs, wl["classes"], wl["hub_class"], Hs, 2, cores, wl["stride"])
    t_all = time.perf_counter()
    for _ in range(args.steps):
        v, hs, dt = cpu_port_rate(tabs, wl["classes"], wl["hub_class"], Hs, 8, cores, wl["stride"])
        vals.append(v)
        if time.perf_counter() - t_all > 150:
            break
    value = float(np.mean(vals))
    sample = "%d hubs x 8 steps per bench step on %d threads (C restatement of CHS.hpp + Python layer)" % (hs, cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(vals), "warmup": args.warmup, "ms_per_step": 1e3 * hs * 8 / value, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64/f32", "data": "synthetic",
        "config": {"workload": wl["desc"]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def run_ppo_loop(args, eng, wl, H, S, rank, world, local, barrier):
    """BASELINE.json configs[4]: PPO-style rollout loop -- segments of 128 env steps; at every step a torch policy reads the
    latest observation and writes the action rows, the engine steps, and obs / reward / done land in the segment's rollout
    buffers.  Policy, step kernel and stores of a whole segment are ONE CUDA graph
    (torch.cuda.graph around chub_step: the engine launches on torch's capture stream).  The policy is a stand-in (one
    linear layer + tanh through torch); it is the trainer's code, not this repo's product."""
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

    obs_buf[T].copy_(eng.reset())
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                       # warm-up outside capture (cuBLAS workspaces, launch shapes)
        segment()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        segment()
    nseg_w, nseg = max(1, args.warmup // T), max(1, args.steps // T)
    for _ in range(nseg_w):
        g.replay()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(nseg):
        g.replay()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    K = nseg * T
    if rank == 0:
        peak, peak_src = measured_peak()
        bytes_env = eng.algorithmic_bytes(0)
        bytes_loop = bytes_env + 13 * 4 + S * 4 + 2 * S * 4   # + policy: read obs, write rows, tanh in place (read + write)
        print(json.dumps({
            "metric": METRIC, "value": H * world * K / (ms_max * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": nseg_w * T, "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 state / f64 arithmetic", "data": "synthetic",
            "config": {"workload": wl["desc"] + "; PPO-style rollout loop: %d-step segments, torch policy (linear + tanh) on "
                                                "DLPack views, rollout buffers written every step, one CUDA graph per segment" % T,
                       "hubs_per_gpu": H, "hubs_total": H * world, "action_stride": S, "segment": T,
                       "l2": "state + per-step slices of the rollout buffers exceed L2",
                       "timer": "CUDA events around the graph replays, max over ranks"},
            "clocks": clocks, "e2e": None,
            "gpu_launches": K * 2,  # graph nodes of ours replayed in the timed region: one step kernel + one tick per env step
            "graph_replays": nseg, "kernels_per_replay_note": "per step: 1 step kernel + 1 tick (ours), 1 GEMM + 1 tanh (torch); results land in the rollout buffers directly",
            "roofline": {"bound": "hbm", "achieved": bytes_loop * H / (ms / K * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": bytes_loop * H / (ms / K * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_hub_step": bytes_loop, "kernel": "whole loop step (policy + chub_step_kernel + stores)"},
            "hub_error_bits": int(eng.counters()["error_or"])}))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=192)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="config5")
    ap.add_argument("--hubs", type=int, default=0, help="override hubs per GPU")
    ap.add_argument("--e2e-steps", type=int, default=24)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--loop", default="step", choices=["step", "ppo"],
                    help="step: bare env steps on pre-drawn actions (default, the headline metric); ppo: configs[4], "
                         "128-step segments with a torch policy and rollout buffers, one CUDA graph per segment")
    ap.add_argument("--segment", type=int, default=128)
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
    if args.impl == "reference":
        return run_reference(args, tabs, wl)

    import torch
    import torch.distributed as dist
    from charginghub_env_b200 import HubEngine
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run)" % (args.gpus, world))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    H, S = wl["H"], wl["stride"]
    K, W = args.steps, max(args.warmup, 3)
    eng = HubEngine(H, wl["classes"], hub_class=wl["hub_class"], device=local, hub_id_base=rank * H,
                    action_stride=S, auto_reset=wl["auto_reset"], diagnostics=False, tables=tabs, seed=0x5EED)
    L = 35040   # a whole year of days: the engine wraps around the trace
    if args.exogenous == "device":
        eng.set_exogenous_generator(tabs, 0.0, 0.0)
        wl["desc"] += ", exogenous inputs generated on the device"
    else:
        eng.set_traces(*year_traces(tabs, L))
    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    npool = 4  # rotating action matrices; state + actions per step exceed the 126 MB L2 for config5/config3
    pool = [torch.rand((H, S), device="cuda", generator=gen) * 2 - 1 for _ in range(npool)]
    step = eng.step
    if args.control == "loads":
        caps = torch.tensor([[40.0 * n for n in c["station_list"]] for c in wl["classes"]], device="cuda")
        caps = caps[torch.as_tensor(wl["hub_class"], device="cuda").long()] if wl["hub_class"] is not None else caps[:1]
        pool = [torch.cat([torch.rand((H, 2), device="cuda", generator=gen) ** 2 * caps,
                           torch.rand((H, 2), device="cuda", generator=gen) * 2 - 1], dim=1).contiguous()
                for _ in range(npool)]
        step = eng.step_loads
        host_rows = torch.rand((H, S), device="cuda", generator=gen) * 2 - 1   # the e2e leg stays the per-pile call
        wl["desc"] += ", scalar-load control (dispatch kernel + step kernel)"
    eng.reset()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def gather_stats():  # the one collective of the path: episode statistics gathered once per rollout
        if world > 1:
            stats = eng.tensor("last_stats")
            gathered = [torch.empty_like(stats) for _ in range(world)] if rank == 0 else None
            dist.gather(stats, gathered, dst=0)

    if args.loop == "ppo":
        del pool
        return run_ppo_loop(args, eng, wl, H, S, rank, world, local, barrier)

    for i in range(W):
        step(pool[i % npool])
    gather_stats()  # warm-up includes the collective: NCCL sets its channels up on first use
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = eng.counters()["launches"]
    barrier()
    ev0.record()
    for i in range(K):
        step(pool[i % npool])
    gather_stats()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.counters()["launches"] - launches0
    tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    value = H * world * K / (ms_max * 1e-3)

    # ---- end to end through the host-buffer C ABI call (pinned host memory both ways) ----
    # three pinned action matrices in rotation (296 MB at config 5: beyond any host cache), fresh values in each
    nhost = 3
    host_pool = [eng.host_buffers()["actions"]] + [eng.pinned_actions() for _ in range(nhost - 1)]
    for j, a in enumerate(host_pool):
        a[...] = (host_rows if args.control == "loads" else pool[j % npool]).cpu().numpy()
    KE = max(2, min(args.e2e_steps, K))
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    host_threads = args.host_threads
    if host_threads < 0:
        host_threads = min(16, ((os.cpu_count() or 2) - 1) // local_world)
        if host_threads < 4:  # too few cores per rank to threshold 98.6 MB of rows faster than PCIe carries them
            host_threads = 0
    eng.set_host_threads(host_threads, args.host_ranges)
    for j in range(3):
        eng.step_host(host_pool[j % nhost], copy=False)
    barrier()
    t0 = time.perf_counter()
    for j in range(KE):
        obs, rew, done = eng.step_host(host_pool[j % nhost], copy=False)
        _ = float(rew[0])
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = H * world * KE / float(te.item())
    err = eng.counters()["error_or"]
    n_ranges = args.host_ranges or (16 if host_threads else 8)
    per = -(-(-(-H // n_ranges)) // 1024) * 1024  # chub_step_host cuts the batch into hub ranges of whole blocks
    host_chunks = -(-H // per)

    if rank == 0:
        peak, peak_src = measured_peak()
        bytes_per_hub = eng.algorithmic_bytes(0) if len(wl["classes"]) == 1 else float(
            np.mean([2 * (16 * sum(c["station_list"]) + 80) + 4 * (sum(c["station_list"]) + 2) + 57
                     for c in wl["classes"]]))
        traffic, traffic_detail = measured_traffic(args.workload, H)
        kernel_ms = ms / K  # this rank's average step duration (step kernel + 1-thread tick), CUDA events
        achieved = bytes_per_hub * H / (kernel_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 state / f64 arithmetic", "data": "synthetic",
            "config": {"workload": wl["desc"], "hubs_per_gpu": H, "hubs_total": H * world, "action_stride": S,
                       "l2": "inputs larger than L2 (state %.0f MB + %d rotating action matrices of %.0f MB)"
                             % (H * (sum(DEFAULT_KW["station_list"]) * 16 + 80) / 1e6, npool, H * S * 4 / 1e6),
                       "timer": "CUDA events on the launch stream, max over ranks"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(H * (16 if host_threads else S * 4)),
                    "d2h_bytes_per_step": int(H * (13 * 4 + 4 + 1)), "steps": KE,
                    "host_rows_bytes_per_step": int(H * S * 4), "host_threads": host_threads,
                    "api": "chub_step_host: float action rows in pinned host memory -> %s -> device, step, obs/reward/done -> "
                           "pinned host, pipelined over %d hub ranges"
                           % ("thresholded into 16-byte records by %d host threads" % host_threads if host_threads
                              else "copied as they are", host_chunks)},
            "gpu_launches": int(launches) * (3 if args.control == "loads" else 2) + KE * (host_chunks + 1),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_detail": traffic_detail, "peak_source": peak_src, "algorithmic_bytes_per_hub_step": bytes_per_hub,
                         "kernel": "chub_step_kernel", "kernel_ms": kernel_ms},
            "hub_error_bits": int(err),
        }
        if world == 1 and not args.no_cpu:
            v, hs, dt = cpu_port_rate(tabs, wl["classes"], wl["hub_class"], 2048, 96, 1, S)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": "%d hubs x 96 steps (%.1f s) of the same hub class on 1 host thread, "
                                              "oracle/hub_oracle.c" % (hs, dt)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
