"""The REAL reference env (unmodified Python layer over the unmodified CHS.hpp through the pybind11 shim, oracle/ref_env.py)
timed the way north_star asks for the CPU figure: one process per core, one env per process (process-global RNG state,
CHS.hpp:23-25), random actions, stdout discarded (test/env_test.py:10-40).  Runs only where /root/reference exists (the
CPU container); writes profiles/r02_reference_env_cpu_container.json, which bench.py quotes -- labelled as measured
here, not on the GPU box -- next to its own CPU figure.
    python tools/reference_env_rate.py [--procs N] [--episodes 20]
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(args):
    idx, episodes = args
    import random
    import numpy as np
    from oracle import ref_env
    Env = ref_env.load_reference_env_class()
    random.seed(idx)
    np.random.seed(idx)
    rs = np.random.RandomState(idx)
    with ref_env.quiet():
        env = Env(**ref_env.DEFAULT_KWARGS)
        env.reset()
        for _ in range(96):                      # warm-up episode
            env.step(rs.uniform(-1, 1, 47))
        t0 = time.perf_counter()
        n = 0
        for _ in range(episodes):
            env.reset()
            done = False
            while not done:
                _, _, done, _ = env.step(rs.uniform(-1, 1, 47))
                n += 1
        dt = time.perf_counter() - t0
    return n, dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    ap.add_argument("--episodes", type=int, default=20)
    a = ap.parse_args()
    out = {}
    for procs in sorted({1, a.procs}):
        t0 = time.perf_counter()
        with mp.get_context("spawn").Pool(procs) as pool:
            res = pool.map(worker, [(i, a.episodes) for i in range(procs)])
        wall = time.perf_counter() - t0
        steps = sum(r[0] for r in res)
        rate = sum(r[0] / r[1] for r in res)
        out["procs_%d" % procs] = {"processes": procs, "env_steps": steps, "env_steps_per_s": rate,
                                   "per_process": rate / procs, "wall_s_incl_start": wall}
        print(procs, "processes:", "%.0f env-steps/s" % rate)
    out["what"] = ("unmodified reference env (evcssp_manager.py over CHS.hpp, pybind11 shim instead of Boost.Python), one process "
                   "per core, random actions, %d episodes per process, stdout discarded" % a.episodes)
    out["where"] = "CPU container of the build (%d vCPUs), NOT the GPU box: /root/reference does not travel" % (os.cpu_count() or 0)
    with open(os.path.join(ROOT, "profiles", "r02_reference_env_cpu_container.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
