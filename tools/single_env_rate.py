"""Steps/s of the drop-in single-hub gym env (EvcsspManagerEnv_v6 over the engine), config 1 of BASELINE.json."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from charginghub_env_b200.env import EvcsspManagerEnv_v6
import bench

kw = {k: v for k, v in bench.DEFAULT_KW.items()}
env = EvcsspManagerEnv_v6(seed_rand=False, **kw)
rs = np.random.RandomState(0)
env.reset()
for _ in range(20):
    env.step(rs.uniform(-1, 1, 47))
n = 0
t0 = time.perf_counter()
for ep in range(5):
    env.reset()
    done = False
    while not done:
        _, _, done, _ = env.step(rs.uniform(-1, 1, 47))
        n += 1
dt = time.perf_counter() - t0
print("single-hub gym env: %d steps in %.3f s = %.0f steps/s (%.3f ms per step)" % (n, dt, n / dt, dt / n * 1e3))
