"""CPU: the exogenous-generator oracle (oracle/exo_oracle.c) against the package's host generator
(charginghub_env_b200/exogenous.py), both fed the same per-hub X stream.  HubExogenous itself is pinned bit for bit on
the reference RUN LIVE (renewable.ReNew / OU_Noise, EvcsspManagerEnv_v6.make_state) in
tests/test_exo_live_reference.py; the device generator is compared with this oracle in tests/test_gpu_exo.py."""
import numpy as np

from charginghub_env_b200 import exogenous, sharding
from oracle import oracle as O


class _XStream(object):
    """random.randint / numpy.random.normal drawn from hub `hub_id`'s X stream."""

    def __init__(self, base_seed, hub_id):
        k0, k1, _ = sharding.hub_streams(base_seed, np.array([hub_id], np.int64))
        self.k0, self.k1, self.ctr = int(k0[0]), int(k1[0]), 0

    def randint(self, a, b):
        v = O.exo_rand31(self.k0, self.k1, self.ctr) % (b - a + 1) + a
        self.ctr += 1
        return v

    def normal(self):
        v, self.ctr = O.exo_normals(self.k0, self.k1, self.ctr, 1)
        return float(v[0])


def test_generator_oracle_equals_host_generator(tables, monkeypatch):
    H, seed = 5, 4242
    for rf, pf in ((0.0, 0.0), (0.3, 0.2)):
        orc = O.ExoOracle(H, tables, rf, pf)
        orc.seed_base(seed, 10)
        gens, streams = [], []
        for h in range(H):
            xs = _XStream(seed, 10 + h)
            streams.append(xs)
            gens.append(exogenous.HubExogenous(tables, renew_fluctuate=rf, price_fluctuate=pf))
        for episode in range(3):
            pv, wd, pr = orc.reset()
            for h in range(H):
                monkeypatch.setattr(exogenous.random, "randint", streams[h].randint)
                monkeypatch.setattr(exogenous.np.random, "normal", streams[h].normal)
                gens[h].new_day()
                got = gens[h].state_values(0, float(tables["price"][95]))
                gens[h].reset_price_count()
                assert got == (pv[h], wd[h], pr[h]), (episode, h, got, (pv[h], wd[h], pr[h]))
            for t in range(96):
                pv, wd, pr = orc.step(t)
                for h in range(H):
                    monkeypatch.setattr(exogenous.random, "randint", streams[h].randint)
                    monkeypatch.setattr(exogenous.np.random, "normal", streams[h].normal)
                    got = gens[h].state_values((t + 1) % 96, float(tables["price"][t]))
                    assert got == (pv[h], wd[h], pr[h]), (episode, t, h)
        pd, wdd, ctr = orc.days()
        assert all(int(ctr[h]) == streams[h].ctr for h in range(H))
        assert all(0 <= pd[h] < len(tables["pv_days"]) and 0 <= wdd[h] < len(tables["wd_days"]) for h in range(H))


def test_x_stream_normals_are_standard():
    v, after = O.exo_normals(123, 456, 0, 200000)
    assert abs(v.mean()) < 0.01 and abs(v.var() - 1.0) < 0.02 and after >= 400000
    assert abs(np.mean(v ** 3)) < 0.03 and abs(np.mean(v ** 4) - 3.0) < 0.1
