"""Exogenous inputs of a hub: PV / wind day traces with Ornstein-Uhlenbeck noise and the noisy price.

The engine consumes traces; this module PRODUCES them on the host the way the reference does
(renewable.py:21-76 ReNew / OU_Noise, evcssp_manager.py:344-361 make_state), drawing from the same
global generators in the same order -- Python's `random` for the day choice, `numpy.random` for the
three OU processes -- so a script that seeds them like test/env_test.py:14-17 sees the reference's
exogenous trajectory.  It holds no hub dynamics.
"""
import random

import numpy as np

SCALE_PV, SCALE_WD = 5, 1  # evcssp_manager.py:346-347


class OUProcess(object):
    """x += theta * (mu - x) + sigma * N(0,1)  (renewable.py:56-76); one np.random.normal() per sample."""

    def __init__(self, theta, sigma, mu=0.0):
        self.theta, self.sigma, self.mu = theta, sigma, mu
        random.seed(1)  # the reference's OU_Noise reseeds Python's `random` on construction (renewable.py:63)
        self.x = mu

    def sample(self):
        self.x = self.x + (self.theta * (self.mu - self.x) + self.sigma * np.random.normal())
        return self.x


class HubExogenous(object):
    """Per-hub generator of (pv x5, wind, price) for successive state times."""

    def __init__(self, tables, renew_fluctuate=0.0, price_fluctuate=0.0):
        self.pv_days, self.wd_days, self.price_table = tables["pv_days"], tables["wd_days"], tables["price"]
        self.renew_fluctuate, self.price_fluctuate = renew_fluctuate, price_fluctuate
        # construction order of the reference: price noise (evcssp_manager.py:30), then pv, wd (renewable.py:34-35)
        self.price_noise = OUProcess(0.1, 0.005)
        self.pv_noise = OUProcess(0.01, 1.0)
        self.wd_noise = OUProcess(0.01, 1.5)
        self.pv_day = self.wd_day = None
        self.price_count = 0
        self.price_hold = 0.0

    def new_day(self):  # renew_reset, renewable.py:51-53
        self.pv_day = random.randint(0, len(self.pv_days) - 1)
        self.wd_day = random.randint(0, len(self.wd_days) - 1)

    def state_values(self, time, last_table_price):
        """Values shown in the state at `time` (make_state): pv, wd first, then the price with held noise."""
        pv = float(self.pv_days[self.pv_day][time])
        if pv > 0 and self.pv_day % 2 == 0:
            pv += float(self.pv_noise.sample()) * (1 + self.renew_fluctuate)
        pv = max(0, pv) * SCALE_PV
        wd = float(self.wd_days[self.wd_day][time])
        wd += float(self.wd_noise.sample()) * (1 + self.renew_fluctuate)
        wd = max(0, wd) * SCALE_WD
        if self.price_count % 4 == 0:
            self.price_hold = float(self.price_noise.sample()) * (1 + self.price_fluctuate)
        price = last_table_price + self.price_hold
        self.price_count += 1
        return pv, wd, price

    def reset_price_count(self):  # evcssp_manager.py:313 (after make_state(0))
        self.price_count = 0


def synthetic_year(tables, length=35040):
    """Shared synthetic traces of SURVEY.md section 8d (config 4): PV, wind, price over `length` steps.  The engine
    addresses traces modulo their length (include/chub.h chub_set_traces), so `length` is a whole number of days: a
    rollout that runs past the year wraps onto slot 0 with the time of day intact."""
    if length % 96:
        raise ValueError("length must be a multiple of 96 (the engine wraps around the trace)")
    t = np.arange(length, dtype=np.float64)
    tau, d = t % 96, t // 96
    pv = 210.0 * np.maximum(0.0, np.sin(np.pi * (tau - 24) / 48.0)) * (1 + 0.3 * np.cos(2 * np.pi * d / 365.0))
    wd = np.clip(46 + 40 * np.sin(2 * np.pi * t / 355.0) + 10 * np.sin(2 * np.pi * t / 37.0), 0, 98)
    price = tables["price"][(t.astype(np.int64) - 1) % 96] * (1 + 0.1 * np.sin(2 * np.pi * d / 7.0))
    return pv, wd, price


def sampled_days(tables, n_days, rs, n_hubs=None):
    """Noise-free traces stitched from random PV / wind days of the data set: [96*n_days] or per hub (a ring: the engine
    wraps around it)."""
    shape = (n_days,) if n_hubs is None else (n_days, n_hubs)
    pd, wdd = rs.randint(0, len(tables["pv_days"]), shape), rs.randint(0, len(tables["wd_days"]), shape)
    pv = np.moveaxis(tables["pv_days"][pd], -1, 1).reshape((96 * n_days,) + shape[1:]) * SCALE_PV
    wd = np.moveaxis(tables["wd_days"][wdd], -1, 1).reshape((96 * n_days,) + shape[1:])
    price = tables["price"][(np.arange(96 * n_days) - 1) % 96]
    price = price.reshape((-1,) + (1,) * (len(shape) - 1)) * np.ones(shape[1:])
    return np.ascontiguousarray(pv), np.ascontiguousarray(wd), np.ascontiguousarray(price)
