"""Regenerate charginghub_env_b200/data/hub_tables.npz from the reference checkout.

The engine needs the reference's exogenous DATA (not code): the 96x301 arrival CDF exactly as the
reference's own CSV reader parses it (CHS.hpp:93-175 parses through fp32, so the values are read
back from the compiled reference via oracle/_ref), the 96 prices, and the PV (100x96) / wind
(150x96) day traces.  Run in the build container only (needs /root/reference):
    python tools/make_data.py
"""
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_env  # noqa: E402

ps = ref_env.load_shim()
flow = np.ascontiguousarray(ps.flow_table(), dtype=np.float64)
assert flow.shape == (96, 301)
d = os.path.join(ref_env.REF_ROOT, "evcssp_env_cpp", "envs", "data_file")
price = np.array(pickle.load(open(os.path.join(d, "price_data", "price_after_MAD_96.pkl"), "rb")), np.float64)
pv = np.array(pickle.load(open(os.path.join(d, "pv_power_100.pkl"), "rb")), np.float64)
wd = np.array(pickle.load(open(os.path.join(d, "wd_power_150.pkl"), "rb")), np.float64)
assert price.shape == (96,) and pv.shape == (100, 96) and wd.shape == (150, 96)
out = os.path.join(ROOT, "charginghub_env_b200", "data", "hub_tables.npz")
np.savez_compressed(out, arrival_cdf=flow, price=price, pv_days=pv, wd_days=wd)
print("wrote", out, os.path.getsize(out), "bytes")
