"""TEST INFRASTRUCTURE ONLY: ctypes front end of oracle/hub_oracle.c (the CPU restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")

MAXP = 32
OBS = 13
NTAB = 102
FAST, SLOW = 0, 1

# field ids of orc_get() -- (id, dtype, per-hub shape)
FIELD = {
    "pile_car": (0, np.int32, (2, MAXP)), "pile_charge": (1, np.int32, (2, MAXP)),
    "pile_urgent": (2, np.int32, (2, MAXP)), "pile_stay": (3, np.int32, (2, MAXP)),
    "pile_already": (4, np.int32, (2, MAXP)),
    "pile_soc": (16, np.float32, (2, MAXP)), "pile_target": (17, np.float32, (2, MAXP)),
    "pile_power": (18, np.float32, (2, MAXP)), "pile_arrive": (19, np.float32, (2, MAXP)),
    "pile_emergency": (20, np.float32, (2, MAXP)),
    "st_line": (32, np.int32, (2,)), "st_flow_in": (33, np.int32, (2,)),
    "st_car_number": (34, np.int32, (2,)), "st_incomplete": (35, np.int32, (2,)),
    "st_min_power": (48, np.float32, (2,)), "st_charge_power": (49, np.float32, (2,)),
    "st_max_power": (50, np.float32, (2,)),
    "hub_t": (64, np.int32, ()), "hub_uctr": (65, np.int32, ()), "hub_lcg": (66, np.int32, ()),
    "fcev_arrive": (67, np.int32, ()), "fcev_qn": (68, np.int32, ()), "hub_done": (69, np.int32, ()),
    "hub_err": (70, np.int32, ()), "fcev_arrive_sum": (71, np.int32, ()),
    "store": (80, np.float64, ()), "store_soc": (81, np.float64, ()), "reward": (82, np.float64, ()),
    "income": (83, np.float64, ()), "hy_flow": (84, np.float64, ()), "h2_power": (85, np.float64, ()),
    "fc_power": (86, np.float64, ()), "hy_use": (87, np.float64, ()), "not_meet": (88, np.float64, ()),
    "hy_to_use": (89, np.float64, ()), "demand": (90, np.float64, ()), "ep_return": (91, np.float64, ()),
    "cum_income": (92, np.float64, ()), "cum_draw": (93, np.float64, ()), "ev_grid0": (94, np.float64, ()),
    "ev_grid1": (95, np.float64, ()), "h2_grid": (96, np.float64, ()), "hy_act": (97, np.float64, ()),
    "test_penalty": (98, np.float64, ()), "deviation": (99, np.float64, ()),
    "obs": (112, np.float64, (OBS,)),
}


def build(force=False):
    src = os.path.join(_HERE, "hub_oracle.c")
    src2 = os.path.join(_HERE, "exo_oracle.c")
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= max(os.path.getmtime(src), os.path.getmtime(src2))):
        return _LIB_PATH
    os.makedirs(os.path.dirname(_LIB_PATH), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-ffp-contract=off", "-fno-builtin-pow",
                           src, src2, "-o", _LIB_PATH, "-lm"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        L.orc_create.restype = ctypes.c_void_p
        L.orc_create.argtypes = [ctypes.c_int, ctypes.c_int]
        L.orc_destroy.argtypes = [ctypes.c_void_p]
        L.orc_set_class.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 6 + [ctypes.c_double] * 6 + [ctypes.c_void_p]
        L.orc_get_class_table.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.orc_get_class_derived.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.orc_set_tables.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double,
                                     ctypes.c_double, ctypes.c_void_p]
        L.orc_set_hub_class.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.orc_seed.argtypes = [ctypes.c_void_p] * 5
        L.orc_seed_base.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64]
        L.orc_reset.argtypes = [ctypes.c_void_p] * 5
        L.orc_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 3
        L.orc_set_load_mode.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_get.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.orc_get.restype = ctypes.c_int
        L.orc_curve_batch.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_int]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def curve_batch(which, x, constant_power=False):
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty_like(x)
    lib().orc_curve_batch(which, _p(x), _p(y), x.size, int(constant_power))
    return y


def class_kwargs(station_list=(20, 25), station_type_list=("fast", "slow"), constant_charging=False,
                 hydro_prod_rate=None, hydro_store_vlt=None, init_soc=0.5, fc_max_power=None,
                 fcev_permeate=0.01, hydro_loss=0, **_ignored):
    """Reference constructor kwargs (evcssp_manager.py:25-27) -> orc_set_class argument tuple."""
    tmap = {"fast": FAST, "slow": SLOW}
    return (int(station_list[0]), int(station_list[1]), tmap[station_type_list[0]], tmap[station_type_list[1]],
            int(bool(constant_charging)),
            float(hydro_prod_rate) if hydro_prod_rate is not None else -1.0,
            float(hydro_store_vlt) if hydro_store_vlt is not None else -1.0,
            float(init_soc), float(fc_max_power) if fc_max_power is not None else -1.0,
            float(fcev_permeate), float(hydro_loss))


class Oracle(object):
    """Batch of H hubs stepped by the scalar C restatement."""

    def __init__(self, H, classes, hub_class=None, tables=None, el_tables=None):
        self.L = lib()
        self.H = int(H)
        self.classes = list(classes)
        self.E = self.L.orc_create(self.H, len(self.classes))
        for k, kw in enumerate(self.classes):
            tab = None
            if el_tables is not None and el_tables[k] is not None:
                tab = np.ascontiguousarray(el_tables[k], dtype=np.float64)
                assert tab.shape == (NTAB,)
            self.L.orc_set_class(self.E, k, *class_kwargs(**kw), _p(tab))
        if hub_class is not None:
            hub_class = np.ascontiguousarray(hub_class, dtype=np.int32)
        self.L.orc_set_hub_class(self.E, _p(hub_class))
        self.hub_class = hub_class
        if tables is not None:
            self.set_tables(**tables)

    def __del__(self):
        try:
            self.L.orc_destroy(self.E)
        except Exception:
            pass

    def set_tables(self, flow, price, price_mean, price_std, sin_table):
        flow = np.ascontiguousarray(flow, dtype=np.float64)
        price = np.ascontiguousarray(price, dtype=np.float64)
        sin_table = np.ascontiguousarray(sin_table, dtype=np.float64)
        assert flow.shape == (96, 301) and price.shape == (96,) and sin_table.shape == (96,)
        self.L.orc_set_tables(self.E, _p(flow), _p(price), float(price_mean), float(price_std), _p(sin_table))

    def class_table(self, k=0):
        out = np.empty(NTAB, np.float64)
        self.L.orc_get_class_table(self.E, k, _p(out))
        return out

    def class_derived(self, k=0):
        out = np.empty(8, np.float64)
        self.L.orc_get_class_derived(self.E, k, _p(out))
        return dict(zip(["v_h_max", "el_cells", "cap_mass", "cpr_w12", "transformer0", "transformer1", "v_M",
                         "fc_cells"], out))

    def seed(self, k0, k1, lcg, uctr=None):
        a = [np.ascontiguousarray(x, dtype=np.uint32) for x in (k0, k1)]
        c = np.ascontiguousarray(uctr, dtype=np.uint32) if uctr is not None else None
        l = np.ascontiguousarray(lcg, dtype=np.uint32)
        self.L.orc_seed(self.E, _p(a[0]), _p(a[1]), _p(c), _p(l))

    def seed_base(self, base_seed, hub_id_base=0):
        self.L.orc_seed_base(self.E, ctypes.c_uint64(base_seed), ctypes.c_uint64(hub_id_base))

    def _tr(self, x):
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, dtype=np.float64), (self.H,)))
        return x

    def reset(self, pv, wd, price, mask=None):
        pv, wd, price = self._tr(pv), self._tr(wd), self._tr(price)
        m = np.ascontiguousarray(mask, dtype=np.uint8) if mask is not None else None
        self.L.orc_reset(self.E, _p(m), _p(pv), _p(wd), _p(price))

    def step(self, actions, pv, wd, price):
        actions = np.ascontiguousarray(actions, dtype=np.float32)
        assert actions.ndim == 2 and actions.shape[0] == self.H
        pv, wd, price = self._tr(pv), self._tr(wd), self._tr(price)
        self.L.orc_step(self.E, _p(actions), actions.shape[1], _p(pv), _p(wd), _p(price))

    def set_load_mode(self, on=True):
        """Rows passed to step() become [load0 kW, load1 kW, action[-2], action[-1]] (CHS.hpp:1169-1186, 1478-1495)."""
        self.L.orc_set_load_mode(self.E, int(bool(on)))

    def get(self, name):
        fid, dt, shp = FIELD[name]
        out = np.empty((self.H,) + shp, dtype=dt)
        rc = self.L.orc_get(self.E, fid, _p(out))
        assert rc == 0, name
        return out


# ---- exogenous generator oracle (oracle/exo_oracle.c) ----
class ExoOracle(object):
    """CPU restatement of the reference's PV / wind / price generation on the per-hub X stream (include/chub_exo.h)."""

    def __init__(self, H, tables, renew_fluctuate=0.0, price_fluctuate=0.0):
        self.L, self.H = lib(), H
        L = self.L
        L.exo_create.restype = ctypes.c_void_p
        L.exo_create.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                 ctypes.c_void_p, ctypes.c_double, ctypes.c_double]
        L.exo_destroy.argtypes = [ctypes.c_void_p]
        L.exo_seed_base.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64]
        L.exo_reset.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p] * 4
        L.exo_step.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p] * 4
        L.exo_get_days.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p] * 3
        pv = np.ascontiguousarray(tables["pv_days"], np.float64)
        wd = np.ascontiguousarray(tables["wd_days"], np.float64)
        pr = np.ascontiguousarray(tables["price"], np.float64)
        self.X = L.exo_create(H, _p(pv), pv.shape[0], _p(wd), wd.shape[0], _p(pr), float(renew_fluctuate),
                              float(price_fluctuate))

    def __del__(self):
        try:
            self.L.exo_destroy(self.X)
        except Exception:
            pass

    def seed_base(self, base_seed, hub_id_base=0):
        self.L.exo_seed_base(self.X, int(base_seed), int(hub_id_base))

    def reset(self, mask=None):
        """Values of the state at time 0 after a reset: (pv, wd, price), each [H]; masked-out hubs keep zeros."""
        out = [np.zeros(self.H, np.float64) for _ in range(3)]
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        self.L.exo_reset(self.X, _p(m), _p(out[0]), _p(out[1]), _p(out[2]))
        return out

    def step(self, time_prev):
        """Values of the state entered by a step taken at per-hub time slot `time_prev`."""
        out = [np.zeros(self.H, np.float64) for _ in range(3)]
        tp = np.ascontiguousarray(np.broadcast_to(np.asarray(time_prev, np.int32), (self.H,)), np.int32)
        self.L.exo_step(self.X, _p(tp), _p(out[0]), _p(out[1]), _p(out[2]))
        return out

    def days(self):
        a, b, c = np.zeros(self.H, np.int32), np.zeros(self.H, np.int32), np.zeros(self.H, np.uint32)
        self.L.exo_get_days(self.X, _p(a), _p(b), _p(c))
        return a, b, c


def exo_normals(k0, k1, ctr0, n):
    L = lib()
    L.exo_normals.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_int, ctypes.c_void_p,
                              ctypes.c_void_p]
    out, after = np.zeros(n, np.float64), np.zeros(1, np.uint32)
    L.exo_normals(int(k0), int(k1), int(ctr0), n, _p(out), _p(after))
    return out, int(after[0])


def exo_rand31(k0, k1, ctr):
    L = lib()
    L.exo_rand31.restype = ctypes.c_uint32
    L.exo_rand31.argtypes = [ctypes.c_uint32] * 3
    return int(L.exo_rand31(int(k0), int(k1), int(ctr)))
