"""ctypes binding of libchub.so (the C ABI declared in include/chub.h).

There is no fallback: if the shared library is missing or cannot be loaded this module raises.
Build it with `python __graft_entry__.py` (or charginghub_env_b200.build.build()).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CHUB_LIB selects another build of the same sources (libchub_checked.so, the self-checking one); never another backend
LIB_PATH = os.environ.get("CHUB_LIB") or os.path.join(_HERE, "csrc", "libchub.so")

MAX_PILES = 32
OBS_DIM = 13
EL_TABLE = 102
FCEV_QCAP = 16
DAY = 96
REPORT_LEN = 29
STATIC_MAX_HUBS = 1048576
FAST, SLOW = 0, 1

BUF_OBS, BUF_REWARD, BUF_DONE, BUF_STATS, BUF_LAST_STATS, BUF_ERROR, BUF_FINAL_SOC = range(7)

# chub_field_id -> (id, numpy dtype name, per-hub shape)
FIELDS = {
    "pile_car": (0, "int32", (2, MAX_PILES)), "pile_charge": (1, "int32", (2, MAX_PILES)),
    "pile_urgent": (2, "int32", (2, MAX_PILES)), "pile_stay": (3, "int32", (2, MAX_PILES)),
    "pile_already": (4, "int32", (2, MAX_PILES)),
    "pile_soc": (16, "float32", (2, MAX_PILES)), "pile_target": (17, "float32", (2, MAX_PILES)),
    "pile_power": (18, "float32", (2, MAX_PILES)), "pile_arrive": (19, "float32", (2, MAX_PILES)),
    "st_line": (32, "int32", (2,)), "st_flow_in": (33, "int32", (2,)), "st_car_number": (34, "int32", (2,)),
    "st_min_power": (48, "float32", (2,)), "st_charge_power": (49, "float32", (2,)),
    "st_max_power": (50, "float32", (2,)),
    "hub_t": (64, "int32", ()), "hub_uctr": (65, "int32", ()), "hub_lcg": (66, "int32", ()),
    "fcev_arrive": (67, "int32", ()), "fcev_qn": (68, "int32", ()),
    "store": (80, "float64", ()),
}
DIAG = ["reward", "income", "hy_flow", "h2_power", "fc_power", "hy_use", "not_meet", "hy_to_use", "demand",
        "ev_grid0", "ev_grid1", "h2_grid", "hy_act", "store_soc"]
for _k, _name in enumerate(DIAG):
    FIELDS["diag_" + _name] = (96 + _k, "float64", ())


class Config(ctypes.Structure):
    _fields_ = [("n_hubs", ctypes.c_int32), ("device", ctypes.c_int32), ("hub_id_base", ctypes.c_int64),
                ("n_classes", ctypes.c_int32), ("action_stride", ctypes.c_int32), ("auto_reset", ctypes.c_int32),
                ("diagnostics", ctypes.c_int32), ("reserved", ctypes.c_int32 * 4)]


class HubClass(ctypes.Structure):
    _fields_ = [("n_piles", ctypes.c_int32 * 2), ("station_type", ctypes.c_int32 * 2),
                ("constant_charging", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("hydro_prod_rate", ctypes.c_double), ("hydro_store_vlt", ctypes.c_double),
                ("init_soc", ctypes.c_double), ("fc_max_power", ctypes.c_double),
                ("fcev_permeate", ctypes.c_double), ("hydro_loss", ctypes.c_double),
                ("el_table", ctypes.POINTER(ctypes.c_double))]


class Tables(ctypes.Structure):
    _fields_ = [("arrival_cdf", ctypes.POINTER(ctypes.c_double)), ("price", ctypes.POINTER(ctypes.c_double)),
                ("price_mean", ctypes.c_double), ("price_std", ctypes.c_double),
                ("time_sin", ctypes.POINTER(ctypes.c_double))]


_vp, _i32, _i64, _u64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_uint64

# every symbol include/chub.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "chub_last_error": (ctypes.c_char_p, []),
    "chub_version": (ctypes.c_int, []),
    "chub_create": (ctypes.c_int, [ctypes.POINTER(Config), ctypes.POINTER(_vp)]),
    "chub_destroy": (ctypes.c_int, [_vp]),
    "chub_set_hub_classes": (ctypes.c_int, [_vp, ctypes.POINTER(HubClass), _i32, _vp]),
    "chub_set_tables": (ctypes.c_int, [_vp, ctypes.POINTER(Tables)]),
    "chub_get_class_table": (ctypes.c_int, [_vp, _i32, _vp]),
    "chub_set_traces": (ctypes.c_int, [_vp, _vp, _vp, _vp, _i32, _i32]),
    "chub_poke_trace": (ctypes.c_int, [_vp, _i32, ctypes.c_double, ctypes.c_double, ctypes.c_double, _vp]),
    "chub_set_exogenous_generator": (ctypes.c_int, [_vp, _vp, _i32, _vp, _i32, ctypes.c_double, ctypes.c_double]),
    "chub_seed": (ctypes.c_int, [_vp, _u64]),
    "chub_seed_explicit": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "chub_reset": (ctypes.c_int, [_vp, _vp, _i64, _vp]),
    "chub_step": (ctypes.c_int, [_vp, _vp, _vp]),
    "chub_step_into": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "chub_step_loads": (ctypes.c_int, [_vp, _vp, _vp]),
    "chub_step_loads_static": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "chub_situation_width": (ctypes.c_int, [_vp, _vp, _vp]),
    "chub_situation": (ctypes.c_int, [_vp, _vp, _vp]),
    "chub_step_host": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "chub_step_host_packed": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "chub_reset_host": (ctypes.c_int, [_vp, _vp, _vp]),
    "chub_step_packed": (ctypes.c_int, [_vp, _vp, _vp]),
    "chub_set_host_threads": (ctypes.c_int, [_vp, _i32, _i32]),
    "chub_pack_actions": (ctypes.c_int, [_vp, _vp, _vp]),
    "chub_pack_rows": (None, [_vp, _i32, _vp, _i32, _i64, _vp]),
    "chub_rollout_capture": (ctypes.c_int, [_vp, _vp, _i32, _vp]),
    "chub_rollout": (ctypes.c_int, [_vp, _vp]),
    "chub_buffer": (ctypes.c_int, [_vp, _i32, ctypes.POINTER(_vp), ctypes.POINTER(_i64)]),
    "chub_tensor": (ctypes.c_int, [_vp, _i32, ctypes.POINTER(_vp)]),
    "chub_hub_report": (ctypes.c_int, [_vp, _i32, _vp, _vp]),
    "chub_get_field": (ctypes.c_int, [_vp, _i32, _vp]),
    "chub_state_bytes": (_i64, [_vp]),
    "chub_get_state": (ctypes.c_int, [_vp, _vp]),
    "chub_set_state": (ctypes.c_int, [_vp, _vp]),
    "chub_counters": (ctypes.c_int, [_vp, ctypes.POINTER(_i64), ctypes.POINTER(_i64), ctypes.POINTER(ctypes.c_uint32)]),
    "chub_gather_stats": (ctypes.c_int, [_vp, _vp, _i32, _vp, _vp, _vp]),
    "chub_gather_wait": (ctypes.c_int, [_vp, _vp]),
    "chub_reduce_stats": (ctypes.c_int, [_vp, _vp, _vp, _vp]),
    "chub_set_timing": (ctypes.c_int, [_vp, _i32]),
    "chub_step_times": (ctypes.c_int, [_vp, _vp, _i32, ctypes.POINTER(_i32)]),
    "chub_checks": (ctypes.c_int, [_vp, ctypes.POINTER(_i32), ctypes.POINTER(_i64)]),
    "chub_debug_curve": (ctypes.c_int, [_vp, _i32, ctypes.c_uint32, _i64, _vp, _vp]),
    "chub_alloc_host": (ctypes.c_int, [_i64, ctypes.POINTER(_vp)]),
    "chub_free_host": (ctypes.c_int, [_vp]),
    "chub_algorithmic_bytes": (_i64, [_vp, _i32]),
    "chub_unit_hubs": (_i32, [_vp]),
}

_lib = None


class ChubError(RuntimeError):
    pass


def load():
    """Load libchub.so and bind every symbol of include/chub.h.  Raises if anything is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ChubError("libchub.so is not built (%s); run `python __graft_entry__.py` -- "
                        "there is no CPU fallback" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise ChubError("libchub error %d: %s" % (rc, load().chub_last_error().decode()))
