"""HubEngine: Python face of the C ABI (include/chub.h) -- one shard of hubs on one B200.

It mirrors, for a batch, what the reference does for one hub through EvcsspManagerEnv_v6
(evcssp_env_cpp/envs/evcssp_manager.py:19-414): constructor kwargs become hub classes, reset() and
step(action) keep their meaning, observations / rewards / done flags come back as torch tensors
that alias engine memory through DLPack (no copy), or as numpy arrays through the host-buffer path.
All stepping happens in the CUDA library; this module contains no simulation arithmetic.
"""
import ctypes
import os

import numpy as np

from . import _lib
from ._lib import ChubError, check

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "hub_tables.npz")
_TYPE = {"fast": _lib.FAST, "slow": _lib.SLOW}


def load_tables(path=None):
    """Static tables of the hub model: arrival CDF, prices, PV / wind day traces (tools/make_data.py)."""
    z = np.load(path or _DATA)
    price = np.ascontiguousarray(z["price"], np.float64)
    k = 2 * np.pi / 96
    return dict(arrival_cdf=np.ascontiguousarray(z["arrival_cdf"], np.float64), price=price,
                price_mean=float(np.mean(price)), price_std=float(np.std(price)),
                time_sin=np.sin(k * np.arange(96, dtype=np.float64)),
                pv_days=np.ascontiguousarray(z["pv_days"], np.float64),
                wd_days=np.ascontiguousarray(z["wd_days"], np.float64))


def make_hub_class(station_list=(20, 25), station_type_list=("fast", "slow"), constant_charging=False,
                   hydro_prod_rate=None, hydro_store_vlt=None, seed_rand=True, init_soc=0.5, fc_max_power=None,
                   fcev_permeate=0.01, use_lagrange=False, renew_fluctuate=0, price_fluctuate=0, hydro_loss=0,
                   el_table=None):
    """The reference's constructor kwargs (evcssp_manager.py:25-27) -> chub_hub_class."""
    if len(station_list) != 2 or len(station_type_list) != 2:
        raise ValueError("station_list and station_type_list must have length 2")  # evcssp_manager.py:37
    c = _lib.HubClass()
    c.n_piles[0], c.n_piles[1] = int(station_list[0]), int(station_list[1])
    try:
        c.station_type[0], c.station_type[1] = _TYPE[station_type_list[0]], _TYPE[station_type_list[1]]
    except KeyError:
        raise ValueError("EVS type must be fast or slow")  # Aggregator_Simple.py:196
    c.constant_charging = int(bool(constant_charging))
    c.hydro_prod_rate = float(hydro_prod_rate) if hydro_prod_rate is not None else -1.0
    c.hydro_store_vlt = float(hydro_store_vlt) if hydro_store_vlt is not None else -1.0
    c.init_soc = float(init_soc)
    c.fc_max_power = float(fc_max_power) if fc_max_power is not None else -1.0
    c.fcev_permeate = float(fcev_permeate)
    c.hydro_loss = float(hydro_loss)
    keep = None
    if el_table is not None:
        keep = np.ascontiguousarray(el_table, np.float64)
        if keep.shape != (_lib.EL_TABLE,):
            raise ValueError("el_table must have %d entries" % _lib.EL_TABLE)
        c.el_table = keep.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
    return c, keep


def _dptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _tensor_ptr(t, dtype=None, device=None):
    """Device pointer of a contiguous torch CUDA tensor (or a raw int).  dtype: torch dtype or tuple of dtypes the
    kernel may reinterpret the bytes as; device: the engine's CUDA ordinal -- the reference would have converted a
    wrong-typed array, the kernel would silently read garbage, so both are checked here."""
    if isinstance(t, int):
        return ctypes.c_void_p(t)
    if not t.is_cuda or not t.is_contiguous():
        raise ValueError("expected a contiguous CUDA tensor")
    if dtype is not None and t.dtype not in (dtype if isinstance(dtype, tuple) else (dtype,)):
        raise ValueError("expected a %s tensor, got %s" % (dtype, t.dtype))
    if device is not None and t.device.index != device:
        raise ValueError("tensor lives on cuda:%s, the engine on cuda:%d" % (t.device.index, device))
    return ctypes.c_void_p(t.data_ptr())


def _current_stream(device):
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class HubEngine(object):
    def __init__(self, n_hubs, classes=None, hub_class=None, device=0, hub_id_base=0, action_stride=None,
                 auto_reset=False, diagnostics=False, tables=None, seed=0):
        self.lib = _lib.load()
        classes = list(classes) if classes is not None else [dict()]
        self.class_kwargs = classes
        self._keep = []
        arr = (_lib.HubClass * len(classes))()
        n_act = 0
        for k, kw in enumerate(classes):
            c, keep = make_hub_class(**kw)
            arr[k] = c
            self._keep.append(keep)
            n_act = max(n_act, c.n_piles[0] + c.n_piles[1] + 2)
        self.n_hubs = int(n_hubs)
        self.device = int(device)
        self.action_dim = n_act
        self.action_stride = int(action_stride) if action_stride is not None else n_act
        cfg = _lib.Config(n_hubs=self.n_hubs, device=self.device, hub_id_base=int(hub_id_base),
                          n_classes=len(classes), action_stride=self.action_stride,
                          auto_reset=int(bool(auto_reset)), diagnostics=int(bool(diagnostics)))
        h = ctypes.c_void_p()
        check(self.lib.chub_create(ctypes.byref(cfg), ctypes.byref(h)))
        self.h = h
        self.diagnostics = bool(diagnostics)
        hc = None
        if hub_class is not None:
            hc = np.ascontiguousarray(hub_class, np.int32)
            if hc.shape != (self.n_hubs,):
                raise ValueError("hub_class must have one entry per hub")
        self.hub_class = hc
        check(self.lib.chub_set_hub_classes(self.h, arr, len(classes), _dptr(hc)))
        self.tables = tables if tables is not None else load_tables()
        t = _lib.Tables()
        self._tab_keep = [np.ascontiguousarray(self.tables[k], np.float64) for k in ("arrival_cdf", "price", "time_sin")]
        t.arrival_cdf = self._tab_keep[0].ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        t.price = self._tab_keep[1].ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        t.time_sin = self._tab_keep[2].ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        t.price_mean, t.price_std = float(self.tables["price_mean"]), float(self.tables["price_std"])
        check(self.lib.chub_set_tables(self.h, ctypes.byref(t)))
        self._tensors = {}
        self._host = None
        if seed is not None:
            self.seed(seed)

    # ---- lifetime ----
    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self._tensors.clear()
            if self._host is not None:
                for p in self._host["ptrs"]:
                    self.lib.chub_free_host(p)
                self._host = None
            self.lib.chub_destroy(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- configuration ----
    def class_table(self, k=0):
        out = np.empty(_lib.EL_TABLE, np.float64)
        check(self.lib.chub_get_class_table(self.h, k, _dptr(out)))
        return out

    def algorithmic_bytes(self, k=0):
        return int(self.lib.chub_algorithmic_bytes(self.h, k))

    def unit_hubs(self):
        """Hubs per scheduling unit of a whole-batch step launch (chub_unit_hubs): hubs of one unit share an SM."""
        return int(self.lib.chub_unit_hubs(self.h))

    def set_traces(self, pv, wd, price, per_hub=False):
        """pv (already x5), wd, price (table[t-1] + noise): [L] shared or [L, n_hubs] per hub."""
        a = [np.ascontiguousarray(x, np.float64) for x in (pv, wd, price)]
        want = (a[0].shape[0], self.n_hubs) if per_hub else (a[0].shape[0],)
        for x in a:
            if x.shape != want:
                raise ValueError("trace shape %s, expected %s" % (x.shape, want))
        check(self.lib.chub_set_traces(self.h, _dptr(a[0]), _dptr(a[1]), _dptr(a[2]), a[0].shape[0], int(per_hub)))

    def set_exogenous_generator(self, tables=None, renew_fluctuate=0.0, price_fluctuate=0.0):
        """Generate PV / wind / price on the device (per hub: random days at every reset + OU noise, the reference's
        ReNew / OU_Noise / make_state on the hub's X stream) instead of reading traces."""
        tabs = tables if tables is not None else load_tables()
        pv = np.ascontiguousarray(tabs["pv_days"], np.float64)
        wd = np.ascontiguousarray(tabs["wd_days"], np.float64)
        if pv.ndim != 2 or wd.ndim != 2 or pv.shape[1] != 96 or wd.shape[1] != 96:
            raise ValueError("pv_days / wd_days must be [n_days, 96]")
        check(self.lib.chub_set_exogenous_generator(self.h, _dptr(pv), pv.shape[0], _dptr(wd), wd.shape[0],
                                                    float(renew_fluctuate), float(price_fluctuate)))

    def poke_trace(self, slot, pv, wd, price, stream=0):
        check(self.lib.chub_poke_trace(self.h, int(slot), float(pv), float(wd), float(price), ctypes.c_void_p(int(stream))))

    def seed(self, base_seed):
        check(self.lib.chub_seed(self.h, ctypes.c_uint64(int(base_seed) & ((1 << 64) - 1))))

    def seed_explicit(self, key0, key1, lcg, uctr=None):
        a = [np.ascontiguousarray(x, np.uint32) for x in (key0, key1, lcg)]
        c = np.ascontiguousarray(uctr, np.uint32) if uctr is not None else None
        check(self.lib.chub_seed_explicit(self.h, _dptr(a[0]), _dptr(a[1]), _dptr(c), _dptr(a[2])))

    # ---- device path ----
    def _f32(self, t):
        import torch
        return _tensor_ptr(t, torch.float32, self.device)

    def _stream(self, stream):
        if stream is not None:
            return ctypes.c_void_p(int(stream))
        return _current_stream(self.device)

    def reset(self, mask=None, step_index=0, stream=None):
        """reset() of every hub (or of hubs whose mask byte is set; mask = CUDA uint8 tensor)."""
        import torch
        if mask is not None and not isinstance(mask, int) and tuple(mask.shape) != (self.n_hubs,):
            raise ValueError("mask must be [%d]" % self.n_hubs)
        m = _tensor_ptr(mask, (torch.uint8, torch.bool), self.device) if mask is not None else None
        check(self.lib.chub_reset(self.h, m, int(step_index), self._stream(stream)))
        return self.tensor("obs")

    def step(self, actions, stream=None):
        """step(action) of every hub.  actions: CUDA float32 [n_hubs, action_stride] in [-1, 1]."""
        if tuple(actions.shape) != (self.n_hubs, self.action_stride):
            raise ValueError("actions must be [%d, %d]" % (self.n_hubs, self.action_stride))  # evcssp_manager.py:148
        check(self.lib.chub_step(self.h, self._f32(actions), self._stream(stream)))
        return self.tensor("obs"), self.tensor("reward"), self.tensor("done")

    def step_into(self, actions, obs_out=None, reward_out=None, done_out=None, stream=None):
        """step(action) with obs / reward / done written straight into the caller's CUDA tensors (float32 [H, 13],
        float32 [H], uint8 [H]; contiguous) -- e.g. the slices of a rollout buffer.  None keeps the engine's own."""
        import torch
        if tuple(actions.shape) != (self.n_hubs, self.action_stride):
            raise ValueError("actions must be [%d, %d]" % (self.n_hubs, self.action_stride))
        for t, shape, dt in ((obs_out, (self.n_hubs, _lib.OBS_DIM), torch.float32), (reward_out, (self.n_hubs,), torch.float32),
                             (done_out, (self.n_hubs,), torch.uint8)):
            if t is not None and (tuple(t.shape) != shape or t.dtype != dt or not t.is_contiguous()):
                raise ValueError("output tensors must be contiguous %r %s" % (shape, dt))
        check(self.lib.chub_step_into(self.h, self._f32(actions), self._f32(obs_out) if obs_out is not None else None,
                                      self._f32(reward_out) if reward_out is not None else None,
                                      _tensor_ptr(done_out, torch.uint8, self.device) if done_out is not None else None,
                                      self._stream(stream)))

    def step_packed(self, packed, stream=None):
        """step(action) on packed records: CUDA int32/uint8 tensor of 16 bytes per hub, {on-bits lo, on-bits hi, a[-2],
        a[-1]} (include/chub.h chub_step_packed; pack_actions builds them from float rows on the host)."""
        import torch
        if packed.numel() * packed.element_size() != 16 * self.n_hubs:
            raise ValueError("packed must hold 16 bytes per hub")
        check(self.lib.chub_step_packed(self.h, _tensor_ptr(packed, (torch.int32, torch.uint8, torch.float32), self.device),
                                        self._stream(stream)))
        return self.tensor("obs"), self.tensor("reward"), self.tensor("done")

    def step_loads(self, loads, stream=None):
        """The stations' scalar-load step, evs_step(float) (CHS.hpp:1169-1186, 1478-1495).  loads: CUDA float32
        [n_hubs, 4], row = [power target of station 0 in kW, of station 1, a[-2], a[-1]]."""
        if tuple(loads.shape) != (self.n_hubs, 4):
            raise ValueError("loads must be [%d, 4]" % self.n_hubs)
        check(self.lib.chub_step_loads(self.h, self._f32(loads), self._stream(stream)))
        return self.tensor("obs"), self.tensor("reward"), self.tensor("done")

    def step_loads_static(self, loads, obs_out=None, reward_out=None, done_out=None, stream=None):
        """step_loads without any allocation (include/chub.h chub_step_loads_static): the dispatch scratch is the library's
        file-scope device array (at most _lib.STATIC_MAX_HUBS hubs; shared per device -- overlapping calls race).  Returns
        nothing new: results are in the given tensors or, for None, in the engine's own obs / reward / done buffers."""
        import torch
        if tuple(loads.shape) != (self.n_hubs, 4):
            raise ValueError("loads must be [%d, 4]" % self.n_hubs)
        for t, shape, dt in ((obs_out, (self.n_hubs, _lib.OBS_DIM), torch.float32), (reward_out, (self.n_hubs,), torch.float32),
                             (done_out, (self.n_hubs,), torch.uint8)):
            if t is not None and (tuple(t.shape) != shape or t.dtype != dt or not t.is_contiguous()):
                raise ValueError("output tensors must be contiguous %r %s" % (shape, dt))
        check(self.lib.chub_step_loads_static(self.h, self._f32(loads), self._f32(obs_out) if obs_out is not None else None,
                                              self._f32(reward_out) if reward_out is not None else None,
                                              _tensor_ptr(done_out, torch.uint8, self.device) if done_out is not None else None,
                                              self._stream(stream)))

    SITUATION_COLUMNS = ("car", "charge", "emergency", "assign", "power", "soc", "init_soc", "target_soc")

    def situation(self, out=None, stream=None):
        """The stations' `situation` table (CHS.hpp:204-231) as CUDA float32 matrices [n_hubs, S0 + S1] per column
        (station 0's piles first).  Returns (dict column -> view, (S0, S1)); `out` may be a reusable [8, H, S] tensor."""
        import torch
        s0, s1 = ctypes.c_int32(0), ctypes.c_int32(0)
        check(self.lib.chub_situation_width(self.h, ctypes.byref(s0), ctypes.byref(s1)))
        shape = (len(self.SITUATION_COLUMNS), self.n_hubs, s0.value + s1.value)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device="cuda:%d" % self.device)
        if tuple(out.shape) != shape or out.dtype != torch.float32:
            raise ValueError("out must be float32 %r" % (shape,))
        check(self.lib.chub_situation(self.h, self._f32(out), self._stream(stream)))
        return {name: out[k] for k, name in enumerate(self.SITUATION_COLUMNS)}, (s0.value, s1.value)

    def capture_rollout(self, actions, n_steps, stream=None):
        if tuple(actions.shape) != (self.n_hubs, self.action_stride):
            raise ValueError("actions must be [%d, %d]" % (self.n_hubs, self.action_stride))
        check(self.lib.chub_rollout_capture(self.h, self._f32(actions), int(n_steps), self._stream(stream)))

    def rollout(self, stream=None):
        check(self.lib.chub_rollout(self.h, self._stream(stream)))

    def tensor(self, name):
        """torch view (DLPack, zero copy) of an engine buffer: obs, reward, done, stats, last_stats, error."""
        if name in self._tensors:
            return self._tensors[name]
        import torch
        which = {"obs": _lib.BUF_OBS, "reward": _lib.BUF_REWARD, "done": _lib.BUF_DONE, "stats": _lib.BUF_STATS,
                 "last_stats": _lib.BUF_LAST_STATS, "error": _lib.BUF_ERROR, "final_soc": _lib.BUF_FINAL_SOC}[name]
        p = ctypes.c_void_p()
        check(self.lib.chub_tensor(self.h, which, ctypes.byref(p)))
        ctypes.pythonapi.PyCapsule_New.restype = ctypes.py_object
        ctypes.pythonapi.PyCapsule_New.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p]
        cap = ctypes.pythonapi.PyCapsule_New(p, b"dltensor", None)
        t = torch.from_dlpack(cap)
        self._tensors[name] = t
        return t

    # ---- host path (what a CPU caller of the reference has) ----
    def host_buffers(self):
        """Pinned numpy arrays: actions [H, stride] f32, obs [H, 13] f32, reward [H] f32, done [H] u8."""
        if self._host is None:
            def pinned(n, ct, dt, shape):
                p = ctypes.c_void_p()
                check(self.lib.chub_alloc_host(n * ctypes.sizeof(ct), ctypes.byref(p)))
                a = np.ctypeslib.as_array((ct * n).from_address(p.value)).view(dt).reshape(shape)
                return p, a
            H, S = self.n_hubs, self.action_stride
            pa, a = pinned(H * S, ctypes.c_float, np.float32, (H, S))
            po, o = pinned(H * _lib.OBS_DIM, ctypes.c_float, np.float32, (H, _lib.OBS_DIM))
            pr, r = pinned(H, ctypes.c_float, np.float32, (H,))
            pd, d = pinned(H, ctypes.c_uint8, np.uint8, (H,))
            self._host = dict(ptrs=[pa, po, pr, pd], actions=a, obs=o, reward=r, done=d)
        return self._host

    def pinned_actions(self):
        """One more pinned [n_hubs, action_stride] float32 array (kept alive by the engine) for callers that rotate
        their action buffers; pass it to step_host(..., copy=False)."""
        p = ctypes.c_void_p()
        n = self.n_hubs * self.action_stride
        check(self.lib.chub_alloc_host(n * 4, ctypes.byref(p)))
        a = np.ctypeslib.as_array((ctypes.c_float * n).from_address(p.value)).reshape(self.n_hubs, self.action_stride)
        self.host_buffers()["ptrs"].append(p)
        return a

    def step_host(self, actions=None, stream=0, copy=True):
        """Host-buffer step: action rows from host memory, step, obs/reward/done back to pinned host arrays, synchronise.
        actions=None steps on host_buffers()["actions"]; copy=False hands the given C-contiguous float32 array to the
        engine as it is (pinned_actions() arrays, or any host memory) instead of copying it into the engine's own."""
        hb = self.host_buffers()
        src = hb["actions"]
        if actions is not None:
            if copy:
                hb["actions"][...] = actions
            else:
                if actions.dtype != np.float32 or not actions.flags["C_CONTIGUOUS"] or actions.shape != src.shape:
                    raise ValueError("copy=False needs a C-contiguous float32 [%d, %d] array" % src.shape)
                src = actions
        check(self.lib.chub_step_host(self.h, _dptr(src), _dptr(hb["obs"]), _dptr(hb["reward"]),
                                      _dptr(hb["done"]), ctypes.c_void_p(int(stream))))
        return hb["obs"], hb["reward"], hb["done"]

    def pinned_records(self):
        """A pinned uint32 [n_hubs, 4] array for step_host_packed (kept alive by the engine)."""
        p = ctypes.c_void_p()
        check(self.lib.chub_alloc_host(self.n_hubs * 16, ctypes.byref(p)))
        a = np.ctypeslib.as_array((ctypes.c_uint32 * (self.n_hubs * 4)).from_address(p.value)).reshape(self.n_hubs, 4)
        self.host_buffers()["ptrs"].append(p)
        return a

    def step_host_packed(self, records, stream=0):
        """Host-buffer step on ready-made 16-byte records (uint32 [n_hubs, 4] host array: on-bits lo, hi, a[-2], a[-1] as
        float bits -- pack_actions() makes them): records host->device, step, obs/reward/done back to the pinned arrays."""
        hb = self.host_buffers()
        if records.dtype != np.uint32 or not records.flags["C_CONTIGUOUS"] or records.shape != (self.n_hubs, 4):
            raise ValueError("records must be a C-contiguous uint32 [%d, 4] array" % self.n_hubs)
        check(self.lib.chub_step_host_packed(self.h, _dptr(records), _dptr(hb["obs"]), _dptr(hb["reward"]),
                                             _dptr(hb["done"]), ctypes.c_void_p(int(stream))))
        return hb["obs"], hb["reward"], hb["done"]

    def set_host_threads(self, n_threads, n_ranges=0):
        """Threads that threshold the action rows inside step_host (0: float rows cross PCIe unpacked) and the number of
        hub ranges of its pipeline (0: keep)."""
        check(self.lib.chub_set_host_threads(self.h, int(n_threads), int(n_ranges)))

    def pack_actions(self, actions):
        """Host float rows [n_hubs, action_stride] -> uint32 [n_hubs, 4] records of step_packed."""
        a = np.ascontiguousarray(actions, np.float32)
        if a.shape != (self.n_hubs, self.action_stride):
            raise ValueError("actions must be [%d, %d]" % (self.n_hubs, self.action_stride))
        out = np.empty((self.n_hubs, 4), np.uint32)
        check(self.lib.chub_pack_actions(self.h, _dptr(a), _dptr(out)))
        return out

    def reset_host(self, stream=0):
        hb = self.host_buffers()
        check(self.lib.chub_reset_host(self.h, _dptr(hb["obs"]), ctypes.c_void_p(int(stream))))
        return hb["obs"]

    # ---- read-back ----
    def hub_report(self, hub=0, stream=0):
        """One hub's scalars in one call (include/chub.h chub_hub_report): dict of the diagnostics columns plus
        st_min_power / st_charge_power / st_max_power / st_line / st_flow_in (pairs), t, store, fcev_arrive, fcev_qn, error."""
        out = np.empty(_lib.REPORT_LEN, np.float64)
        check(self.lib.chub_hub_report(self.h, int(hub), _dptr(out), ctypes.c_void_p(int(stream))))
        rep = {name: float(out[k]) for k, name in enumerate(_lib.DIAG)}
        rep.update(st_min_power=(float(out[14]), float(out[17])), st_charge_power=(float(out[15]), float(out[18])),
                   st_max_power=(float(out[16]), float(out[19])), st_line=(int(out[20]), int(out[21])),
                   st_flow_in=(int(out[22]), int(out[23])), t=int(out[24]), store=float(out[25]),
                   fcev_arrive=int(out[26]), fcev_qn=int(out[27]), error=int(out[28]))
        return rep

    def get(self, name):
        fid, dt, shp = _lib.FIELDS[name]
        out = np.empty((self.n_hubs,) + shp, dtype=dt)
        check(self.lib.chub_get_field(self.h, fid, _dptr(out)))
        return out

    def counters(self):
        s, l, e = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_uint32()
        check(self.lib.chub_counters(self.h, ctypes.byref(s), ctypes.byref(l), ctypes.byref(e)))
        return dict(step_index=s.value, launches=l.value, error_or=e.value)

    # ---- the one collective; timers; parity instrument ----
    def gather_stats(self, comm_ptr, root=0, hubs_per_rank=None, out=None, stream=None):
        """Start the gather of the last-finished-episode statistics on rank `root` (include/chub.h chub_gather_stats).
        comm_ptr: ncclComm_t as an int (torch: ProcessGroupNCCL._comm_ptr()).  out: float32 CUDA [sum of hubs, 4] on the
        root.  The transfer runs on the engine's communication stream; gather_wait() joins it."""
        hp = np.ascontiguousarray(hubs_per_rank, np.int32) if hubs_per_rank is not None else None
        check(self.lib.chub_gather_stats(self.h, ctypes.c_void_p(int(comm_ptr)), int(root), _dptr(hp),
                                         self._f32(out) if out is not None else None, self._stream(stream)))

    def gather_wait(self, stream=None):
        check(self.lib.chub_gather_wait(self.h, self._stream(stream)))

    def reduce_stats(self, comm_ptr=None, out=None, stream=None):
        """Column sums of the statistics over all hubs of all ranks (fp64 [4], CUDA); comm_ptr None = local hubs only."""
        import torch
        if out is None:
            out = torch.empty(4, dtype=torch.float64, device="cuda:%d" % self.device)
        check(self.lib.chub_reduce_stats(self.h, ctypes.c_void_p(int(comm_ptr)) if comm_ptr else None,
                                         _tensor_ptr(out, torch.float64, self.device), self._stream(stream)))
        return out

    def checks(self):
        """(compiled_in, failures) of the self-checking build (include/chub.h chub_checks)."""
        c, n = ctypes.c_int32(0), ctypes.c_int64(0)
        check(self.lib.chub_checks(self.h, ctypes.byref(c), ctypes.byref(n)))
        return bool(c.value), int(n.value)

    def set_timing(self, on=True):
        check(self.lib.chub_set_timing(self.h, int(bool(on))))

    def step_times(self, max_n=128):
        """Durations (ms, oldest first) of the most recent step-kernel launches, from CUDA events on their own stream."""
        out, n = np.zeros(max_n, np.float32), ctypes.c_int32(0)
        check(self.lib.chub_step_times(self.h, _dptr(out), int(max_n), ctypes.byref(n)))
        return out[: n.value].copy()

    def debug_curve(self, which, first_bits, n, out=None, stream=None):
        """The curves as the step kernel compiles them, on n consecutive float32 bit patterns (chub_debug_curve)."""
        import torch
        if out is None:
            out = torch.empty(int(n), dtype=torch.float32, device="cuda:%d" % self.device)
        check(self.lib.chub_debug_curve(self.h, int(which), ctypes.c_uint32(int(first_bits)), int(n), self._f32(out),
                                        self._stream(stream)))
        return out

    def get_state(self):
        n = int(self.lib.chub_state_bytes(self.h))
        blob = np.empty(n, np.uint8)
        check(self.lib.chub_get_state(self.h, _dptr(blob)))
        return blob

    def set_state(self, blob):
        blob = np.ascontiguousarray(blob, np.uint8)
        if blob.size != int(self.lib.chub_state_bytes(self.h)):
            raise ValueError("state blob size mismatch")
        check(self.lib.chub_set_state(self.h, _dptr(blob)))


__all__ = ["HubEngine", "ChubError", "load_tables", "make_hub_class"]
