"""TEST INFRASTRUCTURE ONLY.  Loader for the UNMODIFIED reference environment in this container.

Imports `evcssp_env_cpp.envs.evcssp_manager.EvcsspManagerEnv_v6` straight from the read-only
checkout at /root/reference, with two stand-ins the image lacks:
  * `pyevstation`  -> oracle/_ref/pyevstation*.so  (pybind11 shim over the unmodified CHS.hpp,
                      see oracle/ref_shim.cpp; the reference binds it with Boost.Python, main.cpp:19-290)
  * `gym`          -> a ~30 line stub (gym 0.18.3 is not installed and there is no network)
It cannot run on the GPU box (no /root/reference there): it is used here to validate the C
restatement (oracle/hub_oracle.c) and to generate tests/golden/*.npz (oracle/gen_golden.py).
"""
import contextlib
import importlib
import importlib.util
import os
import sys
import types

REF_ROOT = "/root/reference"
_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, "evcssp_env_cpp"))


def shim_path():
    if not os.path.isdir(_REF_DIR):
        return None
    for f in sorted(os.listdir(_REF_DIR)):
        if f.startswith("pyevstation") and f.endswith(".so"):
            return os.path.join(_REF_DIR, f)
    return None


@contextlib.contextmanager
def quiet():
    """The reference prints from C++ (std::cout) and Python inside reset/step; silence fd 1."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    try:
        yield
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)


class capture(object):
    """Like quiet(), but fd 1 goes to a temporary file that can be read back incrementally: the reference reports
    departures below target only by printing (CHS.hpp:916-921, 1081-1086), so counting them means reading its stdout."""

    def __enter__(self):
        import tempfile
        sys.stdout.flush()
        self._saved = os.dup(1)
        self._tmp = tempfile.TemporaryFile(mode="w+b")
        os.dup2(self._tmp.fileno(), 1)
        self._pos = 0
        return self

    def new_text(self):
        sys.stdout.flush()
        self._tmp.seek(self._pos)
        data = self._tmp.read()
        self._pos += len(data)
        return data.decode("utf-8", "replace")

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self._saved, 1)
        os.close(self._saved)
        self._tmp.close()
        return False


def load_shim(csv_dir=None):
    """Import the pybind11 `pyevstation`.  CHS.hpp:175 reads the arrival CSV from the CWD at load."""
    if "pyevstation" in sys.modules:
        return sys.modules["pyevstation"]
    path = shim_path()
    if path is None:
        raise RuntimeError("oracle/_ref/pyevstation*.so missing: run `make -C oracle ref`")
    cwd = os.getcwd()
    if csv_dir is None and reference_available():
        csv_dir = os.path.join(REF_ROOT, "test")
    try:
        if csv_dir:
            os.chdir(csv_dir)
        with quiet():
            spec = importlib.util.spec_from_file_location("pyevstation", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
    finally:
        os.chdir(cwd)
    sys.modules["pyevstation"] = mod
    sys.modules["lion_cpp20.pyevstation"] = mod
    return mod


def _install_gym_stub():
    if "gym" in sys.modules:
        return
    import numpy as np

    gym = types.ModuleType("gym")

    class Env(object):
        metadata = {}

    class Box(object):
        def __init__(self, low, high, shape=None, dtype=np.float32):
            self.dtype = dtype
            if shape is None:
                shape = np.shape(low)
            self.shape = tuple(shape)
            self.low = np.broadcast_to(np.asarray(low, dtype=dtype), self.shape).copy()
            self.high = np.broadcast_to(np.asarray(high, dtype=dtype), self.shape).copy()

        def sample(self):
            return np.random.uniform(self.low, self.high).astype(self.dtype)

    spaces = types.ModuleType("gym.spaces")
    spaces.Box = Box
    error = types.ModuleType("gym.error")
    utils = types.ModuleType("gym.utils")
    seeding = types.ModuleType("gym.utils.seeding")

    def np_random(seed=None):
        rs = np.random.RandomState(seed if seed is not None else 0)
        return rs, seed

    seeding.np_random = np_random
    utils.seeding = seeding
    envs = types.ModuleType("gym.envs")
    registration = types.ModuleType("gym.envs.registration")
    registration.registry = {}

    def register(id, **kw):
        registration.registry[id] = kw

    registration.register = register
    envs.registration = registration
    gym.Env, gym.spaces, gym.error, gym.utils, gym.envs = Env, spaces, error, utils, envs
    for name, m in [("gym", gym), ("gym.spaces", spaces), ("gym.error", error), ("gym.utils", utils),
                    ("gym.utils.seeding", seeding), ("gym.envs", envs),
                    ("gym.envs.registration", registration)]:
        sys.modules[name] = m


def load_reference_env_class():
    """Return the reference's EvcsspManagerEnv_v6 class (unmodified source, imported in place)."""
    if not reference_available():
        raise RuntimeError("/root/reference is not present on this machine")
    load_shim()
    _install_gym_stub()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    with quiet():
        mod = importlib.import_module("evcssp_env_cpp.envs.evcssp_manager")
    return mod.EvcsspManagerEnv_v6


DEFAULT_KWARGS = dict(  # test/env_test.py:10-30
    station_list=[20, 25], station_type_list=["fast", "slow"], constant_charging=False,
    seed_rand=False, hydro_prod_rate=100, hydro_store_vlt=25.0, init_soc=0.2, fc_max_power=100,
    fcev_permeate=0.01, use_lagrange=False, renew_fluctuate=0.0, price_fluctuate=0.0, hydro_loss=0.0)
