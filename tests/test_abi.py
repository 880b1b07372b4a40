"""CPU: the C-ABI library loads, exports every symbol include/chub.h declares, and refuses to run
without a CUDA device (there is no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "chub.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(chub_[a-z_0-9]+)\s*\(", src)
    return sorted(set(names))


def test_header_and_binding_agree():
    from charginghub_env_b200 import _lib
    decl = declared_functions()
    assert len(decl) >= 25
    assert sorted(_lib.SYMBOLS) == decl


def test_library_exports_every_declared_symbol():
    from charginghub_env_b200 import _lib
    lib = _lib.load()  # binds every symbol; raises AttributeError if one is missing
    for name in declared_functions():
        assert hasattr(lib, name), name
    assert lib.chub_version() >= 100


def test_no_gpu_means_error_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from charginghub_env_b200 import ChubError, HubEngine, _lib
    with pytest.raises(ChubError) as ei:
        HubEngine(8)
    assert "no CUDA device" in str(ei.value) or "CUDA" in str(ei.value)
    lib = _lib.load()
    cfg = _lib.Config(n_hubs=8, device=0, hub_id_base=0, n_classes=1, action_stride=47)
    h = ctypes.c_void_p()
    assert lib.chub_create(ctypes.byref(cfg), ctypes.byref(h)) == -2  # CHUB_ERR_CUDA
    assert not h.value


def test_argument_validation_without_gpu():
    from charginghub_env_b200 import _lib, make_hub_class
    lib = _lib.load()
    assert lib.chub_create(None, None) == -1
    assert b"null" in lib.chub_last_error()
    with pytest.raises(ValueError):
        make_hub_class(station_list=[1, 2, 3])
    with pytest.raises(ValueError):
        make_hub_class(station_type_list=["fast", "turbo"])
    with pytest.raises(ValueError):
        make_hub_class(el_table=[0.0] * 5)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import, link or exec it."""
    pkg = os.path.join(ROOT, "charginghub_env_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower(), (dirpath, f)
