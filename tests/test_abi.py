"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads, and exports every symbol
include/lgb200/lgb200.h declares; error paths that need no GPU behave as documented."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "lgb200", "lgb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lg_[a-z_0-9]+)\s*\(", text)) - {"lg_unit_vector_fn"})


def test_header_symbols_exported(lib):
    from libgran_b200.system import ABI_SYMBOLS

    names = declared_symbols()
    assert len(names) >= 25
    for name in names:
        assert hasattr(lib, name), f"{name} declared in lgb200.h but not exported"
    assert sorted(ABI_SYMBOLS) == names, "python binding list out of sync with the header"


def test_abi_version(lib):
    assert lib.lg_abi_version() == 1


def test_no_cpu_fallback(lib):
    """Without a CUDA device lg_create must fail loudly (LG_ENODEVICE); it must never compute on the CPU."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.lg_create(4, 1.0, 1.0, 1.0, -1, C.byref(h))
    assert rc == -2
    assert b"no CPU fallback" in lib.lg_last_error()
    assert not h.value


def test_bad_arguments(lib):
    h = C.c_void_p()
    assert lib.lg_create(0, 1.0, 1.0, 1.0, -1, C.byref(h)) == -1
    assert lib.lg_create(4, -1.0, 1.0, 1.0, -1, C.byref(h)) == -1
    assert lib.lg_do_step(None, 1e-13) == -1
    assert lib.lg_destroy(None) == 0


def test_halo_bad_arguments(lib):
    """lg_halo_*: argument errors are reported without touching a device."""
    hl = C.c_void_p()
    assert lib.lg_halo_create(None, 1, (C.c_int64 * 1)(8), C.byref(hl)) == -1
    assert lib.lg_halo_exchange(None, None) == -1
    assert lib.lg_run_slab(None, None, 1e-13, 1, 0) == -1
    assert lib.lg_halo_set_lists(None, 0, 0, None, 0, None, 0, None, 0, None) == -1
    assert lib.lg_halo_export(None, 0, C.create_string_buffer(64), (C.c_int64 * 4)()) == -1
    assert lib.lg_halo_connect(None, 0, C.create_string_buffer(64), (C.c_int64 * 4)()) == -1
    assert lib.lg_halo_disconnect(None) == 0 and lib.lg_halo_destroy(None) == 0


def test_size_mismatch_exception(lib):
    """rotational_system.h:69-70: differing x0 / v0 / theta0 / omega0 sizes throw SizeMismatchException."""
    from libgran_b200.system import GranularSystem, SizeMismatchException

    with pytest.raises(SizeMismatchException):
        GranularSystem(np.zeros((3, 3)), np.zeros((2, 3)), np.zeros((3, 3)), np.zeros((3, 3)), 1.0, 1.0, 1.0)


def test_product_does_not_import_oracle():
    """The product package must not reference anything under oracle/."""
    pkg = os.path.join(ROOT, "libgran_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in txt and "from oracle" not in txt and "dem_oracle" not in txt, f
    for dirpath, _, files in os.walk(os.path.join(ROOT, "include")):
        for f in files:
            txt = open(os.path.join(dirpath, f), errors="replace").read()
            assert "dem_oracle" not in txt and "oracle/" not in txt, f
