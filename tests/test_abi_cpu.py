"""CPU: the C-ABI shared library loads without a GPU and exports every symbol include/*.h declares; the ctypes table of the
bridge names exactly those symbols; entry points that need no device behave as the reference's do
(acestep_ggml.cpp:231-236 last_error on NULL; acestep_ggml.h:23-29 status values)."""
import ctypes
import os
import re

import numpy as np

from acestep_b200 import bridge

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for h in ("acestep_ggml.h", "acestep_b200.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names |= set(re.findall(r"\b(ace_(?:ggml|b200)_[a-z0-9_]+)\s*\(", text))
    return names


def test_every_declared_symbol_is_exported_and_bound():
    decl = declared_symbols()
    assert len(decl) >= 30
    lib = ctypes.CDLL(str(bridge.DEFAULT_LIB))
    missing = [n for n in sorted(decl) if not hasattr(lib, n)]
    assert not missing, missing
    assert set(bridge.SIGNATURES) == decl, (sorted(decl - set(bridge.SIGNATURES)), sorted(set(bridge.SIGNATURES) - decl))


def test_no_device_entry_points():
    lib = bridge.load_library()
    assert lib.ace_ggml_last_error(None).decode().endswith("null context")
    assert lib.ace_b200_kernel_launches() == 0  # nothing ran: there is no CPU compute path to fall back to
    out = np.zeros(8, np.float32)
    assert lib.ace_b200_shift_schedule(ctypes.c_float(3.0), 8, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))) == 0
    assert np.allclose(out, [1.0, 0.9545454545, 0.9, 0.8333333333, 0.75, 0.6428571429, 0.5, 0.3], atol=1e-7)
    assert lib.ace_b200_shift_schedule(ctypes.c_float(3.0), 0, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))) == 2  # INVALID_ARG
    assert lib.ace_ggml_dit_forward(None, None, None, None, None, None, 4, 0, ctypes.c_float(0.5), ctypes.c_float(0.5), None, 0) == 2
