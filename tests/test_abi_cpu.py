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


def _probe(path, gguf=True):
    lib = bridge.load_library()
    buf = ctypes.create_string_buffer(512)
    n = lib.ace_b200_debug_probe_tensor_file(str(path).encode(), 1 if gguf else 0, buf, 512)
    return int(n), buf.value.decode()


def _gguf_bytes(tensors, n_kv=0, version=3):
    """minimal GGUF writer: tensors = [(name, dims(list, ggml order), type, offset)]; data section = 4 KiB of zeros"""
    import struct
    b = struct.pack("<IIQQ", 0x46554747, version, len(tensors), n_kv)
    for name, dims, typ, off in tensors:
        nb = name.encode()
        b += struct.pack("<Q", len(nb)) + nb + struct.pack("<I", len(dims))
        for d in dims:
            b += struct.pack("<Q", d & 0xFFFFFFFFFFFFFFFF)
        b += struct.pack("<IQ", typ, off & 0xFFFFFFFFFFFFFFFF)
    b += b"\0" * ((-len(b)) % 32)
    return b + b"\0" * 4096


def test_malformed_weight_files_are_rejected(tmp_path):
    """crafted headers must not wrap the readers' size arithmetic (offset + nbytes, es * n, negative / huge dims)"""
    import struct
    ok = tmp_path / "ok.gguf"
    ok.write_bytes(_gguf_bytes([("a", [64, 4], 0, 0), ("b", [32, 2], 8, 1024)]))
    assert _probe(ok)[0] == 2
    cases = {
        "huge_dim": [("a", [1 << 62, 4], 0, 0)],
        "negative_dim": [("a", [-64, 4], 0, 0)],
        "offset_wraps": [("a", [64, 4], 0, (1 << 64) - 512)],
        "past_end": [("a", [64, 4], 0, 4096 - 64)],
        "numel_wraps": [("a", [1 << 33, 1 << 33], 0, 0)],
        "bad_block": [("a", [33, 2], 8, 0)],
    }
    for name, t in cases.items():
        p = tmp_path / f"{name}.gguf"
        p.write_bytes(_gguf_bytes(t))
        n, msg = _probe(p)
        assert n == -1 and msg, (name, n, msg)
    # implausible counts / an array KV whose length would wrap es * n
    p = tmp_path / "counts.gguf"
    p.write_bytes(struct.pack("<IIQQ", 0x46554747, 3, 1 << 60, 0) + b"\0" * 64)
    assert _probe(p)[0] == -1
    key = b"k"
    kv = struct.pack("<Q", len(key)) + key + struct.pack("<IIQ", 9, 6, (1 << 62) + 1)  # array of f32, n * 4 wraps to 4
    p = tmp_path / "kvwrap.gguf"
    p.write_bytes(struct.pack("<IIQQ", 0x46554747, 3, 0, 1) + kv + b"\0" * 64)
    assert _probe(p)[0] == -1
    # safetensors: negative dim / offsets out of range
    import json
    for hdr in ({"a": {"dtype": "F32", "shape": [-4, 2], "data_offsets": [0, 32]}}, {"a": {"dtype": "F32", "shape": [4, 2], "data_offsets": [0, 1 << 40]}}):
        hj = json.dumps(hdr).encode()
        p = tmp_path / "bad.safetensors"
        p.write_bytes(struct.pack("<Q", len(hj)) + hj + b"\0" * 64)
        assert _probe(p, gguf=False)[0] == -1
