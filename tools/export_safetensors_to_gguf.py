#!/usr/bin/env python
"""safetensors -> GGUF exporter for the DiT / VAE / Qwen3 checkpoints this library loads.

Same rule as the reference's exporter (acestep_ggml/tools/export_safetensors_to_gguf.py:213-265): every float tensor with >= 2
dims whose last dim is a whole number of blocks is quantised (Q8_0 / Q6_K / Q4_K), every other float tensor is stored F16, integer
tensors are kept.  The difference: the reference borrows ggml_quantize_chunk from a built libggml; this tool uses the library's OWN
row quantisers through the C ABI (ace_b200_debug_quantize_rows, CPU-only, no GPU needed) — they are bit-identical to ggml's
(tests/test_wpack_cpu.py), so the files are byte-for-byte what the reference exporter writes for the quantised tensors.

    python tools/export_safetensors_to_gguf.py --input <dir>/model.safetensors --output <dir>/model.gguf --quant Q4_K
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200")]

GGML_TYPE = {"Q8_0": 8, "Q4_K": 12, "Q6_K": 14}
BLOCK = {"Q8_0": (32, 34), "Q4_K": (256, 144), "Q6_K": (256, 210)}  # elements, bytes


def read_safetensors(path: str):
    """yields (name, dtype string, shape, raw little-endian view) — 8-byte header length + JSON + data (safetensors.cpp:46-129)"""
    mm = np.memmap(path, dtype=np.uint8, mode="r")
    n = int(np.frombuffer(mm[:8].tobytes(), dtype="<u8")[0])
    header = json.loads(mm[8:8 + n].tobytes())
    base = 8 + n
    for name, e in header.items():
        if name == "__metadata__":
            continue
        o0, o1 = e["data_offsets"]
        yield name, e["dtype"], tuple(int(x) for x in e["shape"]), mm[base + o0: base + o1]


def to_f32(raw: np.ndarray, dtype: str, shape) -> np.ndarray:
    if dtype == "F32":
        return np.frombuffer(raw.tobytes(), dtype="<f4").reshape(shape)
    if dtype == "F16":
        return np.frombuffer(raw.tobytes(), dtype="<f2").astype(np.float32).reshape(shape)
    if dtype == "BF16":
        u = np.frombuffer(raw.tobytes(), dtype="<u2").astype(np.uint32) << 16
        return u.view(np.float32).reshape(shape)
    raise ValueError(f"unsupported dtype {dtype}")


class RowQuantizer:
    def __init__(self):
        from acestep_b200 import bridge
        self.lib = bridge.load_library()

    def quantize(self, w2d: np.ndarray, quant: str) -> np.ndarray:
        w2d = np.ascontiguousarray(w2d, dtype=np.float32)
        rows, k = w2d.shape
        be, bb = BLOCK[quant]
        rb = k // be * bb
        out = np.empty((rows, rb), dtype=np.uint8)

        def part(lo, hi):
            src = w2d[lo:hi]
            dst = out[lo:hi]
            st = self.lib.ace_b200_debug_quantize_rows(src.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), GGML_TYPE[quant], hi - lo, k,
                                                       dst.ctypes.data_as(ctypes.c_void_p), dst.nbytes)
            if st != 0:
                raise RuntimeError(f"ace_b200_debug_quantize_rows failed: {st}")

        n = min(os.cpu_count() or 1, max(1, rows // 32))
        cuts = [rows * i // n for i in range(n + 1)]
        with ThreadPoolExecutor(n) as ex:  # ctypes drops the GIL during the call
            list(ex.map(lambda i: part(cuts[i], cuts[i + 1]), range(n)))
        return out


def export(src: str, dst: str, quant: str, arch: str = "acestep", verbose: bool = False) -> dict:
    import gguf
    quant = {"Q8": "Q8_0", "Q6": "Q6_K", "Q4": "Q4_K", "Q4_K_M": "Q4_K"}.get(quant.upper(), quant.upper())
    if quant not in ("F16", "Q8_0", "Q6_K", "Q4_K"):
        raise ValueError(f"unsupported --quant {quant}")
    rq = RowQuantizer() if quant != "F16" else None
    w = gguf.GGUFWriter(dst, arch=arch)
    w.add_type("model")
    w.add_quantization_version(gguf.GGML_QUANT_VERSION)
    stats = {"quantized": 0, "f16": 0, "other": 0}
    for name, dtype, shape, raw in read_safetensors(src):
        if dtype in ("I32", "I64"):
            w.add_tensor(name, np.frombuffer(raw.tobytes(), dtype="<i4" if dtype == "I32" else "<i8").reshape(shape).copy())
            stats["other"] += 1
            continue
        a = to_f32(raw, dtype, shape)
        if rq is not None and a.ndim >= 2 and shape[-1] > 0 and shape[-1] % BLOCK[quant][0] == 0:
            q = rq.quantize(a.reshape(-1, shape[-1]), quant)
            w.add_tensor(name, q.reshape(tuple(shape[:-1]) + (q.shape[-1],)), raw_dtype=getattr(gguf.GGMLQuantizationType, quant))
            stats["quantized"] += 1
        else:
            w.add_tensor(name, a.astype(np.float16))
            stats["f16"] += 1
        if verbose:
            print(f"{name} {shape} {dtype}", flush=True)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file(progress=False)
    w.close()
    return stats


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--input", required=True, help="model.safetensors / diffusion_pytorch_model.safetensors")
    ap.add_argument("--output", required=True, help="model.gguf")
    ap.add_argument("--quant", default="Q8_0", help="F16 | Q8_0 | Q6_K | Q4_K")
    ap.add_argument("--arch", default="acestep")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    stats = export(a.input, a.output, a.quant, a.arch, a.verbose)
    print(f"[done] output={a.output}")
    print(f"[stats] quantized={stats['quantized']} f16_fallback={stats['f16']} other={stats['other']}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
