"""TEST INFRASTRUCTURE (oracle). Spawns oracle/ref_runner.py under the aligned-malloc
LD_PRELOAD shim so that the unmodified reference library can run on glibc
(ggml-backend.cpp:2265). Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module."""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile
from typing import Dict, List, Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def _cpu_flags() -> set:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def ref_isa() -> str:
    need = {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512_vnni", "avx512vbmi"}
    return "avx512" if need <= _cpu_flags() else "avx2"


def ref_lib(isa: Optional[str] = None) -> str:
    return os.path.join(REF_DIR, isa or ref_isa(), "libacestep_ggml.so")


def ref_available() -> bool:
    return os.path.exists(ref_lib()) and os.path.exists(os.path.join(REF_DIR, "libalignshim.so"))


def run_ref(calls: List[dict], inputs: Dict[str, np.ndarray], *, dit_dir: Optional[str] = None,
            vae_dir: Optional[str] = None, text_dir: Optional[str] = None, threads: int = 4, compute_buffer_bytes: int = 1 << 30,
            env: Optional[Dict[str, str]] = None, timeout: float = 3600.0, isa: Optional[str] = None):
    """Run the reference. Returns (outputs dict, timings dict)."""
    if not ref_available():
        raise RuntimeError("oracle/_ref not built: run `make -C oracle ref` (needs /root/reference)")
    with tempfile.TemporaryDirectory(prefix="ace_ref_") as td:
        inp = os.path.join(td, "in.npz")
        out = os.path.join(td, "out.npz")
        np.savez(inp, **inputs)
        job = {"lib": ref_lib(isa), "dit_dir": dit_dir, "vae_dir": vae_dir, "text_dir": text_dir, "threads": threads,
               "compute_buffer_bytes": int(compute_buffer_bytes), "inputs": inp, "out": out, "calls": calls}
        jp = os.path.join(td, "job.json")
        json.dump(job, open(jp, "w"))
        e = dict(os.environ)
        e["LD_PRELOAD"] = os.path.join(REF_DIR, "libalignshim.so")
        e.update(env or {})
        r = subprocess.run([sys.executable, os.path.join(HERE, "ref_runner.py"), jp], env=e,
                           capture_output=True, text=True, timeout=timeout)
        if r.returncode != 0:
            raise RuntimeError(f"reference run failed rc={r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}")
        outs = dict(np.load(out))
        tim = json.load(open(out + ".timings.json"))
        tim["stderr"] = r.stderr[-4000:]
        return outs, tim
