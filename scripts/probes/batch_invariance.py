#!/usr/bin/env python
"""Where do a batched forward and the single-sample forwards of the same inputs differ?  (Q4_K turbo DiT, T = 6000)
    python scripts/probes/batch_invariance.py [quant] [T]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]
from tools import fixtures as fx  # noqa: E402
from acestep_b200.bridge import B200Bridge  # noqa: E402
from parity_common import metrics, turbo_dit_dir  # noqa: E402

quant = sys.argv[1] if len(sys.argv) > 1 else "Q4_K"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 6000
br = B200Bridge()
br.load_dit(turbo_dit_dir(quant))
inp = [fx.dit_inputs(fx.TURBO, T, 512, seed=1234 + i) for i in range(3)]
h = np.stack([i["hidden_states"] for i in inp]); c = np.stack([i["context_latents"] for i in inp]); e = np.stack([i["encoder_hidden_states"] for i in inp])


def describe(tag, a, b):
    d = np.abs(a.astype(np.float64) - b)
    per_frame = d.max(axis=-1)
    bad = np.nonzero(per_frame > 1e-5)[0]
    print(f"  {tag}: rel_l2 {metrics(a, b)['rel_l2']:.3e} max_abs {d.max():.3e} frames differing {len(bad)}/{len(per_frame)}"
          + (f" first {bad[0]} last {bad[-1]} argmax {int(per_frame.argmax())}" if len(bad) else ""), flush=True)


def run(ts, label):
    print(f"== {label} ts={ts} env: " + " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("ACE_")), flush=True)
    b1 = br.dit_forward_batched(h, c, e, None, None, ts, ts)
    b2 = br.dit_forward_batched(h, c, e, None, None, ts, ts)
    for s in range(3):
        describe(f"batched run 1 vs run 2, sample {s}", b2[s], b1[s])
    singles = [br.dit_forward_tfirst(h[s], c[s], e[s], None, None, ts[s], ts[s]) for s in range(3)]
    singles2 = [br.dit_forward_tfirst(h[s], c[s], e[s], None, None, ts[s], ts[s]) for s in range(3)]
    for s in range(3):
        describe(f"single run 1 vs run 2, sample {s}", singles2[s], singles[s])
        describe(f"batched vs single, sample {s}", b1[s], singles[s])
    # order swapped: sample 2 first in the batch
    perm = [2, 0, 1]
    bp = br.dit_forward_batched(h[perm], c[perm], e[perm], None, None, [ts[i] for i in perm], [ts[i] for i in perm])
    for k, s in enumerate(perm):
        describe(f"permuted batch slot {k} (sample {s}) vs single", bp[k], singles[s])


run([1.0, 0.7, 0.3], "full depth")
run([0.5, 0.5, 0.5], "full depth, equal t")
for k in (1, 2, 4):
    os.environ["ACE_GGML_DIT_MAX_LAYERS"] = str(k)
    run([1.0, 0.7, 0.3], f"{k} layer(s)")
del os.environ["ACE_GGML_DIT_MAX_LAYERS"]
os.environ["ACE_B200_NORM_ROWS"] = "1"
run([1.0, 0.7, 0.3], "full depth, one row per RMSNorm CTA")
del os.environ["ACE_B200_NORM_ROWS"]
os.environ["ACE_B200_PDL"] = "0"
run([1.0, 0.7, 0.3], "full depth, no PDL")
