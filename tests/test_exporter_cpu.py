"""CPU: the in-tree exporter (tools/export_safetensors_to_gguf.py, the library's own bit-identical quantisers) writes the same
tensors as the reference exporter's rule with ggml_quantize_chunk, and the unmodified reference loads and runs the exported file."""
import os

import numpy as np
import pytest

from tools import export_safetensors_to_gguf as exporter
from tools import fixtures as fx
from oracle import refbridge
from parity_common import metrics


def raw_tensors(path):
    import gguf
    return {t.name: (int(t.tensor_type), np.asarray(t.data).tobytes()) for t in gguf.GGUFReader(path).tensors}


@pytest.mark.parametrize("quant", ["F16", "Q8_0", "Q4_K", "Q6_K"])
def test_exported_gguf_equals_reference_rule(tmp_path, quant):
    if quant != "F16" and not refbridge.ref_available():
        pytest.skip("comparison file needs ggml_quantize_chunk from oracle/_ref")
    d = fx.make_dit_fixture(str(tmp_path / "dit"), fx.TOY, 0)
    out = str(tmp_path / "exported.gguf")
    stats = exporter.export(os.path.join(d, "model.safetensors"), out, quant)
    assert stats["quantized"] > 0 or quant == "F16"
    cmp_path = str(tmp_path / "cmp.gguf")
    # fast=True: every quantised type through ggml_quantize_chunk of the reference library (what the reference exporter calls)
    fx.save_gguf(cmp_path, fx.dit_tensors(fx.TOY, 0), quant, refbridge.ref_lib() if quant != "F16" else None, fast=True)
    mine, theirs = raw_tensors(out), raw_tensors(cmp_path)
    assert set(mine) == set(theirs)
    for name in mine:
        assert mine[name] == theirs[name], name  # same ggml type, same bytes


def test_reference_runs_the_exported_file(tmp_path):
    if not refbridge.ref_available():
        pytest.skip("oracle/_ref not built")
    d = fx.make_dit_fixture(str(tmp_path / "dit"), fx.TOY, 0)
    exporter.export(os.path.join(d, "model.safetensors"), os.path.join(d, "model.gguf"), "Q4_K")
    inp = fx.dit_inputs(fx.TOY, 20, 5, seed=3)
    calls = [{"op": "dit_forward", "t": 0.5, "t_r": 0.5, "out": "v"}]
    v_gguf, _ = refbridge.run_ref(calls, inp, dit_dir=d)
    # the same file written with the reference exporter's own dependency (ggml_quantize_chunk) gives the same bytes, hence the same output
    fx.save_gguf(os.path.join(d, "model.gguf"), fx.dit_tensors(fx.TOY, 0), "Q4_K", refbridge.ref_lib(), fast=True)
    v_ref_export, _ = refbridge.run_ref(calls, inp, dit_dir=d)
    assert np.isfinite(v_gguf["v"]).all() and metrics(v_gguf["v"], v_ref_export["v"])["rel_l2"] <= 1e-6
