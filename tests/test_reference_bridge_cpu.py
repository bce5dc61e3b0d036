"""CPU: the reference's own ctypes bridge — scripts/run_non_ggml_real_case.py: GGMLCAPIBridge (:135-354), imported by file path from
/root/reference when that tree is present (it is not on the GPU box) — can dlopen libacestep_b200.so and bind every symbol it uses;
construction then reaches ace_ggml_create, which refuses without a CUDA device (there is no CPU path)."""
import importlib.util
import os
import sys

import pytest

from acestep_b200 import bridge

SCRIPT = "/root/reference/scripts/run_non_ggml_real_case.py"


def load_reference_module():
    spec = importlib.util.spec_from_file_location("ref_run_non_ggml_real_case", SCRIPT)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.skipif(not os.path.exists(SCRIPT), reason="reference tree not present on this machine")
def test_reference_bridge_binds_this_library():
    mod = load_reference_module()
    Bridge = mod.GGMLCAPIBridge
    bound = {}
    orig_bind = Bridge._bind

    def spy(self):
        orig_bind(self)
        bound["lib"] = self.lib

    Bridge._bind = spy
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    try:
        if has_gpu:
            b = Bridge(bridge.DEFAULT_LIB, n_threads=1, compute_buffer_mb=0)
            b.close()
        else:
            with pytest.raises(RuntimeError, match="ace_ggml_create"):
                Bridge(bridge.DEFAULT_LIB, n_threads=1, compute_buffer_mb=0)
    finally:
        Bridge._bind = orig_bind
    assert "lib" in bound  # every argtypes / restype assignment of the reference's _bind resolved a symbol of this library
    # the method names B200Bridge mirrors exist on the reference class
    for name in ("load_text_encoder", "load_dit", "load_vae", "vae_decode_tfirst", "dit_forward_tfirst", "close"):
        assert hasattr(Bridge, name) and hasattr(bridge.B200Bridge, name), name
