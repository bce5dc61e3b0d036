"""The drop-in boundary exercised by the reference's OWN callers.

GPU: the reference's command-line driver (cpp/ace_ggml_cli.cpp, compiled unmodified by oracle/Makefile) linked against
libacestep_b200.so (oracle/_ref/b200/ace_ggml_cli) runs the stages of the reference's staged smoke test
(tools/run_staged_pipeline_tests.py:94-270: text encoder, --dit, --vae, --vae --encode, --pipeline, --pipeline-style-lyric,
--pipeline-style-lyric-timbre) next to the same driver linked against the reference library (oracle/_ref/<isa>/ace_ggml_cli, CPU).
The reference test only checks exit codes; here the printed values and the written WAV files are compared as well."""
import os
import re
import subprocess
import wave

import numpy as np
import pytest

from tools import fixtures as fx
from oracle import refbridge
from parity_common import make_vae_dir, metrics

pytestmark = pytest.mark.gpu
REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
CLI_B200 = os.path.join(REF, "b200", "ace_ggml_cli")


def cli_ref():
    return os.path.join(REF, refbridge.ref_isa(), "ace_ggml_cli")


@pytest.fixture(scope="module")
def dirs(tmp_path_factory):
    if not (os.path.exists(CLI_B200) and os.path.exists(cli_ref())):
        pytest.fail("oracle/_ref CLIs are not built (make -C oracle ref)")
    root = str(tmp_path_factory.mktemp("cli_fixtures"))
    return {"qwen": fx.make_qwen_fixture(os.path.join(root, "qwen"), fx.QWEN_TOY, 0),
            "dit": fx.make_dit_fixture(os.path.join(root, "dit"), fx.TOY_COND, 0), "vae": make_vae_dir(root), "root": root}


def run(binary, args, shim):
    env = dict(os.environ)
    if shim:
        env["LD_PRELOAD"] = os.path.join(REF, "libalignshim.so")
    r = subprocess.run([binary] + [str(a) for a in args], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, (binary, args, r.stdout[-500:], r.stderr[-1500:])
    head = [l for l in r.stdout.splitlines() if l.startswith("ok ")]
    vals = [float(m.group(1)) for m in (re.match(r"^\d+: (-?[\d.]+(?:e-?\d+)?)$", l) for l in r.stdout.splitlines()) if m]
    return head[0] if head else "", np.asarray(vals)


def both(args, wav=None, root=None):
    outs = []
    for binary, shim, tag in ((cli_ref(), True, "ref"), (CLI_B200, False, "b200")):
        a = list(args)
        if wav:
            a += ["--out-wav", os.path.join(root, f"{wav}_{tag}.wav")]
        outs.append(run(binary, a, shim))
    return outs


def read_wav(path):
    with wave.open(path, "rb") as w:
        return np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16).astype(np.float32) / 32768.0


def test_cli_text_encoder_dit_vae(dirs):
    (h0, v0), (h1, v1) = both(["--text-encoder", dirs["qwen"], "--tokens", "5,9,300,17,44,2,511", "--threads", 4])
    assert h0 == h1 and len(v0) == 8 and np.abs(v0 - v1).max() <= 5e-3 * max(1.0, np.abs(v0).max())
    (h0, v0), (h1, v1) = both(["--text-encoder", dirs["qwen"], "--tokens", "5,9,300,17,44", "--mask", "1,1,1,0,0", "--layers", 2, "--threads", 4])
    assert h0 == h1 and np.abs(v0 - v1).max() <= 5e-3 * max(1.0, np.abs(v0).max())
    (h0, v0), (h1, v1) = both(["--dit", dirs["dit"], "--seq-len", 33, "--enc-len", 7, "--timestep", 0.5, "--timestep-r", 0.5, "--seed", 3, "--threads", 4,
                               "--compute-buffer-mb", 512])
    assert h0 == h1 and len(v0) == 8 and metrics(v1, v0)["rel_l2"] <= 5e-3, (v0, v1)
    (h0, v0), (h1, v1) = both(["--vae", dirs["vae"], "--latent-len", 16, "--seed", 1, "--threads", 4, "--compute-buffer-mb", 1024], wav="dec", root=dirs["root"])
    assert h0 == h1
    m = metrics(read_wav(os.path.join(dirs["root"], "dec_b200.wav")), read_wav(os.path.join(dirs["root"], "dec_ref.wav")))
    assert m["cos"] >= 0.999, m
    (h0, v0), (h1, v1) = both(["--vae", dirs["vae"], "--encode", "--audio-len", 48 * 12, "--latent-len", 12, "--seed", 1, "--threads", 4, "--compute-buffer-mb", 1024])
    assert h0 == h1 and np.abs(v0 - v1).max() <= 2e-3 * max(1.0, np.abs(v0).max()), (v0, v1)


@pytest.mark.parametrize("mode", ["pipeline", "pipeline-style-lyric", "pipeline-style-lyric-timbre"])
def test_cli_pipelines(dirs, mode):
    common = ["--text-encoder", dirs["qwen"], "--dit", dirs["dit"], "--vae", dirs["vae"], "--seq-len", 20, "--shift", 3, "--seed", 5, "--threads", 4,
              "--compute-buffer-mb", 1024]
    if mode == "pipeline":
        args = ["--pipeline", "--tokens", "5,9,300,17,44,2,511"] + common
    elif mode == "pipeline-style-lyric":
        args = ["--pipeline-style-lyric", "--style-tokens", "5,9,300,17", "--lyric-tokens", "30,31,200,33,34,470"] + common
    else:
        args = ["--pipeline-style-lyric-timbre", "--style-tokens", "5,9,300,17", "--lyric-tokens", "30,31,200,33,34,470", "--timbre-rand-n", 2, "--timbre-len", 9] + common
    (h0, v0), (h1, v1) = both(args, wav=mode, root=dirs["root"])
    assert h0 == h1 and h0.startswith("ok pipeline")
    a, r = read_wav(os.path.join(dirs["root"], f"{mode}_b200.wav")), read_wav(os.path.join(dirs["root"], f"{mode}_ref.wav"))
    m = metrics(a, r)
    assert a.shape == r.shape and m["cos"] >= 0.999, m
