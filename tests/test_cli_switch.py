"""The reference's own CLI (unmodified c/main.c + c/*.c) with the backend switch linked in front of its outer API
(adapter/qwen_tts_backend_switch.c, adapter/Makefile -> oracle/_ref/qwen-tts-cuda) - SURVEY.md 8f rank 1.
CPU side: with the switch linked in, the CPU path still produces the plain CLI's bytes, and asking for cuda without a
device falls back to it with a message.  GPU side: QWEN_TTS_BACKEND=cuda produces the same token trace through the
stderr protocol the reference's scripts parse, and the same waveform within one 16-bit step."""
import os
import re
import subprocess
import wave

import numpy as np
import pytest

from conftest import FIXTURE_TOKENS, ROOT
from harness import binding as B, modelgen as mg

CLI = os.path.join(ROOT, "oracle", "_ref", "qwen-tts")
CLI_SWITCH = os.path.join(ROOT, "oracle", "_ref", "qwen-tts-cuda")
pytestmark = pytest.mark.skipif(not (os.path.exists(CLI) and os.path.exists(CLI_SWITCH)), reason="oracle/_ref CLIs not built (make -C oracle ref)")
IDS = ",".join(str(t) for t in FIXTURE_TOKENS)


@pytest.fixture(scope="module")
def model_dir(tmp_path_factory):
    spec = mg.PRESETS["tiny"]
    main, cod = mg.build_tensors(spec, seed=3)
    d = str(tmp_path_factory.mktemp("cli_model"))
    mg.write_model_dir(d, spec, main, cod)
    return d


def run_cli(binary, model_dir, out, backend=None, extra=()):
    env = dict(os.environ)
    env.pop("QWEN_TTS_BACKEND", None)
    if backend:
        env["QWEN_TTS_BACKEND"] = backend
    cmd = [binary, "-d", model_dir, "-t", IDS, "-s", "aiden", "-l", "english", "-o", out, "-v", "-v", "--fixed-codec-tokens", "9", "--seed", "7", *extra]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stderr


def read_wav(path):
    with wave.open(path, "rb") as w:
        assert w.getframerate() == 24000 and w.getnchannels() == 1 and w.getsampwidth() == 2
        return np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16)


def trace(stderr):
    m = re.search(r"Token trace: ([0-9,]+)", stderr)
    assert m, stderr[-1500:]
    return [int(x) for x in m.group(1).split(",")]


def test_switch_keeps_cpu_path_and_falls_back(model_dir, tmp_path):
    a, b, c = str(tmp_path / "a.wav"), str(tmp_path / "b.wav"), str(tmp_path / "c.wav")
    ea = run_cli(CLI, model_dir, a)
    eb = run_cli(CLI_SWITCH, model_dir, b)                       # no QWEN_TTS_BACKEND: the reference's CPU path
    assert open(a, "rb").read() == open(b, "rb").read()
    assert trace(ea) == trace(eb)
    import torch
    if not torch.cuda.is_available():
        ec = run_cli(CLI_SWITCH, model_dir, c, backend="cuda")   # asked for cuda, no device: message + CPU result
        assert "using CPU" in ec
        assert open(a, "rb").read() == open(c, "rb").read()


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [(), ("--temperature", "0", "--top-k", "0", "--subtalker-temperature", "0", "--subtalker-top-k", "0")], ids=["sampled", "greedy"])
def test_cli_on_cuda_backend(model_dir, tmp_path, extra):
    a, b = str(tmp_path / "cpu.wav"), str(tmp_path / "cuda.wav")
    ecpu = run_cli(CLI_SWITCH, model_dir, a, backend="cpu", extra=extra)
    ecuda = run_cli(CLI_SWITCH, model_dir, b, backend="cuda", extra=extra)
    assert "Backend: cuda" in ecuda and "Backend: cuda" not in ecpu
    assert trace(ecuda) == trace(ecpu)                            # codebook-0 sequence, bit-exact
    for pat in (r"Generated 9 codec tokens", r"Stop: max_tokens at step", r"Codec decode: \d+ samples", r"Total: [0-9.]+ ms"):
        assert re.search(pat, ecuda), pat                          # the lines scripts/benchmark_*.py parse
    wa, wb = read_wav(a), read_wav(b)
    assert wa.shape == wb.shape and wa.size == 9 * 1920
    assert np.abs(wa.astype(np.int32) - wb.astype(np.int32)).max() <= 1
