"""`--backend {cpu,cuda}` on the reference's CLI and a `cuda` label for its benchmark scripts.

CPU: the flag is removed before the reference's parser sees argv (c/main.c:126-173 rejects unknown options) and selects the
backend; scripts/benchmark_cuda.py parses the CLI's stderr with the reference's own functions (scripts/benchmark_all.py:97-142)
and its mirror regexes give the same answers on real CLI output.  GPU: the `cuda` label runs on the B200 library and parses."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import FIXTURE_TOKENS, ROOT
from harness import modelgen as mg

sys.path.insert(0, os.path.join(ROOT, "scripts"))
import benchmark_cuda as bc  # noqa: E402

CLI = os.path.join(ROOT, "oracle", "_ref", "qwen-tts-cuda")
pytestmark = pytest.mark.skipif(not os.path.exists(CLI), reason="oracle/_ref/qwen-tts-cuda not built (needs /root/reference at build time)")
IDS = ",".join(str(t) for t in FIXTURE_TOKENS)


@pytest.fixture(scope="module")
def model_dir(tmp_path_factory):
    spec = mg.PRESETS["tiny"]
    main, cod = mg.build_tensors(spec, seed=3)
    d = str(tmp_path_factory.mktemp("bench_model"))
    mg.write_model_dir(d, spec, main, cod)
    return d


def _run(args, env_backend=None):
    env = dict(os.environ)
    env.pop("QWEN_TTS_BACKEND", None)
    if env_backend:
        env["QWEN_TTS_BACKEND"] = env_backend
    return subprocess.run([CLI, *args], env=env, capture_output=True, text=True, timeout=600)


def test_backend_flag_is_stripped_and_validated(model_dir, tmp_path):
    out = str(tmp_path / "o.wav")
    base = ["-d", model_dir, "-t", IDS, "-s", "aiden", "-l", "english", "-o", out, "-v", "--fixed-codec-tokens", "3"]
    r = _run(["--backend", "cpu", *base])
    assert r.returncode == 0 and "Unknown option" not in r.stderr and "Backend: cuda" not in r.stderr
    r = _run([*base, "--backend", "cpu"], env_backend="cuda")          # the flag wins over the environment, wherever it stands
    assert r.returncode == 0 and "Backend: cuda" not in r.stderr and "using CPU" not in r.stderr
    r = _run(["--backend", "tpu", *base])
    assert r.returncode == 1 and "Unknown backend" in r.stderr
    r = _run([*base, "--backend"])
    assert r.returncode == 1


@pytest.mark.skipif(not os.path.isdir("/root/reference/scripts"), reason="reference scripts not present")
def test_mirror_regexes_equal_reference_parsers(model_dir, tmp_path):
    ref = bc.load_reference_parsers("/root/reference/scripts")
    assert ref is not None
    base = ["--backend", "cpu", "-d", model_dir, "-t", IDS, "-s", "aiden", "-l", "english", "-o", str(tmp_path / "o.wav"), "--fixed-codec-tokens", "4"]
    one = _run([*base, "-v"]).stderr
    many = _run([*base, "--benchmark-runs", "2", "--benchmark-warmup", "1"]).stderr
    assert bc.mirror_parse_c_stderr(one) == ref.parse_c_stderr(one)
    assert bc.mirror_parse_c_stderr(one)[1] == 4
    got, want = bc.mirror_parse_persistent_runs(many), ref.parse_persistent_runs(many)
    assert len(got) == len(want) == 2
    for g, w in zip(got, want):
        assert (g.elapsed_ms, g.audio_sec, g.codec_tokens, g.stop_reason, g.internal_total_ms) == \
               (w.elapsed_ms, w.audio_sec, w.codec_tokens, w.stop_reason, w.internal_total_ms)


def test_benchmark_driver_cpu_label(model_dir, tmp_path):
    rep = bc.main(["--model-dir", model_dir, "--labels", "cpu", "--runs", "2", "--warmup", "0", "--persistent", "--fixed-codec-tokens", "4",
                   "--out-dir", str(tmp_path)])
    runs = rep["labels"]["cpu"]["runs"]
    assert len(runs) == 2 and all(r["codec_tokens"] == 4 and abs(r["audio_sec"] - 4 * 0.08) < 1e-6 for r in runs)


@pytest.mark.gpu
def test_benchmark_driver_cuda_label(model_dir, tmp_path):
    rep = bc.main(["--model-dir", model_dir, "--labels", "cuda,cpu", "--runs", "2", "--warmup", "1", "--persistent", "--fixed-codec-tokens", "6",
                   "--out-dir", str(tmp_path)])
    for label in ("cuda", "cpu"):
        runs = rep["labels"][label]["runs"]
        assert len(runs) == 2 and all(r["codec_tokens"] == 6 and abs(r["audio_sec"] - 6 * 0.08) < 1e-6 for r in runs), (label, runs)   # (no -v in persistent mode: no Stop line)
    assert rep["labels"]["cuda"]["rtf_mean"] > 0
    # single-shot mode too (the `Total:` / `Generated N codec tokens` lines)
    rep = bc.main(["--model-dir", model_dir, "--labels", "cuda", "--runs", "1", "--warmup", "0", "--fixed-codec-tokens", "6", "--out-dir", str(tmp_path)])
    r = rep["labels"]["cuda"]["runs"][0]
    assert r["codec_tokens"] == 6 and r["internal_total_ms"] > 0 and abs(r["audio_sec"] - 6 * 0.08) < 1e-6
