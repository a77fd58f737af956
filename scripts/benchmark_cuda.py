#!/usr/bin/env python3
"""`cuda` (and `cpu`) labels for the reference's C-side benchmark.

The reference's scripts/benchmark_all.py runs N binaries ("labels") over the same prompt with per-label environment and
parses the CLI's stderr protocol (scripts/benchmark_all.py:97-142, 206-250).  This driver adds the B200 backend as one more
label without editing that script: it runs oracle/_ref/qwen-tts-cuda (the reference's unmodified CLI + the backend switch,
qwen3-tts-c_b200/adapter/) once per label with `--backend <label>` and the reference's own flag set / persistent mode
(`--benchmark-runs/--benchmark-warmup`, c/main.c:245-272), and parses stderr

  * with the reference's own functions (parse_c_stderr, parse_persistent_runs, RunResult) when its scripts directory is
    present (--reference-scripts, default /root/reference/scripts), and
  * with the mirror of those three regular expressions below when it is not (the GPU box has no /root/reference);
    tests/test_benchmark_cuda.py holds the mirror to the reference's functions on real CLI output.

Output: the reference's per-run line format and a JSON report {label: {runs: [...], rtf_mean, ms_per_token_mean}}.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import re
import subprocess
import sys
import time
import wave
from dataclasses import asdict, dataclass
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CLI = ROOT / "oracle" / "_ref" / "qwen-tts-cuda"
FIXTURE = "151644,77091,198,2354,2244,2355,4041,2244,11752,13,151645,198,151644,77091,198"   # test/tokens_great_power.txt


@dataclass
class RunResult:                      # mirror of scripts/benchmark_all.py:32-57 (fields only)
    elapsed_ms: float
    audio_sec: float
    codec_tokens: int | None = None
    stop_reason: str | None = None
    internal_total_ms: float | None = None


_PERSISTENT_RE = re.compile(
    r"\[persistent\]\s+run\s+(\d+)/(\d+):\s+elapsed=([0-9.]+)\s+ms,\s+audio=([0-9.]+)s,\s+"
    r"talker=([0-9.]+)\s+ms,\s+codec=([0-9.]+)\s+ms,\s+total=([0-9.]+)\s+ms,\s+tokens=(\d+)")


def mirror_parse_c_stderr(stderr: str):
    """scripts/benchmark_all.py:97-115"""
    m = re.search(r"Total:\s+([0-9.]+)\s*ms", stderr)
    internal_ms = float(m.group(1)) if m else None
    m = re.search(r"Generated\s+(\d+)\s+codec tokens", stderr)
    codec_tokens = int(m.group(1)) if m else None
    stop = "eos" if (re.search(r"EOS at step", stderr) or codec_tokens is not None) else None
    return internal_ms, codec_tokens, stop


def mirror_parse_persistent_runs(stderr: str):
    """scripts/benchmark_all.py:124-142"""
    stop = None
    if "Stop: eos" in stderr or "EOS at step" in stderr:
        stop = "eos"
    elif "Stop: max_tokens" in stderr:
        stop = "max_tokens"
    return [RunResult(elapsed_ms=float(m.group(3)), audio_sec=float(m.group(4)), codec_tokens=int(m.group(8)), stop_reason=stop,
                      internal_total_ms=float(m.group(7))) for m in _PERSISTENT_RE.finditer(stderr)]


def load_reference_parsers(scripts_dir: str):
    """the reference's own parse functions, imported from where they lie (nothing is copied); None if absent"""
    p = Path(scripts_dir) / "benchmark_all.py"
    if not p.exists():
        return None
    spec = importlib.util.spec_from_file_location("ref_benchmark_all", str(p))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_benchmark_all"] = mod
    spec.loader.exec_module(mod)
    return mod


def build_cmd(a, label: str, wav: Path, verbose: bool, runs=None, warmup=None):
    """scripts/benchmark_all.py:145-177 + the backend flag"""
    cmd = [str(a.binary), "--backend", label, "-d", a.model_dir, "-t", a.token_ids, "-o", str(wav), "-s", a.speaker, "-l", a.language,
           "--temperature", str(a.temperature), "--top-k", str(a.top_k), "--top-p", str(a.top_p),
           "--repetition-penalty", str(a.repetition_penalty), "--subtalker-temperature", str(a.subtalker_temperature),
           "--subtalker-top-k", str(a.subtalker_top_k), "--subtalker-top-p", str(a.subtalker_top_p),
           "--max-tokens", str(a.max_new_tokens), "--seed", str(a.seed)]
    if a.fixed_codec_tokens > 0:
        cmd += ["--fixed-codec-tokens", str(a.fixed_codec_tokens)]
    if verbose:
        cmd.append("-v")
    if runs is not None:
        cmd += ["--benchmark-runs", str(runs), "--benchmark-warmup", str(warmup or 0)]
    return cmd


def wav_duration(path: Path) -> float:
    with wave.open(str(path), "rb") as wf:
        return float(wf.getnframes()) / float(wf.getframerate())


def bench_label(a, label: str, out_dir: Path, ref):
    parse_one = ref.parse_c_stderr if ref else mirror_parse_c_stderr
    parse_many = ref.parse_persistent_runs if ref else mirror_parse_persistent_runs
    make = ref.RunResult if ref else RunResult
    env = dict(os.environ)
    env.pop("QWEN_TTS_BACKEND", None)           # the flag decides
    wav = out_dir / f"{label}_output.wav"
    results, stderr_all = [], ""
    if a.persistent:
        proc = subprocess.run(build_cmd(a, label, wav, False, a.runs, a.warmup), check=True, text=True, capture_output=True, env=env)
        stderr_all = proc.stderr
        results = parse_many(proc.stderr)
    if not results:
        for i in range(a.warmup + a.runs):
            t0 = time.perf_counter()
            proc = subprocess.run(build_cmd(a, label, wav, True), check=True, text=True, capture_output=True, env=env)
            elapsed = (time.perf_counter() - t0) * 1e3
            stderr_all = proc.stderr
            internal, tokens, stop = parse_one(proc.stderr)
            if i >= a.warmup:
                results.append(make(elapsed_ms=elapsed, audio_sec=wav_duration(wav), codec_tokens=tokens, stop_reason=stop, internal_total_ms=internal))
    for i, rr in enumerate(results):
        rtf = rr.audio_sec / (rr.elapsed_ms / 1e3) if rr.elapsed_ms > 0 else 0.0
        tok = f", {rr.codec_tokens} tokens, {rr.elapsed_ms / rr.codec_tokens:.1f} ms/tok" if rr.codec_tokens else ""
        print(f"  [{label}] run {i + 1}/{len(results)}: {rr.elapsed_ms:.0f} ms, {rr.audio_sec:.2f}s audio, RTF={rtf:.2f}x{tok}")
    return results, stderr_all


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--model-dir", required=True)
    ap.add_argument("--binary", default=str(CLI))
    ap.add_argument("--labels", default="cuda", help="comma-separated: cuda, cpu")
    ap.add_argument("--token-ids", default=FIXTURE)
    ap.add_argument("--speaker", default="aiden")
    ap.add_argument("--language", default="english")
    ap.add_argument("--runs", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--persistent", action="store_true", help="one process per label, model loaded once (c/main.c:245-272)")
    ap.add_argument("--temperature", type=float, default=0.9)
    ap.add_argument("--top-k", type=int, default=50)
    ap.add_argument("--top-p", type=float, default=1.0)
    ap.add_argument("--repetition-penalty", type=float, default=1.05)
    ap.add_argument("--subtalker-temperature", type=float, default=0.9)
    ap.add_argument("--subtalker-top-k", type=int, default=50)
    ap.add_argument("--subtalker-top-p", type=float, default=1.0)
    ap.add_argument("--max-new-tokens", type=int, default=4096)
    ap.add_argument("--fixed-codec-tokens", type=int, default=0)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--out-dir", default="")
    ap.add_argument("--reference-scripts", default="/root/reference/scripts")
    ap.add_argument("--json", default="")
    a = ap.parse_args(argv)
    if not Path(a.binary).exists():
        raise SystemExit(f"{a.binary} is missing: python -c 'import __graft_entry__ as g; g.build()' builds it where /root/reference is present")
    ref = load_reference_parsers(a.reference_scripts)
    out_dir = Path(a.out_dir) if a.out_dir else Path(a.model_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    report = {"parser": "reference scripts/benchmark_all.py" if ref else "mirror (reference scripts not present)", "labels": {}}
    for label in [x for x in a.labels.split(",") if x]:
        runs, _ = bench_label(a, label, out_dir, ref)
        rtfs = [r.audio_sec / (r.elapsed_ms / 1e3) for r in runs if r.elapsed_ms > 0]
        mpt = [r.elapsed_ms / r.codec_tokens for r in runs if r.codec_tokens]
        report["labels"][label] = {"runs": [asdict(r) for r in runs], "rtf_mean": sum(rtfs) / len(rtfs) if rtfs else None,
                                   "ms_per_token_mean": sum(mpt) / len(mpt) if mpt else None}
    text = json.dumps(report, indent=2)
    if a.json:
        Path(a.json).write_text(text)
    print(text)
    return report


if __name__ == "__main__":
    main()
