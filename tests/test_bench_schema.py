"""The bench contract, checked on the committed lines of the last GPU runs (profiles/): every key the driver reads is there,
with the meaning the contract gives it.  CPU-only: bench.py itself needs a B200."""
import glob
import json
import os

import pytest

from conftest import ROOT

PROFILES = os.path.join(ROOT, "profiles")
OWN = sorted(glob.glob(os.path.join(PROFILES, "r01_bench_17b_mode1_v*.json")) + glob.glob(os.path.join(PROFILES, "r01_bench_17b_batch64_pdl1*.json")) +
             glob.glob(os.path.join(PROFILES, "r01_bench_codec_*.json")) + glob.glob(os.path.join(PROFILES, "r01_bench_06b_*.json")))
REF = sorted(glob.glob(os.path.join(PROFILES, "r01_bench_reference_arm*.json")))


def _line(path):
    txt = open(path).read().strip().splitlines()
    return json.loads(txt[-1])


@pytest.mark.parametrize("path", OWN, ids=[os.path.basename(p) for p in OWN])
def test_own_arm_line(path):
    d = _line(path)
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
              "config", "e2e", "gpu_launches", "roofline", "clocks"):
        assert k in d, k
    assert d["unit"] == "audio_s/s" and d["higher_is_better"] is True and d["scaling"] == "weak" and d["data"] == "synthetic"
    assert d["vs_baseline"] is None                       # BASELINE.md publishes no number for this metric
    assert d["warmup"] >= 1 and d["steps"] >= 1 and d["value"] > 0 and d["gpu_launches"] > 0
    assert "workload" in d["config"] and "model" not in d["config"]
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert e["value"] <= d["value"] * 1.001                # host copies and the call overhead are inside the e2e region
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-6 and 0 < r["frac"] < 1
    c = d["clocks"]
    assert c["sm_mhz"] > 0.9 * c["sm_max_mhz"] and not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


@pytest.mark.parametrize("path", REF, ids=[os.path.basename(p) for p in REF])
def test_reference_arm_line(path):
    d = _line(path)
    assert d["impl"] == "reference" and d["unit"] == "audio_s/s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_default_line_carries_cpu_baseline():
    paths = [p for p in OWN if "mode1" in p and "cpu_baseline" in _line(p)]
    assert paths, "no committed default-run line with a cpu_baseline"
    cb = _line(paths[-1])["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] > 0 and cb["sample"]
