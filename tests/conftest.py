"""pytest configuration: registers the `gpu` marker and shares seeded models.

`-m "not gpu"` : oracle vs golden vectors / vs the compiled reference, host logic, C-ABI symbol check.
`-m gpu`       : parity of the CUDA library (through its C ABI) against the oracle.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def orc():
    from oracle import orc_binding as O
    if not os.path.exists(O.ORACLE_LIB):
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    return O.load_oracle()


_MODELS = {}


def get_model(orc, preset, seed=0, with_lm=True, with_codec=True):
    from harness import binding as B, modelgen as mg
    key = (preset, seed, with_lm, with_codec)
    if key not in _MODELS:
        spec = mg.PRESETS[preset]
        main, cod = mg.build_tensors(spec, seed=seed, with_lm=with_lm, with_codec=with_codec)
        _MODELS[key] = B.HostModel(spec, main, cod)
    return _MODELS[key]


@pytest.fixture(scope="session")
def tiny(orc):
    return get_model(orc, "tiny")


@pytest.fixture(scope="session")
def cuda():
    """The product library.  No fallback: a missing .so or a missing device is an error under -m gpu."""
    from harness import binding as B
    lib = B.load_cuda()
    if lib.qwen_tts_cuda_init(0) != 0:
        raise RuntimeError("qwen_tts_cuda_init failed: " + (lib.qwen_tts_cuda_last_error() or b"").decode())
    return lib


FIXTURE_TOKENS = [151644, 77091, 198, 2354, 2244, 2355, 4041, 2244, 11752, 13, 151645, 198, 151644, 77091, 198]
"""test/tokens_great_power.txt of the reference (SURVEY.md section 4): the only fixture it ships"""
