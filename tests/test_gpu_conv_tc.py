"""The vocoder's convolutions on tcgen05 (csrc/qtc_conv_tc.cu) against the oracle's kernel_causal_conv1d /
kernel_transposed_conv1d / kernel_snake_beta restatements, at the vocoder's own shapes: every channel count of the decoder
(1536 .. 96), k = 7 with dilation 1 / 3 / 9, k = 1 with residual, transposed convolutions k = 2r with stride r, ragged lengths
(not a multiple of the 128-sample tile), SnakeBeta on the way in.  Two bf16 pieces per value and three MMAs per product: the
bar is >= 90 dB per layer (measured ~100 dB); the whole decoder's waveform is held to 80 dB elsewhere."""
import ctypes as C

import numpy as np
import pytest

from harness import binding as B
from test_gpu_parity import snr_db

pytestmark = pytest.mark.gpu


def _log(tag, snr):
    import os
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "snr_conv_tc.log"), "a") as f:
        f.write(f"{tag} {snr:.2f} dB\n")


def _setup(cuda):
    cuda.qwen_tts_cuda_conv1d_tc.argtypes = [B.f32p, B.f32p, B.f32p, B.f32p] + [C.c_int] * 6 + [B.f32p, B.f32p, B.f32p]


@pytest.mark.parametrize("ci,co,k,L,dil,act,res", [
    (96, 96, 7, 1000, 9, True, False), (96, 96, 1, 777, 1, True, True), (192, 192, 7, 515, 3, True, False), (384, 384, 7, 300, 1, False, False),
    (768, 768, 7, 260, 9, True, False), (1024, 1536, 7, 130, 1, False, False), (64, 32, 3, 5, 1, False, False), (200, 64, 7, 129, 2, True, True)])
def test_conv_tc_vs_oracle(cuda, orc, ci, co, k, L, dil, act, res):
    _setup(cuda)
    rng = np.random.default_rng(ci + co + k)
    x = (rng.standard_normal((ci, L)) * 1.5).astype(np.float32)
    w = (rng.standard_normal((co, ci, k)) / np.sqrt(ci * k)).astype(np.float32)
    b = rng.standard_normal(co).astype(np.float32)
    a, ib = np.exp(0.2 * rng.standard_normal(ci)).astype(np.float32), np.exp(0.2 * rng.standard_normal(ci)).astype(np.float32)
    xin = x
    if act:
        xin = np.zeros_like(x)
        orc.orc_snake_beta(B.fptr(xin), B.fptr(x), B.fptr(a), B.fptr(ib), ci, L)
    want = np.zeros((co, L), dtype=np.float32)
    orc.orc_causal_conv1d(B.fptr(want), B.fptr(xin), B.fptr(w), B.fptr(b), ci, co, k, L, dil, 1)
    r = rng.standard_normal((L, co)).astype(np.float32) if res else None
    if res:
        want = want + r.T
    got = np.zeros((L, co), dtype=np.float32)
    assert cuda.qwen_tts_cuda_conv1d_tc(B.fptr(got), B.fptr(x), B.fptr(w), B.fptr(b), ci, co, k, L, dil, 1, B.fptr(a) if act else None,
                                        B.fptr(ib) if act else None, B.fptr(r) if res else None) == 0, cuda.qwen_tts_cuda_last_error()
    snr = snr_db(want, np.ascontiguousarray(got.T))
    _log(f"conv {ci}->{co} k={k} dil={dil} L={L}", snr)
    assert snr >= 90.0, snr


@pytest.mark.parametrize("ci,co,s,L", [(1536, 768, 8, 70), (768, 384, 5, 200), (384, 192, 4, 333), (192, 96, 3, 1000), (96, 64, 8, 77)])
def test_tconv_tc_vs_oracle(cuda, orc, ci, co, s, L):
    _setup(cuda)
    rng = np.random.default_rng(ci + s)
    x = rng.standard_normal((ci, L)).astype(np.float32)
    w = (rng.standard_normal((ci, co, 2 * s)) / np.sqrt(ci * 2)).astype(np.float32)
    b = rng.standard_normal(co).astype(np.float32)
    a, ib = np.exp(0.2 * rng.standard_normal(ci)).astype(np.float32), np.exp(0.2 * rng.standard_normal(ci)).astype(np.float32)
    xin = np.zeros_like(x)
    orc.orc_snake_beta(B.fptr(xin), B.fptr(x), B.fptr(a), B.fptr(ib), ci, L)
    want = np.zeros((co, L * s + 2 * s), dtype=np.float32)
    nl = C.c_int(0)
    orc.orc_transposed_conv1d(B.fptr(want), B.fptr(xin), B.fptr(w), B.fptr(b), ci, co, 2 * s, s, L, C.byref(nl))
    assert nl.value == L * s
    want = want.reshape(-1)[: co * nl.value].reshape(co, nl.value)
    got = np.zeros((L * s, co), dtype=np.float32)
    assert cuda.qwen_tts_cuda_conv1d_tc(B.fptr(got), B.fptr(x), B.fptr(w), B.fptr(b), ci, co, 2 * s, L, 1, s, B.fptr(a), B.fptr(ib), None) == 0, \
        cuda.qwen_tts_cuda_last_error()
    snr = snr_db(want, np.ascontiguousarray(got.T))
    _log(f"tconv {ci}->{co} s={s} L={L}", snr)
    assert snr >= 90.0, snr
