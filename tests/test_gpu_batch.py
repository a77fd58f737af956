"""Batched (lock-step) decode path: the tcgen05 GEMM against numpy, then whole batched steps against the oracle."""
import ctypes as C

import numpy as np
import pytest

from harness import binding as B, modelgen as mg

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(180)]


def _bf16_round(a):
    u = a.astype(np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) >> 16
    return (u.astype(np.uint32) << 16).view(np.float32), u.astype(np.uint16)


@pytest.mark.parametrize("M,N,K", [(64, 4096, 2048), (5, 300, 128), (64, 2048, 6144), (17, 3072, 1024), (1, 128, 64)])
def test_gemm_tc_matches_numpy(cuda, M, N, K):
    """C[M,N] = bf16(A)[M,K] . W[N,K]^T on the tcgen05 path (TMA + UMMA + TMEM epilogue, split-K reductions in L2)"""
    rng = np.random.default_rng(M * 1000 + N + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    Wf, Wb = _bf16_round(rng.standard_normal((N, K)).astype(np.float32) * 0.05)
    Ar, _ = _bf16_round(A)
    want = Ar.astype(np.float64) @ Wf.astype(np.float64).T
    got = np.zeros((M, N), dtype=np.float32)
    cuda.qwen_tts_cuda_gemm_bf16_tc.argtypes = [B.f32p, B.f32p, C.POINTER(C.c_uint16), C.c_int, C.c_int, C.c_int]
    rc = cuda.qwen_tts_cuda_gemm_bf16_tc(B.fptr(got), B.fptr(A), Wb.ctypes.data_as(C.POINTER(C.c_uint16)), M, N, K)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    err = np.abs(got - want).max() / np.abs(want).max()
    assert err <= 2e-5, err          # products are exact in fp32; only the accumulation order differs
