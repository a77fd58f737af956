"""qwen_tts_cuda_sample_top_k_noalloc: the allocation-free variant of the sampling entry point (scratch in a static
device array, device pointers in and out) against qwen_tts_cuda_sample_top_k through the C ABI - token and RNG bits
must be identical (integer work: bit-exact), on two calls in a row that share the static scratch.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from harness import binding as B

pytestmark = pytest.mark.gpu

MAX_VOCAB = 16384       # QTC_NOALLOC_MAX_VOCAB in include/qwen_tts_cuda.h

# (vocab, top_k, top_p, temperature): radix-select path (no scratch), top-p slow path (p[] + ord[] in the scratch),
# literal insertion list (k > 1024: val[] / idx[] in the scratch), full-vocabulary slow path, the maximum size
CASES = [(3072, 50, 1.0, 0.9), (3072, 50, 0.9, 1.0), (3072, 1500, 1.0, 0.9), (2048, 0, 0.5, 1.0), (3072, 50, 0.8, 0.9),
         (MAX_VOCAB, 40, 0.95, 1.0)]


def _draw_pair(cuda, torch, lg, vocab, k, top_p, temp, seed):
    """one draw through each entry point from the same RNG state -> ((token, rng bits) existing, (token, rng bits) no-alloc)"""
    st = np.array([seed], dtype=np.uint32).view(np.float32)
    want_tok = cuda.qwen_tts_cuda_sample_top_k(B.fptr(lg), vocab, k, C.c_float(top_p), C.c_float(temp), B.fptr(st))
    assert want_tok >= 0, cuda.qwen_tts_cuda_last_error()
    want_rng = int(st.view(np.uint32)[0])

    d_lg = torch.from_numpy(lg).cuda()
    d_rng = torch.from_numpy(np.array([seed], dtype=np.uint32).view(np.int32)).cuda()
    d_tok = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = cuda.qwen_tts_cuda_sample_top_k_noalloc(d_lg.data_ptr(), vocab, k, C.c_float(top_p), C.c_float(temp),
                                                 d_rng.data_ptr(), d_tok.data_ptr(), None)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    torch.cuda.synchronize()
    got_tok = int(d_tok.cpu()[0])
    got_rng = int(d_rng.cpu().numpy().view(np.uint32)[0])
    return (want_tok, want_rng), (got_tok, got_rng)


def test_noalloc_sampler_equals_existing_entry_point(cuda):
    import torch
    assert torch.cuda.is_available()
    rng = np.random.default_rng(11)
    for rep in range(2):                      # two passes over every case: the array keeps the previous call's contents
        for (vocab, k, top_p, temp) in CASES:
            lg = (rng.standard_normal(vocab) * 3).astype(np.float32)
            lg[rng.integers(0, vocab, 8)] = lg.max()          # ties at the top: order must still match
            for seed in (42 + rep, 977):
                want, got = _draw_pair(cuda, torch, lg, vocab, k, top_p, temp, seed)
                assert got == want, (rep, vocab, k, top_p, temp, seed, got, want)


def test_noalloc_sampler_two_calls_in_a_row_share_the_scratch(cuda):
    """back to back on one stream without a host sync in between: a large top-p draw leaves the array full, a smaller
    one follows; both equal the existing entry point, and the RNG state chains across the calls"""
    import torch
    rng = np.random.default_rng(5)
    big = (rng.standard_normal(8192) * 2).astype(np.float32)
    small = (rng.standard_normal(1024) * 2).astype(np.float32)
    st = np.array([1234], dtype=np.uint32).view(np.float32)
    want = [cuda.qwen_tts_cuda_sample_top_k(B.fptr(big), 8192, 100, C.c_float(0.9), C.c_float(1.0), B.fptr(st)),
            cuda.qwen_tts_cuda_sample_top_k(B.fptr(small), 1024, 0, C.c_float(0.7), C.c_float(0.8), B.fptr(st))]
    want_rng = int(st.view(np.uint32)[0])

    d_big, d_small = torch.from_numpy(big).cuda(), torch.from_numpy(small).cuda()
    d_rng = torch.from_numpy(np.array([1234], dtype=np.uint32).view(np.int32)).cuda()
    d_tok = torch.full((2,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    assert cuda.qwen_tts_cuda_sample_top_k_noalloc(d_big.data_ptr(), 8192, 100, C.c_float(0.9), C.c_float(1.0),
                                                   d_rng.data_ptr(), d_tok.data_ptr(), None) == 0
    assert cuda.qwen_tts_cuda_sample_top_k_noalloc(d_small.data_ptr(), 1024, 0, C.c_float(0.7), C.c_float(0.8),
                                                   d_rng.data_ptr(), d_tok.data_ptr() + 4, None) == 0
    torch.cuda.synchronize()
    assert d_tok.cpu().tolist() == want
    assert int(d_rng.cpu().numpy().view(np.uint32)[0]) == want_rng


def test_noalloc_sampler_rejects_sizes_beyond_the_static_scratch(cuda):
    import torch
    d = torch.zeros(8, dtype=torch.float32, device="cuda")
    t = torch.zeros(1, dtype=torch.int32, device="cuda")
    for vocab in (MAX_VOCAB + 1, 0, -3):
        rc = cuda.qwen_tts_cuda_sample_top_k_noalloc(d.data_ptr(), vocab, 50, C.c_float(1.0), C.c_float(1.0), d.data_ptr(), t.data_ptr(), None)
        assert rc == -1
        assert b"noalloc" in cuda.qwen_tts_cuda_last_error() and b"16384" in cuda.qwen_tts_cuda_last_error()
    assert cuda.qwen_tts_cuda_sample_top_k_noalloc(None, 128, 50, C.c_float(1.0), C.c_float(1.0), d.data_ptr(), t.data_ptr(), None) == -1
