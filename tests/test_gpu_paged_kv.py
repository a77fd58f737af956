"""The talker's paged KV cache (north_star: "GQA attention over an HBM-resident paged KV cache"): one pool per layer allocated at
upload, pages of 64 positions handed out by a free list, per-utterance page tables read by every attention kernel.

  * steady state: a second utterance allocates nothing (qtc_perf_t.device_mallocs) and re-uses its pages;
  * batch rows take their pages interleaved (row b: pages b, B + b, ...), so no row's cache is contiguous - codes still equal
    the oracle's solo run of each row, across a page boundary;
  * a batch-1 utterance after a batch gets its pages back from the idle rows; an over-long request or a batch wider than the
    context was uploaded for fails with a message instead of allocating."""
import ctypes as C

import numpy as np
import pytest

from conftest import FIXTURE_TOKENS, get_model
from harness import binding as B, modelgen as mg
from test_gpu_parity import CSV, TOK

pytestmark = pytest.mark.gpu


def _perf(cuda, ctx):
    p = B.Perf()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(p))
    return p


def _gen(cuda, ctx, params, n, csv=CSV):
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(params))
    ns, nf = C.c_int(0), C.c_int(0)
    codes = np.zeros((n, 16), dtype=np.int32)
    ptr = cuda.qwen_tts_cuda_generate(ctx, csv, b"aiden", b"english", C.byref(ns), B.iptr(codes), n, C.byref(nf))
    assert ptr, cuda.qwen_tts_cuda_last_error()
    cuda.qwen_tts_cuda_free_host(ptr)
    return codes[: nf.value]


def test_steady_state_allocates_nothing(cuda, tiny):
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 1)
    assert ctx
    p0 = _perf(cuda, ctx)
    assert p0.device_mallocs == 0 and p0.kv_pool_pages >= 66 and p0.kv_pages_in_use == 0
    a = _gen(cuda, ctx, B.default_params(fixed_codec_tokens=130), 130)          # prompt 10 + 130 frames: 3 pages
    p1 = _perf(cuda, ctx)
    assert p1.kv_pages_in_use == 3 and p1.kv_page_allocs == 3
    b = _gen(cuda, ctx, B.default_params(fixed_codec_tokens=130), 130)
    p2 = _perf(cuda, ctx)
    assert np.array_equal(a, b)
    assert p2.device_mallocs == p1.device_mallocs and p2.kv_page_allocs == 3, "the second utterance must not allocate"
    c = _gen(cuda, ctx, B.default_params(fixed_codec_tokens=200), 200)          # grows by two pages; workspaces grow once
    p3 = _perf(cuda, ctx)
    assert np.array_equal(c[:130], a) and p3.kv_pages_in_use == 4 and p3.kv_page_allocs == 4
    cuda.qwen_tts_cuda_release(ctx)


def test_batch_rows_with_interleaved_pages_equal_oracle(cuda, orc, tiny):
    cuda.qwen_tts_cuda_generate_batch.argtypes = [C.c_void_p, C.c_int, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.f32p, C.c_int]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 4)
    assert ctx
    nb, frames, G = 4, 70, 16                             # 10-13 prompt positions + 70 frames: every row crosses into a second page
    base = list(FIXTURE_TOKENS)
    prompts = [base, base[:4] + [2500, 31, 77] + base[4:], base[:9] + base[10:], base[:4] + [900] + base[4:]]
    toks = np.array(sum(prompts, []), dtype=np.int32)
    offs = np.array([0] + list(np.cumsum([len(q) for q in prompts])), dtype=np.int32)
    spk = np.full(nb, mg.SPEAKERS["aiden"], dtype=np.int32)
    lang = np.full(nb, mg.LANGUAGES["english"], dtype=np.int32)
    p = B.default_params(fixed_codec_tokens=frames, sample_seed=9)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
    codes = np.zeros((nb, frames, G), dtype=np.int32)
    nfr = np.zeros(nb, dtype=np.int32)
    assert cuda.qwen_tts_cuda_generate_batch(ctx, nb, B.iptr(toks), B.iptr(offs), B.iptr(spk), B.iptr(lang), B.iptr(codes), B.iptr(nfr), None, frames) == 0, \
        cuda.qwen_tts_cuda_last_error()
    pf = _perf(cuda, ctx)
    assert pf.kv_pages_in_use == 2 * nb                   # two pages per row, handed out page by page across the rows
    for b in range(nb):
        s = orc.orc_create(tiny.desc)
        pb = B.default_params(fixed_codec_tokens=frames, sample_seed=9 + b)
        orc.orc_set_params(s, C.byref(pb))
        want = np.zeros((frames, G), dtype=np.int32)
        ns, nf = C.c_int(0), C.c_int(0)
        tk = np.array(prompts[b], dtype=np.int32)
        ptr = orc.orc_generate(s, B.iptr(tk), len(tk), 3010, 2050, C.byref(ns), B.iptr(want), frames, C.byref(nf))
        orc.orc_free(ptr)
        orc.orc_destroy(s)
        assert np.array_equal(codes[b], want), f"row {b}: first mismatch at frame {np.argwhere(codes[b] != want)[0]}"
    # a single utterance afterwards: slot 0 keeps its pages, grows from the free list, same codes as row 0
    one = _gen(cuda, ctx, B.default_params(fixed_codec_tokens=frames, sample_seed=9), frames)
    assert np.array_equal(one, codes[0])
    cuda.qwen_tts_cuda_release(ctx)


def test_pool_limits_are_errors_not_allocations(cuda, tiny):
    cuda.qwen_tts_cuda_generate_batch.argtypes = [C.c_void_p, C.c_int, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.f32p, C.c_int]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 1)
    assert ctx
    H = tiny.spec["talker"]["hidden"]
    pool_positions = _perf(cuda, ctx).kv_pool_pages * 64
    emb = np.zeros((pool_positions + 8, H), dtype=np.float32)
    cuda.qwen_tts_cuda_reset(ctx)
    assert cuda.qwen_tts_cuda_talker_prefill(ctx, B.fptr(emb), pool_positions + 8) == -1
    assert b"pool" in cuda.qwen_tts_cuda_last_error() or b"positions" in cuda.qwen_tts_cuda_last_error()
    # a batch of two on a context uploaded for one
    toks = np.array(list(FIXTURE_TOKENS) * 2, dtype=np.int32)
    offs = np.array([0, len(FIXTURE_TOKENS), 2 * len(FIXTURE_TOKENS)], dtype=np.int32)
    codes, nfr = np.zeros((2, 4, 16), dtype=np.int32), np.zeros(2, dtype=np.int32)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.default_params(fixed_codec_tokens=4)))
    assert cuda.qwen_tts_cuda_generate_batch(ctx, 2, B.iptr(toks), B.iptr(offs), None, None, B.iptr(codes), B.iptr(nfr), None, 4) == -1
    assert b"max_batch" in cuda.qwen_tts_cuda_last_error()
    # the context still works
    assert len(_gen(cuda, ctx, B.default_params(fixed_codec_tokens=5), 5)) == 5
    cuda.qwen_tts_cuda_release(ctx)
