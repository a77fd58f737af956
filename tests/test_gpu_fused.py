"""Mode 1 (one persistent cooperative kernel per decoder-stack step, weights of the next phase landing in shared memory
across the grid barriers) against the same golden vectors and oracle as the per-op path, with the landing zone on and off
(QTC_FUSED_FLAGS & 128).  Fast arithmetic: logits within 1e-2 relative (north_star), codes equal."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import FIXTURE_TOKENS, ROOT, get_model
from harness import binding as B
from test_gpu_parity import GEN_CASES, GOLD, TOK, _generate, _stage_run, snr_db

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=[1], ids=["fused"])
def fused_ctx(request, cuda, tiny):
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    assert cuda.qwen_tts_cuda_set_mode(ctx, request.param) == 0
    yield ctx
    cuda.qwen_tts_cuda_release(ctx)


def test_fused_stage_api(cuda, tiny, fused_ctx):
    hid, logits, hids, subs = _stage_run(cuda, fused_ctx, tiny.spec, 0)
    want = GOLD["stage_logits"]
    rel = np.abs(logits - want).max() / np.abs(want).max()
    assert rel <= 2e-4, rel
    assert np.abs(hid - GOLD["stage_prefill_hidden"]).max() <= 2e-4 * np.abs(hid).max()
    assert np.abs(hids - GOLD["stage_hiddens"]).max() <= 2e-4 * np.abs(hids).max()
    assert np.array_equal(subs, GOLD["stage_sub_codes"])


@pytest.mark.parametrize("name,params,spk,lang", GEN_CASES)
def test_fused_generate(cuda, fused_ctx, name, params, spk, lang):
    want_codes, want_pcm = GOLD[f"gen_{name}_codes"], GOLD[f"gen_{name}_pcm"]
    cuda.qwen_tts_cuda_set_arith(fused_ctx, 0)
    codes, pcm = _generate(cuda, fused_ctx, params, spk, lang, len(want_codes))
    assert np.array_equal(codes, want_codes)
    assert snr_db(want_pcm, pcm) >= 80.0


@pytest.mark.parametrize("mode,n_pre,n_steps", [(1, 60, 150), (2, 60, 150), (2, 490, 40)], ids=["fused", "cluster", "cluster-past-512"])
def test_fused_long_context_splits(cuda, orc, mode, n_pre, n_steps):
    """KV length past 64 / 128 / ... positions exercises the split-KV attention and its combine step:
    teacher-forced talker steps on the `small` preset (16 Q / 8 KV heads of 128) against the oracle.  The cluster kernel stages a
    CTA's eighth of the context in shared memory up to 8 x 64 = 512 positions and reads longer contexts through L2: the third
    case crosses that boundary at step 22."""
    hm = get_model(orc, "small")
    H, V = hm.spec["talker"]["hidden"], hm.spec["talker"]["vocab"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx and cuda.qwen_tts_cuda_set_mode(ctx, mode) == 0
    s = orc.orc_create(hm.desc)
    rng = np.random.default_rng(0)
    emb = (rng.standard_normal((n_pre + n_steps, H)) * 0.5).astype(np.float32)
    orc.orc_reset(s)
    orc.orc_talker_prefill(s, B.fptr(emb), n_pre)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.greedy_params()))
    cuda.qwen_tts_cuda_reset(ctx)
    assert cuda.qwen_tts_cuda_talker_prefill(ctx, B.fptr(emb), n_pre) == 0
    worst = 0.0
    for i in range(n_steps):
        want, got = np.zeros(V, dtype=np.float32), np.zeros(V, dtype=np.float32)
        orc.orc_talker_forward(s, B.fptr(emb[n_pre + i]), B.fptr(want))
        assert cuda.qwen_tts_cuda_talker_forward(ctx, B.fptr(emb[n_pre + i]), B.fptr(got)) == 0
        worst = max(worst, float(np.abs(got - want).max() / np.abs(want).max()))
        assert int(got.argmax()) == int(want.argmax()), i
    assert worst <= 5e-4, worst
    assert cuda.qwen_tts_cuda_get_kv_len(ctx) == n_pre + n_steps
    orc.orc_destroy(s)
    cuda.qwen_tts_cuda_release(ctx)


@pytest.mark.parametrize("n_pre", [1 + 15, 64, 65, 100], ids=["skinny16", "skinny64", "wide65", "wide100"])
def test_prefill_both_gemm_paths_vs_oracle(cuda, orc, n_pre):
    """Batched prefill (c/qwen_tts_talker.c:254-472): prompts of <= 64 rows go through the skinny tcgen05 GEMM (weights as the MMA's
    128-row operand, K-split partial sums added in a fixed order), longer ones through the convolution kernel.  Either way the
    hidden state of the last prompt token and the logits of the decode step that follows (which reads every K/V row the prefill
    wrote) are within 2e-4 of the oracle's, and two runs agree bit for bit."""
    hm = get_model(orc, "small")
    H, V = hm.spec["talker"]["hidden"], hm.spec["talker"]["vocab"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    s = orc.orc_create(hm.desc)
    emb = (np.random.default_rng(n_pre).standard_normal((n_pre + 1, H)) * 0.5).astype(np.float32)
    orc.orc_reset(s)
    orc.orc_talker_prefill(s, B.fptr(emb), n_pre)
    want = np.zeros(V, dtype=np.float32)
    want_hid = np.ctypeslib.as_array(orc.orc_hidden(s), shape=(H,)).copy()
    orc.orc_talker_forward(s, B.fptr(emb[n_pre]), B.fptr(want))
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.greedy_params()))
    runs = []
    for _ in range(2):
        cuda.qwen_tts_cuda_reset(ctx)
        assert cuda.qwen_tts_cuda_talker_prefill(ctx, B.fptr(emb), n_pre) == 0, cuda.qwen_tts_cuda_last_error()
        hid, got = np.zeros(H, dtype=np.float32), np.zeros(V, dtype=np.float32)
        assert cuda.qwen_tts_cuda_get_hidden(ctx, B.fptr(hid)) == 0
        assert cuda.qwen_tts_cuda_talker_forward(ctx, B.fptr(emb[n_pre]), B.fptr(got)) == 0
        assert np.abs(hid - want_hid).max() <= 2e-4 * np.abs(want_hid).max()
        assert np.abs(got - want).max() <= 5e-4 * np.abs(want).max()
        assert int(got.argmax()) == int(want.argmax())
        runs.append((hid, got))
    assert np.array_equal(runs[0][0].view(np.uint32), runs[1][0].view(np.uint32))
    assert np.array_equal(runs[0][1].view(np.uint32), runs[1][1].view(np.uint32))
    orc.orc_destroy(s)
    cuda.qwen_tts_cuda_release(ctx)
