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


@pytest.mark.parametrize("pieces", [1, 3])
@pytest.mark.parametrize("M,N,K", [(64, 4096, 2048), (5, 300, 128), (64, 2048, 6144), (17, 3072, 1024), (1, 128, 64)])
def test_gemm_tc_matches_numpy(cuda, M, N, K, pieces):
    """C[M,N] = bf16(A)[M,K] . W[N,K]^T on the tcgen05 path (TMA + UMMA + TMEM epilogue, split-K reductions in L2)"""
    rng = np.random.default_rng(M * 1000 + N + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    Wf, Wb = _bf16_round(rng.standard_normal((N, K)).astype(np.float32) * 0.05)
    Ar, _ = _bf16_round(A)
    want = (Ar if pieces == 1 else A).astype(np.float64) @ Wf.astype(np.float64).T
    got = np.zeros((M, N), dtype=np.float32)
    cuda.qwen_tts_cuda_gemm_bf16_tc.argtypes = [B.f32p, B.f32p, C.POINTER(C.c_uint16), C.c_int, C.c_int, C.c_int, C.c_int]
    rc = cuda.qwen_tts_cuda_gemm_bf16_tc(B.fptr(got), B.fptr(A), Wb.ctypes.data_as(C.POINTER(C.c_uint16)), M, N, K, pieces)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    err = np.abs(got - want).max() / np.abs(want).max()
    assert err <= 2e-5, err          # products are exact in fp32; only the accumulation order differs


# ---------------------------------------------------------------- batched stage API vs the oracle
from conftest import FIXTURE_TOKENS, get_model  # noqa: E402


def _setup_batch_api(cuda):
    cuda.qwen_tts_cuda_batch_reset.argtypes = [C.c_void_p, C.c_int, C.c_int]
    cuda.qwen_tts_cuda_batch_forward.argtypes = [C.c_void_p, C.c_int, B.f32p, B.f32p, B.f32p]
    cuda.qwen_tts_cuda_batch_subtalker.argtypes = [C.c_void_p, C.c_int, B.f32p, B.i32p, B.i32p]
    cuda.qwen_tts_cuda_set_batch_pieces.argtypes = [C.c_void_p, C.c_int]
    cuda.qwen_tts_cuda_generate_batch.argtypes = [C.c_void_p, C.c_int, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.f32p, C.c_int]
    cuda.qwen_tts_cuda_generate_queue.argtypes = [C.c_void_p, C.c_int, C.c_int, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.f32p, C.c_int]


@pytest.mark.parametrize("pieces,tol", [(3, 2e-4), (1, 1e-2)])
def test_batch_talker_steps_vs_oracle(cuda, orc, pieces, tol):
    """5 utterances with different inputs take 40 lock-step decode steps; each row's logits and hidden state against
    its own oracle context (talker logits within 1e-2 relative in bf16 - north_star; ~1e-5 with the exact split)"""
    _setup_batch_api(cuda)
    hm = get_model(orc, "small")
    H, V = hm.spec["talker"]["hidden"], hm.spec["talker"]["vocab"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 8)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    nb, steps = 5, 40
    assert cuda.qwen_tts_cuda_set_batch_pieces(ctx, pieces) == 0
    assert cuda.qwen_tts_cuda_batch_reset(ctx, nb, steps + 4) == 0, cuda.qwen_tts_cuda_last_error()
    rng = np.random.default_rng(7)
    emb = (rng.standard_normal((steps, nb, H)) * 0.5).astype(np.float32)
    states = [orc.orc_create(hm.desc) for _ in range(nb)]
    for s in states:
        orc.orc_reset(s)
    worst = worst_h = 0.0
    for t in range(steps):
        got, hid = np.zeros((nb, V), dtype=np.float32), np.zeros((nb, H), dtype=np.float32)
        assert cuda.qwen_tts_cuda_batch_forward(ctx, nb, B.fptr(np.ascontiguousarray(emb[t])), B.fptr(got), B.fptr(hid)) == 0, cuda.qwen_tts_cuda_last_error()
        for b, s in enumerate(states):
            want = np.zeros(V, dtype=np.float32)
            orc.orc_talker_forward(s, B.fptr(np.ascontiguousarray(emb[t, b])), B.fptr(want))
            wh = np.ctypeslib.as_array(orc.orc_hidden(s), shape=(H,))
            worst = max(worst, float(np.abs(got[b] - want).max() / np.abs(want).max()))
            worst_h = max(worst_h, float(np.abs(hid[b] - wh).max() / np.abs(wh).max()))
            if pieces == 3:
                assert int(got[b].argmax()) == int(want.argmax()), (t, b)
    assert worst <= tol and worst_h <= tol, (worst, worst_h)
    for s in states:
        orc.orc_destroy(s)
    cuda.qwen_tts_cuda_release(ctx)


def test_batch_subtalker_codes_equal_oracle(cuda, orc, tiny):
    """the 15-codebook loop for 4 utterances at once, default sampling, seeds 42 + b: codes equal the oracle's"""
    _setup_batch_api(cuda)
    H, G = tiny.spec["talker"]["hidden"], tiny.spec["groups"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 4)
    assert ctx
    nb = 4
    p = B.default_params()
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
    assert cuda.qwen_tts_cuda_batch_reset(ctx, nb, 32) == 0, cuda.qwen_tts_cuda_last_error()
    rng = np.random.default_rng(3)
    hidden = rng.standard_normal((nb, H)).astype(np.float32)
    first = np.array([5, 900, 2047, 17], dtype=np.int32)
    got = np.zeros((nb, G), dtype=np.int32)
    assert cuda.qwen_tts_cuda_batch_subtalker(ctx, nb, B.fptr(hidden), B.iptr(first), B.iptr(got)) == 0, cuda.qwen_tts_cuda_last_error()
    for b in range(nb):
        s = orc.orc_create(tiny.desc)
        pb = B.default_params(sample_seed=42 + b)
        orc.orc_set_params(s, C.byref(pb))
        want = np.zeros(32, dtype=np.int32)
        orc.orc_subtalker_generate(s, B.fptr(np.ascontiguousarray(hidden[b])), int(first[b]), B.iptr(want))
        orc.orc_destroy(s)
        assert np.array_equal(got[b], want[:G]), (b, got[b], want[:G])
    cuda.qwen_tts_cuda_release(ctx)


def test_generate_batch_equals_per_utterance_oracle(cuda, orc, tiny):
    """whole utterances: 3 prompts of different lengths, speakers and languages, default sampling, 6 frames; codes of
    utterance b equal the oracle's run with seed 42 + b, waveform SNR >= 80 dB"""
    _setup_batch_api(cuda)
    G = tiny.spec["groups"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 4)
    assert ctx
    frames = 6
    p = B.default_params(fixed_codec_tokens=frames)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
    base = list(FIXTURE_TOKENS)
    prompts = [base, base[:4] + [2500, 31, 77] + base[4:], base[:9] + base[10:]]
    spk = np.array([mg.SPEAKERS["aiden"], -1, mg.SPEAKERS["vivian"]], dtype=np.int32)
    lang = np.array([mg.LANGUAGES["english"], mg.LANGUAGES["chinese"], -1], dtype=np.int32)
    toks = np.array(sum(prompts, []), dtype=np.int32)
    offs = np.array([0] + list(np.cumsum([len(q) for q in prompts])), dtype=np.int32)
    nb = len(prompts)
    codes = np.zeros((nb, frames, G), dtype=np.int32)
    nfr = np.zeros(nb, dtype=np.int32)
    pcm = np.zeros((nb, frames * 1920), dtype=np.float32)
    rc = cuda.qwen_tts_cuda_generate_batch(ctx, nb, B.iptr(toks), B.iptr(offs), B.iptr(spk), B.iptr(lang), B.iptr(codes), B.iptr(nfr), B.fptr(pcm), frames)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    assert list(nfr) == [frames] * nb
    for b in range(nb):
        s = orc.orc_create(tiny.desc)
        pb = B.default_params(fixed_codec_tokens=frames, sample_seed=42 + b)
        orc.orc_set_params(s, C.byref(pb))
        want = np.zeros((frames, G), dtype=np.int32)
        ns, nf = C.c_int(0), C.c_int(0)
        tk = np.array(prompts[b], dtype=np.int32)
        ptr = orc.orc_generate(s, B.iptr(tk), len(tk), int(spk[b]), int(lang[b]), C.byref(ns), B.iptr(want), frames, C.byref(nf))
        want_pcm = B.take_floats(orc.orc_free, ptr, ns.value)
        orc.orc_destroy(s)
        assert np.array_equal(codes[b], want), (b, codes[b], want)
        err = ((pcm[b].astype(np.float64) - want_pcm) ** 2).sum()
        snr = 10 * np.log10((want_pcm.astype(np.float64) ** 2).sum() / max(err, 1e-300))
        assert snr >= 80.0, (b, snr)
    cuda.qwen_tts_cuda_release(ctx)


def test_generate_batch_mixed_eos_equals_per_utterance_oracle(cuda, orc):
    """no frame budget: the codec head's EOS row is scaled up so that the four utterances of the batch stop at different frames
    (the oracle: 11, 4, 19 and 7) - rows that are done keep taking lock-step steps as no-ops until the slowest one ends; every
    utterance's frame count and codes must equal the oracle's solo run with its seed (the all_done / stop handling)"""
    _setup_batch_api(cuda)
    spec = mg.PRESETS["tiny"]
    main, cod = mg.build_tensors(spec, seed=0)
    w = mg.bf16_to_f32(main["talker.codec_head.weight"]).copy()
    w[mg.CODEC_IDS["codec_eos_token_id"]] *= 3.0
    main["talker.codec_head.weight"] = mg.f32_to_bf16(w)
    hm = B.HostModel(spec, main, cod)
    G = spec["groups"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 4)
    assert ctx
    cap = 30
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.default_params(max_new_tokens=cap)))
    base = list(FIXTURE_TOKENS)
    prompts = [base, base[:4] + [2500, 31, 77] + base[4:], base[:9] + base[10:], base[:6] + [99] + base[6:]]
    nb = len(prompts)
    spk = np.full(nb, mg.SPEAKERS["aiden"], dtype=np.int32)
    lang = np.full(nb, mg.LANGUAGES["english"], dtype=np.int32)
    toks = np.array(sum(prompts, []), dtype=np.int32)
    offs = np.array([0] + list(np.cumsum([len(q) for q in prompts])), dtype=np.int32)
    codes = np.zeros((nb, cap, G), dtype=np.int32)
    nfr = np.zeros(nb, dtype=np.int32)
    rc = cuda.qwen_tts_cuda_generate_batch(ctx, nb, B.iptr(toks), B.iptr(offs), B.iptr(spk), B.iptr(lang), B.iptr(codes), B.iptr(nfr), None, cap)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    want_n = []
    for b in range(nb):
        s = orc.orc_create(hm.desc)
        orc.orc_set_params(s, C.byref(B.default_params(max_new_tokens=cap, sample_seed=42 + b)))
        want = np.zeros((cap, G), dtype=np.int32)
        ns, nf = C.c_int(0), C.c_int(0)
        tk = np.array(prompts[b], dtype=np.int32)
        ptr = orc.orc_generate(s, B.iptr(tk), len(tk), int(spk[b]), int(lang[b]), C.byref(ns), B.iptr(want), cap, C.byref(nf))
        if ptr:
            orc.orc_free(ptr)
        orc.orc_destroy(s)
        want_n.append(nf.value)
        assert int(nfr[b]) == nf.value, (b, list(nfr), want_n)
        assert np.array_equal(codes[b][: nf.value], want[: nf.value]), b
    assert len(set(want_n)) > 2 and min(want_n) < cap, want_n       # the utterances really stop at different frames
    cuda.qwen_tts_cuda_release(ctx)


def _queue_prompts(n):
    base = list(FIXTURE_TOKENS)
    variants = [base, base[:4] + [2500, 31, 77] + base[4:], base[:9] + base[10:], base[:6] + [99] + base[6:],
                base[:5] + [1234, 4321] + base[5:], base[:8] + base[9:], base[:7] + [555, 556, 557, 558] + base[7:],
                base[:3] + [42] + base[3:], base[:10] + [7, 8] + base[10:]]
    return [variants[i % len(variants)] + ([] if i < len(variants) else [100 + i]) for i in range(n)]


def _oracle_solo(orc, desc, params, prompt, spk, lang, cap, G):
    s = orc.orc_create(desc)
    orc.orc_set_params(s, C.byref(params))
    want = np.zeros((cap, G), dtype=np.int32)
    ns, nf = C.c_int(0), C.c_int(0)
    tk = np.array(prompt, dtype=np.int32)
    ptr = orc.orc_generate(s, B.iptr(tk), len(tk), spk, lang, C.byref(ns), B.iptr(want), cap, C.byref(nf))
    pcm = B.take_floats(orc.orc_free, ptr, ns.value) if ptr else np.zeros(0, dtype=np.float32)
    orc.orc_destroy(s)
    return nf.value, want, pcm


def test_generate_queue_admits_into_freed_rows_eos(cuda, orc):
    """Continuous batching (qwen_tts_cuda_generate_queue): 9 utterances queue for 3 lock-step rows, no frame budget, the EOS row
    of the codec head boosted so that utterances end at different frames.  A row whose utterance has ended is retired at the next
    poll, its KV pages go back to the free list and the next utterance is admitted into it (prefilled on its own while the other
    rows are mid-utterance).  Every utterance's frame count and codes equal the oracle's solo run with seed 42 + u."""
    _setup_batch_api(cuda)
    spec = mg.PRESETS["tiny"]
    main, cod = mg.build_tensors(spec, seed=0)
    w = mg.bf16_to_f32(main["talker.codec_head.weight"]).copy()
    w[mg.CODEC_IDS["codec_eos_token_id"]] *= 3.0
    main["talker.codec_head.weight"] = mg.f32_to_bf16(w)
    hm = B.HostModel(spec, main, cod)
    G = spec["groups"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 4)
    assert ctx
    cap, n_utt, rows = 30, 9, 3
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.default_params(max_new_tokens=cap)))
    prompts = _queue_prompts(n_utt)
    spk = np.full(n_utt, mg.SPEAKERS["aiden"], dtype=np.int32)
    lang = np.full(n_utt, mg.LANGUAGES["english"], dtype=np.int32)
    toks = np.array(sum(prompts, []), dtype=np.int32)
    offs = np.array([0] + list(np.cumsum([len(q) for q in prompts])), dtype=np.int32)
    codes = np.zeros((n_utt, cap, G), dtype=np.int32)
    nfr = np.full(n_utt, -1, dtype=np.int32)
    perf0 = B.Perf()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf0))
    rc = cuda.qwen_tts_cuda_generate_queue(ctx, n_utt, rows, B.iptr(toks), B.iptr(offs), B.iptr(spk), B.iptr(lang), None, B.iptr(codes), B.iptr(nfr), None, cap)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    perf = B.Perf()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    assert perf.queue_admitted == n_utt
    assert perf.kv_page_allocs - perf0.kv_page_allocs >= n_utt      # every admission took its pages from the free list again
    want_n = []
    for u in range(n_utt):
        nf, want, _ = _oracle_solo(orc, hm.desc, B.default_params(max_new_tokens=cap, sample_seed=42 + u), prompts[u], int(spk[u]), int(lang[u]), cap, G)
        want_n.append(nf)
        assert int(nfr[u]) == nf, (u, list(nfr), want_n)
        assert np.array_equal(codes[u][:nf], want[:nf]), u
    assert len(set(want_n)) > 2 and min(want_n) < cap, want_n       # the utterances really stop at different frames
    cuda.qwen_tts_cuda_release(ctx)


def test_generate_queue_fixed_budget_with_pcm(cuda, orc, tiny):
    """the same with fixed frame budgets - one per utterance (6, 3, 5, 6, 2 frames: rows retire at different frames, exactly at
    their budget) - and PCM: 5 utterances through 2 rows; codes equal the oracle's, waveform >= 80 dB; then a plain
    generate_batch on the same context still equals the oracle (the queue leaves rows parked, the batch call re-arms them)"""
    _setup_batch_api(cuda)
    G = tiny.spec["groups"]
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 2)
    assert ctx
    frames, n_utt, rows = 6, 5, 2
    budgets = np.array([6, 3, 5, 6, 2], dtype=np.int32)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.default_params(fixed_codec_tokens=frames)))
    prompts = _queue_prompts(n_utt)
    spk = np.array([mg.SPEAKERS["aiden"], -1, mg.SPEAKERS["vivian"], mg.SPEAKERS["aiden"], -1], dtype=np.int32)
    lang = np.array([mg.LANGUAGES["english"], mg.LANGUAGES["chinese"], -1, -1, mg.LANGUAGES["english"]], dtype=np.int32)
    toks = np.array(sum(prompts, []), dtype=np.int32)
    offs = np.array([0] + list(np.cumsum([len(q) for q in prompts])), dtype=np.int32)
    codes = np.zeros((n_utt, frames, G), dtype=np.int32)
    nfr = np.zeros(n_utt, dtype=np.int32)
    pcm = np.zeros((n_utt, frames * 1920), dtype=np.float32)
    rc = cuda.qwen_tts_cuda_generate_queue(ctx, n_utt, rows, B.iptr(toks), B.iptr(offs), B.iptr(spk), B.iptr(lang), B.iptr(budgets), B.iptr(codes), B.iptr(nfr), B.fptr(pcm), frames)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    assert list(nfr) == list(budgets)
    for u in range(n_utt):
        k = int(budgets[u])
        nf, want, want_pcm = _oracle_solo(orc, tiny.desc, B.default_params(fixed_codec_tokens=k, sample_seed=42 + u), prompts[u], int(spk[u]), int(lang[u]), frames, G)
        assert nf == k and np.array_equal(codes[u][:k], want[:k]), (u, codes[u], want)
        got_pcm = pcm[u][: len(want_pcm)]
        err = ((got_pcm.astype(np.float64) - want_pcm) ** 2).sum()
        snr = 10 * np.log10((want_pcm.astype(np.float64) ** 2).sum() / max(err, 1e-300))
        assert len(want_pcm) == k * 1920 and snr >= 80.0, (u, snr)
    codes2 = np.zeros((2, frames, G), dtype=np.int32)
    rc = cuda.qwen_tts_cuda_generate_batch(ctx, 2, B.iptr(toks), B.iptr(offs), B.iptr(spk), B.iptr(lang), B.iptr(codes2), B.iptr(nfr), None, frames)
    assert rc == 0, cuda.qwen_tts_cuda_last_error()
    for b in range(2):
        nf, want, _ = _oracle_solo(orc, tiny.desc, B.default_params(fixed_codec_tokens=frames, sample_seed=42 + b), prompts[b], int(spk[b]), int(lang[b]), frames, G)
        assert np.array_equal(codes2[b], want), b
    cuda.qwen_tts_cuda_release(ctx)
