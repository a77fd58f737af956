"""Parity of the CUDA library (through its C ABI, include/qwen_tts_cuda.h) against the oracle and the
golden vectors the unmodified reference produced.  Needs a B200: run with -m gpu.

Bars (BASELINE.json north_star):
  * strict arithmetic: logits, hidden states, sampled codes BIT-IDENTICAL to the C path;
  * fast arithmetic: talker logits within 1e-2 relative (measured ~1e-5), codes equal on these cases;
  * waveform: SNR >= 80 dB against the C path's PCM for the same codes.
"""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import FIXTURE_TOKENS, ROOT, get_model
from harness import binding as B, modelgen as mg

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "ref_tiny_seed0.npz"))
CSV = ",".join(str(t) for t in FIXTURE_TOKENS).encode()
TOK = np.array(FIXTURE_TOKENS, dtype=np.int32)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def snr_db(ref, got):
    ref, got = ref.astype(np.float64), got.astype(np.float64)
    return 10 * np.log10((ref ** 2).sum() / max(((ref - got) ** 2).sum(), 1e-300))


@pytest.fixture(scope="module")
def tiny_ctx(cuda, tiny):
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    assert cuda.qwen_tts_cuda_set_mode(ctx, 0) == 0      # this module: one kernel per op (mode 1 is tests/test_gpu_fused.py)
    yield ctx
    cuda.qwen_tts_cuda_release(ctx)


# ---------------------------------------------------------------- kernel level
@pytest.mark.parametrize("rows,cols", [(96, 256), (2048, 1024), (1024, 2048), (3072, 1024), (1024, 3072), (40, 64), (2048, 2048)])
def test_matvec_bf16(cuda, orc, rows, cols):
    rng = np.random.default_rng(rows * 7 + cols)
    A = mg.f32_to_bf16(rng.standard_normal((rows, cols)).astype(np.float32) / np.sqrt(cols))
    x = rng.standard_normal(cols).astype(np.float32)
    want = np.zeros(rows, dtype=np.float32)
    orc.orc_matvec_bf16(B.fptr(want), B.uptr(A), B.fptr(x), rows, cols)
    got = np.zeros(rows, dtype=np.float32)
    cuda.qwen_tts_cuda_kernel_arith(1)
    assert cuda.qwen_tts_cuda_matvec_bf16(B.fptr(got), B.uptr(A), B.fptr(x), rows, cols) == 0
    assert np.array_equal(bits(got), bits(want)), "strict GEMV must reproduce the reference's sequential sum bit for bit"
    cuda.qwen_tts_cuda_kernel_arith(0)
    assert cuda.qwen_tts_cuda_matvec_bf16(B.fptr(got), B.uptr(A), B.fptr(x), rows, cols) == 0
    assert np.abs(got - want).max() <= 1e-5 * max(1.0, np.abs(want).max())


def test_matvec_gold(cuda):
    y = np.zeros(96, dtype=np.float32)
    cuda.qwen_tts_cuda_kernel_arith(1)
    assert cuda.qwen_tts_cuda_matvec_bf16(B.fptr(y), B.uptr(np.ascontiguousarray(GOLD["mv_A"])), B.fptr(np.ascontiguousarray(GOLD["mv_x"])), 96, 256) == 0
    cuda.qwen_tts_cuda_kernel_arith(0)
    assert np.array_equal(bits(y), bits(GOLD["mv_y"]))


@pytest.mark.parametrize("inter,hidden", [(512, 256), (3072, 1024), (200, 128)])
def test_swiglu_and_rmsnorm(cuda, orc, inter, hidden):
    rng = np.random.default_rng(inter)
    W = mg.f32_to_bf16(rng.standard_normal((2 * inter, hidden)).astype(np.float32) / np.sqrt(hidden))
    x = rng.standard_normal(hidden).astype(np.float32)
    w = (1 + 0.1 * rng.standard_normal(hidden)).astype(np.float32)
    want, got = np.zeros(inter, dtype=np.float32), np.zeros(inter, dtype=np.float32)
    orc.orc_swiglu_matvec_bf16(B.fptr(want), B.uptr(W), B.fptr(x), inter, hidden)
    nw, ng = np.zeros(hidden, dtype=np.float32), np.zeros(hidden, dtype=np.float32)
    orc.orc_rms_norm(B.fptr(nw), B.fptr(x), B.fptr(w), hidden, C.c_float(1e-6))
    cuda.qwen_tts_cuda_kernel_arith(1)
    if hidden % 64 == 0:
        assert cuda.qwen_tts_cuda_swiglu_matvec_bf16(B.fptr(got), B.uptr(W), B.fptr(x), inter, hidden) == 0
        assert np.array_equal(bits(got), bits(want))
    assert cuda.qwen_tts_cuda_rms_norm(B.fptr(ng), B.fptr(x), B.fptr(w), hidden, C.c_float(1e-6)) == 0
    assert np.array_equal(bits(ng), bits(nw))
    cuda.qwen_tts_cuda_kernel_arith(0)
    assert cuda.qwen_tts_cuda_swiglu_matvec_bf16(B.fptr(got), B.uptr(W), B.fptr(x), inter, hidden) == 0
    assert np.abs(got - want).max() <= 2e-5 * max(1.0, np.abs(want).max())
    assert cuda.qwen_tts_cuda_rms_norm(B.fptr(ng), B.fptr(x), B.fptr(w), hidden, C.c_float(1e-6)) == 0
    assert np.abs(ng - nw).max() <= 1e-5 * np.abs(nw).max()


def test_sampler_known_answers(cuda):
    """token AND RNG state after six consecutive draws, against kernel_sample_top_k of the reference"""
    lg = np.ascontiguousarray(GOLD["sample_logits"])
    cfgs = ((1, 1.0, 1.0, 42), (50, 1.0, 0.9, 42), (50, 1.0, 0.9, 43), (5, 1.0, 2.0, 1),
            (50, 0.8, 0.9, 42), (0, 0.5, 1.0, 9), (3072, 1.0, 1.0, 3), (50, 1.0, 0.0, 42))
    for row, (k, tp, temp, seed) in zip(GOLD["sample_picks"], cfgs):
        st = np.array([seed], dtype=np.float32)
        seq = [cuda.qwen_tts_cuda_sample_top_k(B.fptr(lg), 3072, k, C.c_float(tp), C.c_float(temp), B.fptr(st)) for _ in range(6)]
        assert seq == list(row[:6]) and int(st.view(np.uint32)[0]) == row[6], (k, tp, temp, seed, seq, list(row))


@pytest.mark.parametrize("k,vocab,seed", [(50, 2048, 1), (50, 3072, 2), (7, 2048, 3), (1000, 3072, 4), (1500, 2048, 5), (2, 100, 6)])
def test_sampler_vs_oracle_random(cuda, orc, k, vocab, seed):
    rng = np.random.default_rng(seed)
    for trial in range(4):
        lg = (rng.standard_normal(vocab) * 3).astype(np.float32)
        if trial == 1:
            lg[rng.integers(0, vocab, 40)] = lg.max()          # many exact ties at the top
        if trial == 2:
            lg = np.round(lg)                                    # ties everywhere, also at the k-th place
        if trial == 3:
            lg[: vocab // 2] = -1e9
        a, b = np.array([seed + trial], dtype=np.float32), np.array([seed + trial], dtype=np.float32)
        for _ in range(3):
            want = orc.orc_sample(B.fptr(lg), vocab, k, C.c_float(1.0), C.c_float(0.9), B.fptr(a))
            got = cuda.qwen_tts_cuda_sample_top_k(B.fptr(lg), vocab, k, C.c_float(1.0), C.c_float(0.9), B.fptr(b))
            assert got == want and a.view(np.uint32)[0] == b.view(np.uint32)[0], (trial, k, vocab)


def test_conv_kernels_vs_reference_vectors(cuda):
    for name in ("conv_k7d3", "conv_dw", "conv_k1", "conv_short"):
        ci, co, k, L, dil, grp = (int(v) for v in GOLD[f"{name}_p"])
        yo = np.zeros((co, L), dtype=np.float32)
        assert cuda.qwen_tts_cuda_causal_conv1d(B.fptr(yo), B.fptr(np.ascontiguousarray(GOLD[f"{name}_x"])),
                                                B.fptr(np.ascontiguousarray(GOLD[f"{name}_w"])),
                                                B.fptr(np.ascontiguousarray(GOLD[f"{name}_b"])), ci, co, k, L, dil, grp) == 0
        want = GOLD[f"{name}_y"]
        assert np.abs(yo - want).max() <= 1e-5 * max(1.0, np.abs(want).max()), name
    for name in ("tconv_16_8", "tconv_2_2"):
        ci, co, k, s, L = (int(v) for v in GOLD[f"{name}_p"])
        want = GOLD[f"{name}_y"]
        yo = np.zeros((co, L * s), dtype=np.float32)
        nl = C.c_int(0)
        assert cuda.qwen_tts_cuda_transposed_conv1d(B.fptr(yo), B.fptr(np.ascontiguousarray(GOLD[f"{name}_x"])),
                                                    B.fptr(np.ascontiguousarray(GOLD[f"{name}_w"])),
                                                    B.fptr(np.ascontiguousarray(GOLD[f"{name}_b"])), ci, co, k, s, L, C.byref(nl)) == 0
        assert nl.value == L * s
        assert np.abs(yo - want).max() <= 1e-5 * max(1.0, np.abs(want).max()), name


@pytest.mark.parametrize("ci,co,k,L,dil", [(96, 96, 7, 1000, 9), (192, 96, 7, 333, 1), (64, 200, 3, 129, 1), (130, 70, 1, 500, 1)])
def test_conv_shapes_vs_oracle(cuda, orc, ci, co, k, L, dil):
    rng = np.random.default_rng(ci + co)
    x = rng.standard_normal((ci, L)).astype(np.float32)
    w = (rng.standard_normal((co, ci, k)) / np.sqrt(ci * k)).astype(np.float32)
    b = rng.standard_normal(co).astype(np.float32)
    want, got = np.zeros((co, L), dtype=np.float32), np.zeros((co, L), dtype=np.float32)
    orc.orc_causal_conv1d(B.fptr(want), B.fptr(x), B.fptr(w), B.fptr(b), ci, co, k, L, dil, 1)
    assert cuda.qwen_tts_cuda_causal_conv1d(B.fptr(got), B.fptr(x), B.fptr(w), B.fptr(b), ci, co, k, L, dil, 1) == 0
    assert snr_db(want, got) > 110


@pytest.mark.parametrize("ci,co,s,L", [(96, 48, 8, 77), (64, 32, 5, 130), (40, 40, 3, 64)])
def test_tconv_shapes_vs_oracle(cuda, orc, ci, co, s, L):
    rng = np.random.default_rng(ci + s)
    x = rng.standard_normal((ci, L)).astype(np.float32)
    w = (rng.standard_normal((ci, co, 2 * s)) / np.sqrt(ci * 2)).astype(np.float32)
    b = rng.standard_normal(co).astype(np.float32)
    want, got = np.zeros((co, L * s + 2 * s), dtype=np.float32), np.zeros((co, L * s), dtype=np.float32)
    nl = C.c_int(0)
    orc.orc_transposed_conv1d(B.fptr(want), B.fptr(x), B.fptr(w), B.fptr(b), ci, co, 2 * s, s, L, C.byref(nl))
    want = want.reshape(-1)[: co * nl.value].reshape(co, nl.value)
    assert cuda.qwen_tts_cuda_transposed_conv1d(B.fptr(got), B.fptr(x), B.fptr(w), B.fptr(b), ci, co, 2 * s, s, L, C.byref(nl)) == 0
    assert snr_db(want, got) > 110


def test_snake_and_matmul(cuda, orc):
    rng = np.random.default_rng(5)
    x = (rng.standard_normal((24, 301)) * 2).astype(np.float32)
    a, ib = np.exp(0.2 * rng.standard_normal(24)).astype(np.float32), np.exp(0.2 * rng.standard_normal(24)).astype(np.float32)
    want, got = np.zeros_like(x), np.zeros_like(x)
    orc.orc_snake_beta(B.fptr(want), B.fptr(x), B.fptr(a), B.fptr(ib), 24, 301)
    assert cuda.qwen_tts_cuda_snake_beta(B.fptr(got), B.fptr(x), B.fptr(a), B.fptr(ib), 24, 301) == 0
    assert np.abs(got - want).max() < 1e-5
    A, Bm = rng.standard_normal((77, 64)).astype(np.float32), rng.standard_normal((130, 64)).astype(np.float32)
    cw, cg = np.zeros((77, 130), dtype=np.float32), np.zeros((77, 130), dtype=np.float32)
    orc.orc_matmul_f32(B.fptr(cw), B.fptr(A), B.fptr(Bm), 77, 130, 64)
    assert cuda.qwen_tts_cuda_matmul_f32(B.fptr(cg), B.fptr(A), B.fptr(Bm), 77, 130, 64) == 0
    assert np.abs(cg - cw).max() < 1e-4
    Bb = mg.f32_to_bf16(Bm)
    orc.orc_matmul_bf16(B.fptr(cw), B.fptr(A), B.uptr(Bb), 77, 130, 64)
    cuda.qwen_tts_cuda_kernel_arith(1)
    assert cuda.qwen_tts_cuda_matmul_bf16(B.fptr(cg), B.fptr(A), B.uptr(Bb), 77, 130, 64) == 0
    cuda.qwen_tts_cuda_kernel_arith(0)
    assert np.array_equal(bits(cg), bits(cw))


# ---------------------------------------------------------------- stage API
def _stage_run(cuda, ctx, spec, strict):
    H, V = spec["talker"]["hidden"], spec["talker"]["vocab"]
    cuda.qwen_tts_cuda_set_arith(ctx, strict)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.greedy_params()))
    cuda.qwen_tts_cuda_reset(ctx)
    emb = np.ascontiguousarray(GOLD["stage_embeds"])
    assert cuda.qwen_tts_cuda_talker_prefill(ctx, B.fptr(emb), len(emb)) == 0
    hid = np.zeros(H, dtype=np.float32)
    assert cuda.qwen_tts_cuda_get_hidden(ctx, B.fptr(hid)) == 0
    assert cuda.qwen_tts_cuda_get_kv_len(ctx) == len(emb)
    steps = np.ascontiguousarray(GOLD["stage_steps"])
    logits, hids = np.zeros((len(steps), V), dtype=np.float32), np.zeros((len(steps), H), dtype=np.float32)
    for i in range(len(steps)):
        assert cuda.qwen_tts_cuda_talker_forward(ctx, B.fptr(steps[i]), B.fptr(logits[i])) == 0
        assert cuda.qwen_tts_cuda_get_hidden(ctx, B.fptr(hids[i])) == 0
    assert cuda.qwen_tts_cuda_get_kv_len(ctx) == len(emb) + len(steps)
    subs = []
    for i, (pp, fc) in enumerate(((B.greedy_params(), 17), (B.default_params(), 901),
                                  (B.default_params(sample_seed=5, subtalker_top_k=4), 2047))):
        cuda.qwen_tts_cuda_set_params(ctx, C.byref(pp))
        out = np.zeros(32, dtype=np.int32)
        assert cuda.qwen_tts_cuda_subtalker_generate(ctx, B.fptr(np.ascontiguousarray(GOLD["stage_hiddens"][i])), fc, B.iptr(out)) == 0
        subs.append(out[:16].copy())
    return hid, logits, hids, np.array(subs)


def test_stage_api_strict_bit_exact(cuda, tiny, tiny_ctx):
    hid, logits, hids, subs = _stage_run(cuda, tiny_ctx, tiny.spec, 1)
    assert np.array_equal(bits(hid), bits(GOLD["stage_prefill_hidden"]))
    assert np.array_equal(bits(logits), bits(GOLD["stage_logits"]))
    assert np.array_equal(bits(hids), bits(GOLD["stage_hiddens"]))
    assert np.array_equal(subs, GOLD["stage_sub_codes"])


def test_stage_api_fast_within_tolerance(cuda, tiny, tiny_ctx):
    hid, logits, hids, subs = _stage_run(cuda, tiny_ctx, tiny.spec, 0)
    want = GOLD["stage_logits"]
    rel = np.abs(logits - want).max() / np.abs(want).max()
    assert rel <= 1e-2, rel          # north_star bar
    assert rel <= 2e-4, rel          # what fp32 reordering actually costs
    assert np.abs(hid - GOLD["stage_prefill_hidden"]).max() <= 2e-4 * np.abs(hid).max()
    assert np.array_equal(subs, GOLD["stage_sub_codes"])


# ---------------------------------------------------------------- whole utterance
GEN_CASES = [
    ("greedy", B.greedy_params(fixed_codec_tokens=24), b"aiden", b"english"),
    ("sampled", B.default_params(fixed_codec_tokens=24), b"aiden", b"english"),
    ("sampled_seed7", B.default_params(fixed_codec_tokens=16, sample_seed=7, temperature=1.3, top_k=8,
                                       subtalker_top_k=3, repetition_penalty=1.3), b"AIDEN", b"English"),
    ("nospeaker", B.greedy_params(fixed_codec_tokens=8), None, None),
    ("maxtok", B.greedy_params(max_new_tokens=5), b"aiden", b"english"),
]


def _generate(cuda, ctx, params, spk, lang, n):
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(params))
    ns, nf = C.c_int(0), C.c_int(0)
    codes = np.zeros((n, 16), dtype=np.int32)
    ptr = cuda.qwen_tts_cuda_generate(ctx, CSV, spk, lang, C.byref(ns), B.iptr(codes), n, C.byref(nf))
    assert ptr, cuda.qwen_tts_cuda_last_error()
    pcm = B.take_floats(cuda.qwen_tts_cuda_free_host, ptr, ns.value)
    return codes[: nf.value], pcm


@pytest.mark.parametrize("name,params,spk,lang", GEN_CASES)
@pytest.mark.parametrize("strict", [1, 0])
def test_generate_codes_bit_exact(cuda, tiny_ctx, name, params, spk, lang, strict):
    want_codes, want_pcm = GOLD[f"gen_{name}_codes"], GOLD[f"gen_{name}_pcm"]
    cuda.qwen_tts_cuda_set_arith(tiny_ctx, strict)
    codes, pcm = _generate(cuda, tiny_ctx, params, spk, lang, len(want_codes))
    assert codes.shape == want_codes.shape
    assert np.array_equal(codes, want_codes), f"first mismatch at frame {np.argwhere(codes != want_codes)[0]}"
    assert pcm.shape == want_pcm.shape
    assert snr_db(want_pcm, pcm) >= 80.0
    perf = B.Perf()
    cuda.qwen_tts_cuda_get_perf(tiny_ctx, C.byref(perf))
    assert perf.codec_tokens == len(want_codes) and perf.stop_reason == 2


def test_generate_frames_host_prompt(cuda, orc, tiny, tiny_ctx):
    """generate_frames with the prompt built on the host by the oracle: same codes as the whole-utterance path"""
    s = orc.orc_create(tiny.desc)
    pre, tr, pad = C.c_void_p(), C.c_void_p(), C.c_void_p()
    plen, ntr = C.c_int(0), C.c_int(0)
    assert orc.orc_build_prompt(s, B.iptr(TOK), len(TOK), 3010, 2050, C.byref(pre), C.byref(plen), C.byref(tr), C.byref(ntr), C.byref(pad)) == 0
    p = B.default_params(fixed_codec_tokens=24)
    cuda.qwen_tts_cuda_set_arith(tiny_ctx, 1)
    cuda.qwen_tts_cuda_set_params(tiny_ctx, C.byref(p))
    codes = np.zeros((24, 16), dtype=np.int32)
    n = cuda.qwen_tts_cuda_generate_frames(tiny_ctx, C.cast(pre, B.f32p), plen.value, C.cast(tr, B.f32p), ntr.value, C.cast(pad, B.f32p),
                                           B.iptr(codes), 24)
    assert n == 24 and np.array_equal(codes, GOLD["gen_sampled_codes"])
    for q in (pre, tr, pad):
        orc.orc_free(q)
    orc.orc_destroy(s)


def test_eos_stop(cuda, orc, tiny, tiny_ctx):
    """EOS path: no fixed budget; the stop step and the codes before it must match the oracle's"""
    # make EOS likely: strong repetition penalty pushes mass around; compare against the oracle whatever happens
    p = B.default_params(max_new_tokens=40, temperature=3.0, top_k=3000, sample_seed=11)
    s = orc.orc_create(tiny.desc)
    orc.orc_set_params(s, C.byref(p))
    ns, nf = C.c_int(0), C.c_int(0)
    want = np.zeros((40, 16), dtype=np.int32)
    ptr = orc.orc_generate(s, B.iptr(TOK), len(TOK), 3010, 2050, C.byref(ns), B.iptr(want), 40, C.byref(nf))
    if ptr:
        orc.orc_free(ptr)
    orc.orc_destroy(s)
    cuda.qwen_tts_cuda_set_arith(tiny_ctx, 1)
    cuda.qwen_tts_cuda_set_params(tiny_ctx, C.byref(p))
    codes = np.zeros((40, 16), dtype=np.int32)
    ns2, nf2 = C.c_int(0), C.c_int(0)
    ptr = cuda.qwen_tts_cuda_generate(tiny_ctx, CSV, b"aiden", b"english", C.byref(ns2), B.iptr(codes), 40, C.byref(nf2))
    if ptr:
        cuda.qwen_tts_cuda_free_host(ptr)
    assert nf2.value == nf.value and np.array_equal(codes[: nf.value], want[: nf.value])


def test_codec_only(cuda, tiny_ctx):
    codes = np.ascontiguousarray(GOLD["codec_codes"])
    ns = C.c_int(0)
    ptr = cuda.qwen_tts_cuda_codec_decode(tiny_ctx, B.iptr(codes), len(codes), C.byref(ns))
    assert ptr and ns.value == len(codes) * 1920
    pcm = B.take_floats(cuda.qwen_tts_cuda_free_host, ptr, ns.value)
    assert snr_db(GOLD["codec_pcm"], pcm) >= 80.0
    assert not cuda.qwen_tts_cuda_codec_decode(tiny_ctx, B.iptr(codes), 0, C.byref(ns)) and ns.value == 0


def test_error_behaviour(cuda, tiny_ctx):
    ns = C.c_int(7)
    assert not cuda.qwen_tts_cuda_generate(tiny_ctx, b"1,2,3", b"aiden", b"english", C.byref(ns), None, 0, None) and ns.value == 0
    assert not cuda.qwen_tts_cuda_generate(tiny_ctx, b"1,2,x", b"aiden", b"english", C.byref(ns), None, 0, None)
    assert cuda.qwen_tts_cuda_talker_forward(tiny_ctx, None, None) == -1


# ---------------------------------------------------------------- real head geometry
def test_small_model_strict_and_fast(cuda, orc):
    """H=1024, 16 Q / 8 KV heads of 128, 4+2 layers: the real GQA geometry at a size the oracle runs in seconds"""
    hm = get_model(orc, "small")
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx
    p = B.default_params(fixed_codec_tokens=6)
    s = orc.orc_create(hm.desc)
    orc.orc_set_params(s, C.byref(p))
    ns, nf = C.c_int(0), C.c_int(0)
    want = np.zeros((6, 16), dtype=np.int32)
    ptr = orc.orc_generate(s, B.iptr(TOK), len(TOK), 3010, 2050, C.byref(ns), B.iptr(want), 6, C.byref(nf))
    want_pcm = B.take_floats(orc.orc_free, ptr, ns.value)
    orc.orc_destroy(s)
    for strict in (1, 0):
        cuda.qwen_tts_cuda_set_arith(ctx, strict)
        codes, pcm = _generate(cuda, ctx, p, b"aiden", b"english", 6)
        assert np.array_equal(codes, want), strict
        assert snr_db(want_pcm, pcm) >= 80.0
    cuda.qwen_tts_cuda_release(ctx)
