"""The oracle (oracle/liboracle.so, my CPU restatement) against golden vectors produced by the
UNMODIFIED reference (scripts/make_golden.py -> tests/golden/ref_tiny_seed0.npz).  Bit-exact.
Runs without /root/reference, so it also pins the oracle on the GPU box."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import FIXTURE_TOKENS, ROOT
from harness import binding as B

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "ref_tiny_seed0.npz"))
TOK = np.array(FIXTURE_TOKENS, dtype=np.int32)


def _generate(orc, hm, params, n, spk=3010, lang=2050):
    s = orc.orc_create(hm.desc)
    orc.orc_set_params(s, C.byref(params))
    ns, nf = C.c_int(0), C.c_int(0)
    codes = np.zeros((n, 16), dtype=np.int32)
    ptr = orc.orc_generate(s, B.iptr(TOK), len(TOK), spk, lang, C.byref(ns), B.iptr(codes), n, C.byref(nf))
    pcm = B.take_floats(orc.orc_free, ptr, ns.value)
    orc.orc_destroy(s)
    return codes[:nf.value], pcm


@pytest.mark.parametrize("name,params,spk,lang", [
    ("greedy", B.greedy_params(fixed_codec_tokens=24), 3010, 2050),
    ("sampled", B.default_params(fixed_codec_tokens=24), 3010, 2050),
    ("sampled_seed7", B.default_params(fixed_codec_tokens=16, sample_seed=7, temperature=1.3, top_k=8,
                                       subtalker_top_k=3, repetition_penalty=1.3), 3010, 2050),
    ("nospeaker", B.greedy_params(fixed_codec_tokens=8), -1, -1),
    ("maxtok", B.greedy_params(max_new_tokens=5), 3010, 2050),
])
def test_generate_matches_reference(orc, tiny, name, params, spk, lang):
    want_codes, want_pcm = GOLD[f"gen_{name}_codes"], GOLD[f"gen_{name}_pcm"]
    codes, pcm = _generate(orc, tiny, params, len(want_codes), spk, lang)
    assert np.array_equal(codes, want_codes)
    assert pcm.shape == want_pcm.shape and np.array_equal(pcm.view(np.uint32), want_pcm.view(np.uint32))


def test_stage_api_matches_reference(orc, tiny):
    H, V = tiny.spec["talker"]["hidden"], tiny.spec["talker"]["vocab"]
    s = orc.orc_create(tiny.desc)
    orc.orc_set_params(s, C.byref(B.greedy_params()))
    orc.orc_reset(s)
    emb = np.ascontiguousarray(GOLD["stage_embeds"])
    orc.orc_talker_prefill(s, B.fptr(emb), len(emb))
    hid = np.ctypeslib.as_array(orc.orc_hidden(s), shape=(H,)).copy()
    assert np.array_equal(hid.view(np.uint32), GOLD["stage_prefill_hidden"].view(np.uint32))
    assert orc.orc_kv_len(s) == len(emb)
    steps = np.ascontiguousarray(GOLD["stage_steps"])
    hids = []
    for i in range(len(steps)):
        lg = np.zeros(V, dtype=np.float32)
        orc.orc_talker_forward(s, B.fptr(steps[i]), B.fptr(lg))
        assert np.array_equal(lg.view(np.uint32), GOLD["stage_logits"][i].view(np.uint32)), i
        hids.append(np.ctypeslib.as_array(orc.orc_hidden(s), shape=(H,)).copy())
    assert orc.orc_kv_len(s) == len(emb) + len(steps)
    for i, (pp, fc) in enumerate(((B.greedy_params(), 17), (B.default_params(), 901),
                                  (B.default_params(sample_seed=5, subtalker_top_k=4), 2047))):
        orc.orc_set_params(s, C.byref(pp))
        out = np.zeros(32, dtype=np.int32)
        orc.orc_subtalker_generate(s, B.fptr(hids[i]), fc, B.iptr(out))
        assert np.array_equal(out[:16], GOLD["stage_sub_codes"][i]), i
    orc.orc_destroy(s)


def test_codec_matches_reference(orc, tiny):
    s = orc.orc_create(tiny.desc)
    codes = np.ascontiguousarray(GOLD["codec_codes"])
    ns = C.c_int(0)
    ptr = orc.orc_codec_decode(s, B.iptr(codes), len(codes), C.byref(ns))
    pcm = B.take_floats(orc.orc_free, ptr, ns.value)
    assert ns.value == len(codes) * 1920
    assert np.array_equal(pcm.view(np.uint32), GOLD["codec_pcm"].view(np.uint32))
    # error behaviour of qwen_tts_codec_decode: NULL and 0 samples on empty input (c/qwen_tts_codec.c:583-587)
    assert not orc.orc_codec_decode(s, B.iptr(codes), 0, C.byref(ns)) and ns.value == 0
    orc.orc_destroy(s)


def test_sampler_known_answers(orc):
    lg = np.ascontiguousarray(GOLD["sample_logits"])
    cfgs = ((1, 1.0, 1.0, 42), (50, 1.0, 0.9, 42), (50, 1.0, 0.9, 43), (5, 1.0, 2.0, 1),
            (50, 0.8, 0.9, 42), (0, 0.5, 1.0, 9), (3072, 1.0, 1.0, 3), (50, 1.0, 0.0, 42))
    for row, (k, tp, temp, seed) in zip(GOLD["sample_picks"], cfgs):
        st = np.array([seed], dtype=np.float32)
        seq = [orc.orc_sample(B.fptr(lg), 3072, k, C.c_float(tp), C.c_float(temp), B.fptr(st)) for _ in range(6)]
        assert seq == list(row[:6]) and int(st.view(np.uint32)[0]) == row[6], (k, tp, temp, seed)
    # greedy tie: the lower index wins (c/qwen_tts_kernels.c:437-449)
    assert GOLD["sample_picks"][0][0] == 100


def test_kernel_known_answers(orc):
    y = np.zeros(96, dtype=np.float32)
    orc.orc_matvec_bf16(B.fptr(y), B.uptr(np.ascontiguousarray(GOLD["mv_A"])), B.fptr(np.ascontiguousarray(GOLD["mv_x"])), 96, 256)
    assert np.array_equal(y.view(np.uint32), GOLD["mv_y"].view(np.uint32))
    for name in ("conv_k7d3", "conv_dw", "conv_k1", "conv_short"):
        ci, co, k, L, dil, grp = (int(v) for v in GOLD[f"{name}_p"])
        yo = np.zeros((co, L), dtype=np.float32)
        orc.orc_causal_conv1d(B.fptr(yo), B.fptr(np.ascontiguousarray(GOLD[f"{name}_x"])),
                              B.fptr(np.ascontiguousarray(GOLD[f"{name}_w"])),
                              B.fptr(np.ascontiguousarray(GOLD[f"{name}_b"])), ci, co, k, L, dil, grp)
        assert np.array_equal(yo.view(np.uint32), GOLD[f"{name}_y"].view(np.uint32)), name
    for name in ("tconv_16_8", "tconv_2_2"):
        ci, co, k, s, L = (int(v) for v in GOLD[f"{name}_p"])
        want = GOLD[f"{name}_y"]
        yo = np.zeros((co, L * s + k), dtype=np.float32)
        nl = C.c_int(0)
        orc.orc_transposed_conv1d(B.fptr(yo), B.fptr(np.ascontiguousarray(GOLD[f"{name}_x"])),
                                  B.fptr(np.ascontiguousarray(GOLD[f"{name}_w"])),
                                  B.fptr(np.ascontiguousarray(GOLD[f"{name}_b"])), ci, co, k, s, L, C.byref(nl))
        assert nl.value == want.shape[1] == L * s
        got = yo.reshape(-1)[:co * nl.value].reshape(co, nl.value)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), name


def test_host_prep_matches_oracle(orc):
    """harness/binding.py does the reference's load-time preprocessing (c/qwen_tts.c:577-602) in numpy + libm without the
    checker; hold it bit for bit to the oracle's restatement (orc_prep_codebook / orc_prep_snake)."""
    import ctypes as C
    from harness import binding as B
    rng = np.random.default_rng(3)
    es = rng.standard_normal((2048, 64)).astype(np.float32)
    us = rng.uniform(0.0, 1.5, size=2048).astype(np.float32)
    us[:7] = [0.0, 1e-6, 1e-5, 9.9e-6, 1.0, 3e-5, 2.0]
    want = np.empty_like(es)
    orc.orc_prep_codebook.argtypes = [B.f32p, B.f32p, B.f32p, C.c_int, C.c_int]
    orc.orc_prep_codebook(B.fptr(want), B.fptr(es), B.fptr(us), 2048, 64)
    assert np.array_equal(B.prep_codebook(es, us).view(np.uint32), want.view(np.uint32))
    a = (rng.standard_normal(1536) * 0.7).astype(np.float32)
    b = (rng.standard_normal(1536) * 0.7).astype(np.float32)
    wa, wb = a.copy(), b.copy()
    orc.orc_prep_snake.argtypes = [B.f32p, B.f32p, C.c_int]
    orc.orc_prep_snake(B.fptr(wa), B.fptr(wb), a.size)
    ga, gb = B.prep_snake(a, b)
    assert np.array_equal(ga.view(np.uint32), wa.view(np.uint32)) and np.array_equal(gb.view(np.uint32), wb.view(np.uint32))
