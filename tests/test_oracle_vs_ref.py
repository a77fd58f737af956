"""The oracle against the UNMODIFIED reference compiled into oracle/_ref/libqtts_ref.so
(oracle/Makefile; sources stay in /root/reference).  Both run on the SAME weights: the reference
loads a model directory written by harness/modelgen.py and the oracle is handed a descriptor
filled from the reference's own ctx (adapter/qwen_tts_cuda_adapter.c).  Bit-exact.
Skipped where the prebuilt library is absent."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import FIXTURE_TOKENS
from harness import binding as B, modelgen as mg
from oracle import orc_binding as O

pytestmark = pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref/libqtts_ref.so not built (make -C oracle ref)")
CSV = ",".join(str(t) for t in FIXTURE_TOKENS).encode()


@pytest.fixture(scope="module")
def loaded(orc, tmp_path_factory):
    ref = O.load_ref()
    spec = mg.PRESETS["tiny"]
    main, cod = mg.build_tensors(spec, seed=3)
    d = str(tmp_path_factory.mktemp("model"))
    mg.write_model_dir(d, spec, main, cod)
    ctx = ref.ref_load(d.encode(), 0)
    assert ctx, "reference failed to load the generated model directory"
    desc = B.ModelDesc()
    ref.ref_fill_desc(ctx, C.byref(desc))
    yield ref, ctx, desc, spec
    ref.ref_free(ctx)


@pytest.mark.parametrize("params", [B.greedy_params(fixed_codec_tokens=10),
                                    B.default_params(fixed_codec_tokens=10),
                                    B.default_params(fixed_codec_tokens=6, sample_seed=1234, top_k=5, temperature=1.5)])
def test_generate_bit_exact(orc, loaded, params):
    ref, ctx, desc, spec = loaded
    ref.ref_set_params(ctx, C.byref(params))
    ns = C.c_int(0)
    want_pcm = B.take_floats(ref.ref_free_buf, ref.ref_generate(ctx, CSV, b"Aiden", b"ENGLISH", C.byref(ns)), ns.value)
    T = ref.ref_last_frames()
    want = np.zeros((T, 16), dtype=np.int32)
    ref.ref_last_codes(B.iptr(want))

    s = orc.orc_create(C.byref(desc))
    orc.orc_set_params(s, C.byref(params))
    tok = np.array(FIXTURE_TOKENS, dtype=np.int32)
    codes = np.zeros((T, 16), dtype=np.int32)
    nf = C.c_int(0)
    pcm = B.take_floats(orc.orc_free, orc.orc_generate(s, B.iptr(tok), len(tok), mg.SPEAKERS["aiden"], mg.LANGUAGES["english"],
                                                       C.byref(ns), B.iptr(codes), T, C.byref(nf)), ns.value)
    orc.orc_destroy(s)
    assert nf.value == T and np.array_equal(codes, want)
    assert np.array_equal(pcm.view(np.uint32), want_pcm.view(np.uint32))


def test_descriptor_from_reference_equals_harness_descriptor(orc, loaded):
    """the reference's load-time preprocessing (codebook normalisation, SnakeBeta exp) is reproduced by
    harness.binding.HostModel, so descriptors built without the reference carry the same numbers"""
    ref, ctx, desc, spec = loaded
    main, cod = mg.build_tensors(spec, seed=3)
    hm = B.HostModel(spec, main, cod)
    vq = spec["codec"]["cb_dim"] // 2
    for q in (0, 1, 15):
        a = np.ctypeslib.as_array(desc.codec.codebooks[q], shape=(2048 * vq,))
        b = np.ctypeslib.as_array(hm.desc.codec.codebooks[q], shape=(2048 * vq,))
        assert np.array_equal(a, b)
    n = spec["codec"]["dec"]
    for f in ("act_alpha", "act_beta"):
        a = np.ctypeslib.as_array(getattr(desc.codec.block[0], f), shape=(n,))
        b = np.ctypeslib.as_array(getattr(hm.desc.codec.block[0], f), shape=(n,))
        assert np.array_equal(a, b)
    assert desc.talker.hidden == hm.desc.talker.hidden and desc.sub.layers == hm.desc.sub.layers
    assert desc.codec_eos_id == hm.desc.codec_eos_id and desc.tts_pad_id == hm.desc.tts_pad_id


def test_text_embedding_bit_exact(orc, loaded):
    """embed_text_token (c/qwen_tts.c:823-847) is static in the reference; it is covered through
    qwen_tts_generate above.  Here: the oracle's prompt builder is deterministic and finite."""
    ref, ctx, desc, spec = loaded
    s = orc.orc_create(C.byref(desc))
    out = np.zeros(spec["talker"]["hidden"], dtype=np.float32)
    orc.orc_embed_text_token(s, 151671, B.fptr(out))
    assert np.isfinite(out).all() and np.abs(out).max() > 0
    orc.orc_destroy(s)
