"""ctypes mirror of include/qwen_tts_cuda.h and the loader of the product library
(libqwen_tts_cuda.so, qwen3-tts-c_b200/csrc).  The loaders of the CPU checker live with the checker
(oracle/orc_binding.py) and are imported only by tests/, __graft_entry__.smoke() and bench.py's CPU legs.

Python is harness only: no arithmetic of the hot path happens here.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CUDA_LIB = os.environ.get("QTC_LIB") or os.path.join(ROOT, "qwen3-tts-c_b200", "csrc", "libqwen_tts_cuda.so")   # QTC_LIB: A/B against an older build

MAX_LM_LAYERS, MAX_CODEC_LAYERS, MAX_GROUPS, MAX_Q, MAX_NAMES = 32, 12, 32, 16, 64
u16p, f32p, i32p = C.POINTER(C.c_uint16), C.POINTER(C.c_float), C.POINTER(C.c_int)


class LmLayer(C.Structure):
    _fields_ = [("wq", u16p), ("wk", u16p), ("wv", u16p), ("wo", u16p),
                ("q_norm", f32p), ("k_norm", f32p), ("input_norm", f32p), ("post_attn_norm", f32p),
                ("gate", u16p), ("up", u16p), ("down", u16p)]


class Lm(C.Structure):
    _fields_ = [("vocab", C.c_int), ("hidden", C.c_int), ("intermediate", C.c_int), ("layers", C.c_int),
                ("heads", C.c_int), ("kv_heads", C.c_int), ("head_dim", C.c_int),
                ("rms_eps", C.c_float), ("rope_theta", C.c_float),
                ("layer", LmLayer * MAX_LM_LAYERS), ("final_norm", f32p)]


class CodecLayer(C.Structure):
    _fields_ = [(n, f32p) for n in ("input_norm", "post_attn_norm", "attn_layer_scale", "mlp_layer_scale",
                                    "wq", "wk", "wv", "wo", "gate", "up", "down")]


class ConvNext(C.Structure):
    _fields_ = [(n, f32p) for n in ("dwconv_w", "dwconv_b", "norm_w", "norm_b", "pw1_w", "pw1_b",
                                    "pw2_w", "pw2_b", "gamma")]


class ResUnit(C.Structure):
    _fields_ = [(n, f32p) for n in ("act1_alpha", "act1_beta", "conv1_w", "conv1_b",
                                    "act2_alpha", "act2_beta", "conv2_w", "conv2_b")]


class VocBlock(C.Structure):
    _fields_ = [("act_alpha", f32p), ("act_beta", f32p), ("tconv_w", f32p), ("tconv_b", f32p),
                ("res", ResUnit * 3)]


class Codec(C.Structure):
    _fields_ = [("num_quantizers", C.c_int), ("codebook_size", C.c_int), ("codebook_dim", C.c_int),
                ("hidden", C.c_int), ("latent", C.c_int), ("layers", C.c_int), ("heads", C.c_int),
                ("kv_heads", C.c_int), ("intermediate", C.c_int), ("sliding_window", C.c_int),
                ("decoder_dim", C.c_int), ("rms_eps", C.c_float),
                ("upsample_rates", C.c_int * 4), ("upsampling_ratios", C.c_int * 2),
                ("codebooks", f32p * MAX_Q), ("semantic_out_proj", f32p), ("acoustic_out_proj", f32p),
                ("pre_conv_w", f32p), ("pre_conv_b", f32p), ("in_proj_w", f32p), ("in_proj_b", f32p),
                ("out_proj_w", f32p), ("out_proj_b", f32p), ("layer", CodecLayer * MAX_CODEC_LAYERS),
                ("final_norm", f32p), ("up_tconv_w", f32p * 2), ("up_tconv_b", f32p * 2),
                ("convnext", ConvNext * 2), ("voc_pre_w", f32p), ("voc_pre_b", f32p),
                ("block", VocBlock * 4), ("final_alpha", f32p), ("final_beta", f32p),
                ("final_conv_w", f32p), ("final_conv_b", f32p)]


class ModelDesc(C.Structure):
    _fields_ = [("talker", Lm), ("text_hidden", C.c_int), ("text_vocab", C.c_int), ("num_code_groups", C.c_int),
                ("mrope_section", C.c_int * 3),
                ("codec_embedding", u16p), ("text_embedding", u16p), ("text_fc1_w", u16p), ("text_fc1_b", f32p),
                ("text_fc2_w", u16p), ("text_fc2_b", f32p), ("codec_head", u16p),
                ("sub", Lm), ("sub_codec_embeddings", u16p * (MAX_GROUPS - 1)),
                ("sub_input_proj_w", u16p), ("sub_input_proj_b", f32p), ("sub_lm_heads", u16p * (MAX_GROUPS - 1)),
                ("codec", Codec),
                ("codec_pad_id", C.c_int), ("codec_bos_id", C.c_int), ("codec_eos_id", C.c_int),
                ("codec_nothink_id", C.c_int), ("codec_think_id", C.c_int), ("codec_think_bos_id", C.c_int),
                ("codec_think_eos_id", C.c_int), ("tts_pad_id", C.c_int), ("tts_bos_id", C.c_int),
                ("tts_eos_id", C.c_int),
                ("n_speakers", C.c_int), ("speaker_names", C.c_char_p * MAX_NAMES), ("speaker_ids", C.c_int * MAX_NAMES),
                ("n_languages", C.c_int), ("language_names", C.c_char_p * MAX_NAMES), ("language_ids", C.c_int * MAX_NAMES)]


class GenParams(C.Structure):
    _fields_ = [("temperature", C.c_float), ("subtalker_temperature", C.c_float),
                ("top_k", C.c_int), ("subtalker_top_k", C.c_int),
                ("top_p", C.c_float), ("subtalker_top_p", C.c_float), ("repetition_penalty", C.c_float),
                ("max_new_tokens", C.c_int), ("fixed_codec_tokens", C.c_int), ("sample_seed", C.c_int)]


class Perf(C.Structure):
    _fields_ = [("total_ms", C.c_double), ("talker_ms", C.c_double), ("codec_ms", C.c_double),
                ("prefill_ms", C.c_double), ("codec_tokens", C.c_int), ("stop_reason", C.c_int),
                ("stop_step", C.c_int), ("kernel_launches", C.c_longlong),
                ("device_ms", C.c_double), ("ttff_ms", C.c_double), ("ttfa_ms", C.c_double),
                ("frame_ms_p50", C.c_double), ("frame_ms_p99", C.c_double),
                ("device_mallocs", C.c_longlong), ("kv_page_allocs", C.c_longlong),
                ("kv_pool_pages", C.c_int), ("kv_pages_in_use", C.c_int),
                ("upload_ms", C.c_double), ("upload_bytes", C.c_double), ("queue_admitted", C.c_int)]


PcmCallback = C.CFUNCTYPE(None, C.POINTER(C.c_float), C.c_int, C.c_int, C.c_void_p)


def default_params(**kw) -> GenParams:
    """c/qwen_tts.c:871-880"""
    p = GenParams(0.9, 0.9, 50, 50, 1.0, 1.0, 1.05, 4096, 0, 42)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def greedy_params(**kw) -> GenParams:
    """the reference's deterministic flag set (test/test_eos_regression.py:62-68)"""
    p = GenParams(1.0, 1.0, 1, 1, 1.0, 1.0, 1.0, 4096, 0, 42)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _p(a, typ):
    return None if a is None else a.ctypes.data_as(typ)


def load_cuda():
    """The product library.  Fails loudly if it was not built: there is no fallback."""
    if not os.path.exists(CUDA_LIB):
        raise RuntimeError(f"{CUDA_LIB} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(CUDA_LIB)
    vp = C.c_void_p
    lib.qwen_tts_cuda_init.argtypes = [C.c_int]
    lib.qwen_tts_cuda_upload.restype = vp
    lib.qwen_tts_cuda_upload.argtypes = [C.POINTER(ModelDesc), C.c_int]
    lib.qwen_tts_cuda_release.argtypes = [vp]
    lib.qwen_tts_cuda_set_params.argtypes = [vp, C.POINTER(GenParams)]
    lib.qwen_tts_cuda_get_perf.argtypes = [vp, C.POINTER(Perf)]
    lib.qwen_tts_cuda_set_mode.argtypes = [vp, C.c_int]
    lib.qwen_tts_cuda_set_arith.argtypes = [vp, C.c_int]
    lib.qwen_tts_cuda_kernel_arith.argtypes = [C.c_int]
    lib.qwen_tts_cuda_reset.argtypes = [vp]
    lib.qwen_tts_cuda_talker_prefill.argtypes = [vp, f32p, C.c_int]
    lib.qwen_tts_cuda_talker_forward.argtypes = [vp, f32p, f32p]
    lib.qwen_tts_cuda_subtalker_generate.argtypes = [vp, f32p, C.c_int, i32p]
    lib.qwen_tts_cuda_codec_decode.restype = vp
    lib.qwen_tts_cuda_codec_decode.argtypes = [vp, i32p, C.c_int, i32p]
    lib.qwen_tts_cuda_get_hidden.argtypes = [vp, f32p]
    lib.qwen_tts_cuda_get_kv_len.argtypes = [vp]
    lib.qwen_tts_cuda_head_logits.argtypes = [vp, f32p]
    lib.qwen_tts_cuda_embed_text_tokens.argtypes = [vp, i32p, C.c_int, f32p]
    lib.qwen_tts_cuda_generate_frames.argtypes = [vp, f32p, C.c_int, f32p, C.c_int, f32p, i32p, C.c_int]
    lib.qwen_tts_cuda_generate.restype = vp
    lib.qwen_tts_cuda_generate.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_char_p, i32p, i32p, C.c_int, i32p]
    lib.qwen_tts_cuda_set_streaming.argtypes = [vp, C.c_int, C.c_int]
    lib.qwen_tts_cuda_set_codec_tc.argtypes = [vp, C.c_int]
    lib.qwen_tts_cuda_set_stream_callback.argtypes = [vp, PcmCallback, vp]
    lib.qwen_tts_cuda_set_stream_callback.restype = None
    lib.qwen_tts_cuda_codec_stream_begin.argtypes = [vp, C.c_int]
    lib.qwen_tts_cuda_codec_stream_push.argtypes = [vp, i32p, C.c_int, f32p]
    lib.qwen_tts_cuda_generate_batch.argtypes = [vp, C.c_int, i32p, i32p, i32p, i32p, i32p, i32p, f32p, C.c_int]
    lib.qwen_tts_cuda_generate_queue.argtypes = [vp, C.c_int, C.c_int, i32p, i32p, i32p, i32p, i32p, i32p, i32p, f32p, C.c_int]
    lib.qwen_tts_cuda_matvec_bf16.argtypes = [f32p, u16p, f32p, C.c_int, C.c_int]
    lib.qwen_tts_cuda_swiglu_matvec_bf16.argtypes = [f32p, u16p, f32p, C.c_int, C.c_int]
    lib.qwen_tts_cuda_rms_norm.argtypes = [f32p, f32p, f32p, C.c_int, C.c_float]
    lib.qwen_tts_cuda_sample_top_k.argtypes = [f32p, C.c_int, C.c_int, C.c_float, C.c_float, f32p]
    lib.qwen_tts_cuda_sample_top_k_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.qwen_tts_cuda_last_pcm_s16.argtypes = [vp, C.c_void_p, C.c_int]
    lib.qwen_tts_cuda_causal_conv1d.argtypes = [f32p, f32p, f32p, f32p] + [C.c_int] * 6
    lib.qwen_tts_cuda_transposed_conv1d.argtypes = [f32p, f32p, f32p, f32p] + [C.c_int] * 5 + [i32p]
    lib.qwen_tts_cuda_snake_beta.argtypes = [f32p, f32p, f32p, f32p, C.c_int, C.c_int]
    lib.qwen_tts_cuda_matmul_f32.argtypes = [f32p, f32p, f32p, C.c_int, C.c_int, C.c_int]
    lib.qwen_tts_cuda_matmul_bf16.argtypes = [f32p, f32p, u16p, C.c_int, C.c_int, C.c_int]
    lib.qwen_tts_cuda_bench_talker_step.restype = C.c_double
    lib.qwen_tts_cuda_bench_talker_step.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    lib.qwen_tts_cuda_bench_subtalker.restype = C.c_double
    lib.qwen_tts_cuda_bench_subtalker.argtypes = [vp, C.c_int, C.c_int]
    lib.qwen_tts_cuda_talker_step_bytes.restype = C.c_double
    lib.qwen_tts_cuda_talker_step_bytes.argtypes = [vp, C.c_int]
    lib.qwen_tts_cuda_subtalker_frame_bytes.restype = C.c_double
    lib.qwen_tts_cuda_subtalker_frame_bytes.argtypes = [vp]
    lib.qwen_tts_cuda_last_error.restype = C.c_char_p
    lib.qwen_tts_cuda_free_host = C.CDLL(None).free
    lib.qwen_tts_cuda_free_host.argtypes = [vp]
    return lib


_LIBM = None


def _expf(a: np.ndarray) -> np.ndarray:
    """glibc expf, element by element: numpy's float32 exp is a different implementation and may differ in the last bit
    from what the reference's loader computes (c/qwen_tts.c:596-602).  The SnakeBeta vectors are a few thousand elements."""
    global _LIBM
    if _LIBM is None:
        _LIBM = C.CDLL("libm.so.6")
        _LIBM.expf.restype = C.c_float
        _LIBM.expf.argtypes = [C.c_float]
    return np.array([_LIBM.expf(float(v)) for v in a.ravel()], dtype=np.float32).reshape(a.shape)


def prep_codebook(embedding_sum: np.ndarray, usage: np.ndarray) -> np.ndarray:
    """c/qwen_tts.c:577-594: E[c] = embedding_sum[c] * (1.0f / max(usage[c], 1e-5f)); IEEE fp32 ops, so numpy is bit-identical"""
    inv = np.float32(1.0) / np.maximum(usage.astype(np.float32), np.float32(1e-5))
    return np.ascontiguousarray(embedding_sum.astype(np.float32) * inv[:, None], dtype=np.float32)


def prep_snake(alpha: np.ndarray, beta: np.ndarray):
    """c/qwen_tts.c:596-602: a = expf(alpha), inv_b = 1.0f / (expf(beta) + 1e-9f)"""
    a = _expf(alpha.astype(np.float32))
    b = (np.float32(1.0) / (_expf(beta.astype(np.float32)) + np.float32(1e-9))).astype(np.float32)
    return np.ascontiguousarray(a), np.ascontiguousarray(b)


class HostModel:
    """Tensors of modelgen.build_tensors bound into a ModelDesc (keeps the arrays alive).
    Applies the reference's load-time preprocessing (c/qwen_tts.c:577-602) in numpy / libm, so values are bit-identical
    to a model loaded by the reference; the checker (oracle/) is not involved (tests/test_oracle_golden.py holds the
    two to each other)."""

    def __init__(self, spec, main, cod, oracle_lib=None):
        from . import modelgen as mg
        self.spec, self.keep = spec, []
        d = ModelDesc()

        def b16(name):
            a = main.get(name)
            return None if a is None else a.ctypes.data_as(u16p)

        def f32(t, name):
            a = t.get(name)
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=np.float32)
            self.keep.append(a)
            return a.ctypes.data_as(f32p)

        def fill_lm(lm, L, prefix):
            lm.vocab, lm.hidden, lm.intermediate, lm.layers = L["vocab"], L["hidden"], L["inter"], L["layers"]
            lm.heads, lm.kv_heads, lm.head_dim = L["heads"], L["kv"], L["hd"]
            lm.rms_eps, lm.rope_theta = 1e-6, spec["theta"]
            for i in range(L["layers"]):
                p, l = f"{prefix}.layers.{i}.", lm.layer[i]
                l.wq, l.wk = b16(p + "self_attn.q_proj.weight"), b16(p + "self_attn.k_proj.weight")
                l.wv, l.wo = b16(p + "self_attn.v_proj.weight"), b16(p + "self_attn.o_proj.weight")
                l.q_norm, l.k_norm = f32(main, p + "self_attn.q_norm.weight"), f32(main, p + "self_attn.k_norm.weight")
                l.input_norm = f32(main, p + "input_layernorm.weight")
                l.post_attn_norm = f32(main, p + "post_attention_layernorm.weight")
                l.gate, l.up, l.down = (b16(p + "mlp.gate_proj.weight"), b16(p + "mlp.up_proj.weight"),
                                        b16(p + "mlp.down_proj.weight"))
            lm.final_norm = f32(main, f"{prefix}.norm.weight")

        if main:
            fill_lm(d.talker, spec["talker"], "talker.model")
            fill_lm(d.sub, spec["sub"], "talker.code_predictor.model")
            d.codec_embedding = b16("talker.model.codec_embedding.weight")
            d.text_embedding = b16("talker.model.text_embedding.weight")
            d.text_fc1_w, d.text_fc1_b = b16("talker.text_projection.linear_fc1.weight"), f32(main, "talker.text_projection.linear_fc1.bias")
            d.text_fc2_w, d.text_fc2_b = b16("talker.text_projection.linear_fc2.weight"), f32(main, "talker.text_projection.linear_fc2.bias")
            d.codec_head = b16("talker.codec_head.weight")
            for g in range(spec["groups"] - 1):
                d.sub_codec_embeddings[g] = b16(f"talker.code_predictor.model.codec_embedding.{g}.weight")
                d.sub_lm_heads[g] = b16(f"talker.code_predictor.lm_head.{g}.weight")
            d.sub_input_proj_w = b16("talker.code_predictor.small_to_mtp_projection.weight")
            d.sub_input_proj_b = f32(main, "talker.code_predictor.small_to_mtp_projection.bias")
        else:   # codec-only model: dims still needed by upload()
            for lm, L in ((d.talker, spec["talker"]), (d.sub, spec["sub"])):
                lm.vocab, lm.hidden, lm.intermediate, lm.layers = L["vocab"], L["hidden"], L["inter"], 0
                lm.heads, lm.kv_heads, lm.head_dim, lm.rms_eps, lm.rope_theta = L["heads"], L["kv"], L["hd"], 1e-6, spec["theta"]
        d.text_hidden, d.text_vocab, d.num_code_groups = spec["text_hidden"], spec["text_vocab"], spec["groups"]
        for i in range(3):
            d.mrope_section[i] = spec["mrope"][i]

        Cc, k = spec["codec"], d.codec
        k.num_quantizers, k.codebook_size, k.codebook_dim = Cc["nq"], Cc["cb_size"], Cc["cb_dim"]
        k.hidden, k.latent, k.layers, k.heads, k.kv_heads = Cc["hidden"], Cc["latent"], Cc["layers"], Cc["heads"], Cc["kv"]
        k.intermediate, k.sliding_window, k.decoder_dim, k.rms_eps = Cc["inter"], Cc["window"], Cc["dec"], 1e-5
        for i in range(4):
            k.upsample_rates[i] = Cc["rates"][i]
        for i in range(2):
            k.upsampling_ratios[i] = Cc["ratios"][i]
        if cod:
            vq = Cc["cb_dim"] // 2
            for q in range(Cc["nq"]):
                p = ("decoder.quantizer.rvq_first.vq.layers.0._codebook." if q == 0
                     else f"decoder.quantizer.rvq_rest.vq.layers.{q - 1}._codebook.")
                emb = prep_codebook(cod[p + "embedding_sum"].reshape(Cc["cb_size"], vq), cod[p + "cluster_usage"])
                self.keep.append(emb)
                k.codebooks[q] = emb.ctypes.data_as(f32p)

            def snake(a_name, b_name):
                a, b = prep_snake(cod[a_name], cod[b_name])
                self.keep += [a, b]
                return a.ctypes.data_as(f32p), b.ctypes.data_as(f32p)

            k.semantic_out_proj = f32(cod, "decoder.quantizer.rvq_first.output_proj.weight")
            k.acoustic_out_proj = f32(cod, "decoder.quantizer.rvq_rest.output_proj.weight")
            k.pre_conv_w, k.pre_conv_b = f32(cod, "decoder.pre_conv.conv.weight"), f32(cod, "decoder.pre_conv.conv.bias")
            pt = "decoder.pre_transformer."
            k.in_proj_w, k.in_proj_b = f32(cod, pt + "input_proj.weight"), f32(cod, pt + "input_proj.bias")
            k.out_proj_w, k.out_proj_b = f32(cod, pt + "output_proj.weight"), f32(cod, pt + "output_proj.bias")
            k.final_norm = f32(cod, pt + "norm.weight")
            for i in range(Cc["layers"]):
                p, l = f"{pt}layers.{i}.", k.layer[i]
                l.input_norm, l.post_attn_norm = f32(cod, p + "input_layernorm.weight"), f32(cod, p + "post_attention_layernorm.weight")
                l.attn_layer_scale, l.mlp_layer_scale = f32(cod, p + "self_attn_layer_scale.scale"), f32(cod, p + "mlp_layer_scale.scale")
                l.wq, l.wk = f32(cod, p + "self_attn.q_proj.weight"), f32(cod, p + "self_attn.k_proj.weight")
                l.wv, l.wo = f32(cod, p + "self_attn.v_proj.weight"), f32(cod, p + "self_attn.o_proj.weight")
                l.gate, l.up, l.down = f32(cod, p + "mlp.gate_proj.weight"), f32(cod, p + "mlp.up_proj.weight"), f32(cod, p + "mlp.down_proj.weight")
            for s in range(2):
                k.up_tconv_w[s], k.up_tconv_b[s] = f32(cod, f"decoder.upsample.{s}.0.conv.weight"), f32(cod, f"decoder.upsample.{s}.0.conv.bias")
                p, c = f"decoder.upsample.{s}.1.", k.convnext[s]
                c.dwconv_w, c.dwconv_b = f32(cod, p + "dwconv.conv.weight"), f32(cod, p + "dwconv.conv.bias")
                c.norm_w, c.norm_b = f32(cod, p + "norm.weight"), f32(cod, p + "norm.bias")
                c.pw1_w, c.pw1_b = f32(cod, p + "pwconv1.weight"), f32(cod, p + "pwconv1.bias")
                c.pw2_w, c.pw2_b, c.gamma = f32(cod, p + "pwconv2.weight"), f32(cod, p + "pwconv2.bias"), f32(cod, p + "gamma")
            k.voc_pre_w, k.voc_pre_b = f32(cod, "decoder.decoder.0.conv.weight"), f32(cod, "decoder.decoder.0.conv.bias")
            for b in range(4):
                p, vb = f"decoder.decoder.{b + 1}.block.", k.block[b]
                vb.act_alpha, vb.act_beta = snake(p + "0.alpha", p + "0.beta")
                vb.tconv_w, vb.tconv_b = f32(cod, p + "1.conv.weight"), f32(cod, p + "1.conv.bias")
                for u in range(3):
                    q, r = f"{p}{u + 2}.", vb.res[u]
                    r.act1_alpha, r.act1_beta = snake(q + "act1.alpha", q + "act1.beta")
                    r.conv1_w, r.conv1_b = f32(cod, q + "conv1.conv.weight"), f32(cod, q + "conv1.conv.bias")
                    r.act2_alpha, r.act2_beta = snake(q + "act2.alpha", q + "act2.beta")
                    r.conv2_w, r.conv2_b = f32(cod, q + "conv2.conv.weight"), f32(cod, q + "conv2.conv.bias")
            k.final_alpha, k.final_beta = snake("decoder.decoder.5.alpha", "decoder.decoder.5.beta")
            k.final_conv_w, k.final_conv_b = f32(cod, "decoder.decoder.6.conv.weight"), f32(cod, "decoder.decoder.6.conv.bias")

        ids = mg.CODEC_IDS
        d.codec_pad_id, d.codec_bos_id, d.codec_eos_id = ids["codec_pad_id"], ids["codec_bos_id"], ids["codec_eos_token_id"]
        d.codec_nothink_id, d.codec_think_id = ids["codec_nothink_id"], ids["codec_think_id"]
        d.codec_think_bos_id, d.codec_think_eos_id = ids["codec_think_bos_id"], ids["codec_think_eos_id"]
        d.tts_pad_id, d.tts_bos_id, d.tts_eos_id = mg.TTS_PAD, mg.TTS_BOS, mg.TTS_EOS
        d.n_speakers = len(mg.SPEAKERS)
        for i, (n, v) in enumerate(mg.SPEAKERS.items()):
            d.speaker_names[i], d.speaker_ids[i] = n.encode(), v
        d.n_languages = len(mg.LANGUAGES)
        for i, (n, v) in enumerate(mg.LANGUAGES.items()):
            d.language_names[i], d.language_ids[i] = n.encode(), v
        self.desc, self.main, self.cod = d, main, cod


def fptr(a):
    return a.ctypes.data_as(f32p)


def iptr(a):
    return a.ctypes.data_as(i32p)


def uptr(a):
    return a.ctypes.data_as(u16p)


def take_floats(lib_free, ptr, n):
    """copy a malloc'd float buffer returned by a C call and free it"""
    if not ptr:
        return None
    out = np.ctypeslib.as_array(C.cast(ptr, f32p), shape=(n,)).copy()
    lib_free(ptr)
    return out
