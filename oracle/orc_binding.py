"""ctypes loaders of the CPU checker (TEST INFRASTRUCTURE ONLY):

  * oracle/liboracle.so        - the CPU restatement of the reference's algorithm
  * oracle/_ref/libqtts_ref.so - the unmodified reference + shim (may be absent on a box without /root/reference)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module;
nothing under qwen3-tts-c_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
from harness.binding import GenParams, ModelDesc, f32p, i32p, u16p  # noqa: E402  (struct mirrors of include/qwen_tts_cuda.h)

ORACLE_LIB = os.path.join(HERE, "liboracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libqtts_ref.so")
REF_CLI = os.path.join(HERE, "_ref", "qwen-tts")


def load_oracle():
    lib = C.CDLL(ORACLE_LIB)
    lib.orc_create.restype = C.c_void_p
    lib.orc_create.argtypes = [C.POINTER(ModelDesc)]
    for n in ("orc_destroy", "orc_reset"):
        getattr(lib, n).argtypes = [C.c_void_p]
        getattr(lib, n).restype = None
    lib.orc_set_params.argtypes = [C.c_void_p, C.POINTER(GenParams)]
    lib.orc_talker_prefill.argtypes = [C.c_void_p, f32p, C.c_int]
    lib.orc_talker_forward.argtypes = [C.c_void_p, f32p, f32p]
    lib.orc_subtalker_generate.argtypes = [C.c_void_p, f32p, C.c_int, i32p]
    lib.orc_codec_decode.restype = C.c_void_p
    lib.orc_codec_decode.argtypes = [C.c_void_p, i32p, C.c_int, i32p]
    lib.orc_hidden.restype = f32p
    lib.orc_hidden.argtypes = [C.c_void_p]
    lib.orc_kv_len.argtypes = [C.c_void_p]
    lib.orc_head_logits.argtypes = [C.c_void_p, f32p]
    lib.orc_embed_text_token.argtypes = [C.c_void_p, C.c_int, f32p]
    lib.orc_generate_frames.argtypes = [C.c_void_p, f32p, C.c_int, f32p, C.c_int, f32p, i32p, C.c_int, i32p, i32p]
    lib.orc_build_prompt.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p), i32p,
                                     C.POINTER(C.c_void_p), i32p, C.POINTER(C.c_void_p)]
    lib.orc_generate.restype = C.c_void_p
    lib.orc_generate.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_int, i32p, i32p, C.c_int, i32p]
    lib.orc_free.argtypes = [C.c_void_p]
    lib.orc_sample.argtypes = [f32p, C.c_int, C.c_int, C.c_float, C.c_float, f32p]
    lib.orc_matvec_bf16.argtypes = [f32p, u16p, f32p, C.c_int, C.c_int]
    lib.orc_swiglu_matvec_bf16.argtypes = [f32p, u16p, f32p, C.c_int, C.c_int]
    lib.orc_rms_norm.argtypes = [f32p, f32p, f32p, C.c_int, C.c_float]
    lib.orc_matmul_f32.argtypes = [f32p, f32p, f32p, C.c_int, C.c_int, C.c_int]
    lib.orc_matmul_bf16.argtypes = [f32p, f32p, u16p, C.c_int, C.c_int, C.c_int]
    lib.orc_snake_beta.argtypes = [f32p, f32p, f32p, f32p, C.c_int, C.c_int]
    lib.orc_causal_conv1d.argtypes = [f32p, f32p, f32p, f32p] + [C.c_int] * 6
    lib.orc_transposed_conv1d.argtypes = [f32p, f32p, f32p, f32p] + [C.c_int] * 5 + [i32p]
    lib.orc_prep_snake.argtypes = [f32p, f32p, C.c_int]
    lib.orc_prep_codebook.argtypes = [f32p, f32p, f32p, C.c_int, C.c_int]
    lib.orc_repetition_penalty.argtypes = [f32p, i32p, C.c_int, C.c_int, C.c_float]
    return lib


def have_ref() -> bool:
    return os.path.exists(REF_LIB)


def load_ref():
    lib = C.CDLL(REF_LIB)
    lib.ref_load.restype = C.c_void_p
    lib.ref_load.argtypes = [C.c_char_p, C.c_int]
    lib.ref_free.argtypes = [C.c_void_p]
    lib.ref_fill_desc.argtypes = [C.c_void_p, C.POINTER(ModelDesc)]
    lib.ref_set_params.argtypes = [C.c_void_p, C.POINTER(GenParams)]
    lib.ref_reset.argtypes = [C.c_void_p]
    lib.ref_prefill.argtypes = [C.c_void_p, f32p, C.c_int]
    lib.ref_forward.argtypes = [C.c_void_p, f32p, f32p]
    lib.ref_subtalker.argtypes = [C.c_void_p, f32p, C.c_int, i32p]
    lib.ref_codec_decode.restype = C.c_void_p
    lib.ref_codec_decode.argtypes = [C.c_void_p, i32p, C.c_int, i32p]
    lib.ref_hidden.restype = f32p
    lib.ref_hidden.argtypes = [C.c_void_p]
    lib.ref_kv_len.argtypes = [C.c_void_p]
    lib.ref_generate.restype = C.c_void_p
    lib.ref_generate.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, i32p]
    lib.ref_perf.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    lib.ref_free_buf.argtypes = [C.c_void_p]
    lib.ref_last_codes.argtypes = [i32p]
    lib.kernel_matvec_bf16.argtypes = [f32p, u16p, f32p, C.c_int, C.c_int]
    lib.kernel_sample_top_k.argtypes = [f32p, C.c_int, C.c_int, C.c_float, C.c_float, f32p]
    lib.kernel_causal_conv1d.argtypes = [f32p, f32p, f32p, f32p] + [C.c_int] * 6
    lib.kernel_transposed_conv1d.argtypes = [f32p, f32p, f32p, f32p] + [C.c_int] * 5 + [i32p]
    lib.kernel_snake_beta.argtypes = [f32p, f32p, f32p, f32p, C.c_int, C.c_int]
    return lib
