"""The C-ABI library loads (no GPU needed: cudart is linked statically) and exports every symbol that
include/qwen_tts_cuda.h declares.  No compute calls."""
import ctypes as C
import os
import re
import subprocess

from conftest import ROOT
from harness import binding as B


def _declared():
    src = open(os.path.join(ROOT, "include", "qwen_tts_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(qwen_tts_cuda_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(B.CUDA_LIB):
        subprocess.check_call(["make", "-j8", "-C", os.path.dirname(B.CUDA_LIB)])
    lib = C.CDLL(B.CUDA_LIB)
    names = _declared()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_no_device_means_loud_failure():
    """without a CUDA device init must fail (-1) and say why; nothing falls back to a CPU path"""
    import torch
    if torch.cuda.is_available():
        return
    lib = B.load_cuda()
    assert lib.qwen_tts_cuda_init(0) == -1
    assert lib.qwen_tts_cuda_is_available() == 0
    assert b"CUDA" in lib.qwen_tts_cuda_last_error()
    assert not lib.qwen_tts_cuda_upload(None, 1)


def test_struct_layout_matches_header():
    """ctypes mirrors vs sizeof() from the C compiler"""
    code = r'''
#include <stdio.h>
#include "qwen_tts_cuda.h"
int main(void){ printf("%zu %zu %zu %zu %zu\n", sizeof(qtc_lm_t), sizeof(qtc_codec_t), sizeof(qtc_model_desc_t), sizeof(qtc_gen_params_t), sizeof(qtc_perf_t)); return 0; }
'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(code)
        subprocess.check_call(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "s"), os.path.join(d, "s.c")])
        got = [int(x) for x in subprocess.check_output([os.path.join(d, "s")]).split()]
    want = [C.sizeof(B.Lm), C.sizeof(B.Codec), C.sizeof(B.ModelDesc), C.sizeof(B.GenParams), C.sizeof(B.Perf)]
    assert got == want


def test_binding_arity_matches_header():
    """every function the ctypes harness declares argtypes for takes as many arguments as its prototype in the header (the queue
    entry point gained one in the course of round 2: a mirror that lags would pass garbage pointers)"""
    src = open(os.path.join(ROOT, "include", "qwen_tts_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(qwen_tts_cuda_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        depth, n = 0, 0 if args in ("", "void") else 1
        for ch in args:
            depth += ch == "("
            depth -= ch == ")"
            n += ch == "," and depth == 0
        protos[m.group(1)] = n
    lib = B.load_cuda()
    checked = 0
    for name, n in protos.items():
        fn = getattr(lib, name)
        if fn.argtypes is not None:
            assert len(fn.argtypes) == n, (name, len(fn.argtypes), n)
            checked += 1
    assert checked >= 20, checked
