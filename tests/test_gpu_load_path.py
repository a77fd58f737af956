"""Load path: weights go up through pinned staging buffers (time and bytes in qtc_perf_t), and the 16-bit
PCM the reference's WAV writer produces on the host (c/qwen_tts_audio.c:70-82) comes off the device already converted."""
import ctypes as C

import numpy as np
import pytest

from harness import binding as B
from test_gpu_parity import CSV

pytestmark = pytest.mark.gpu


def test_upload_reports_time_and_bytes(cuda, tiny):
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    perf = B.Perf()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    main_bytes = sum(a.nbytes for a in tiny.keep if a.dtype == np.uint16)      # the LM's bf16 tensors are copied as they are, once
    assert perf.upload_ms > 0.0
    assert perf.upload_bytes >= main_bytes, (perf.upload_bytes, main_bytes)
    cuda.qwen_tts_cuda_release(ctx)


def test_s16_pcm_from_device_equals_host_conversion(cuda, tiny):
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    p = B.default_params(fixed_codec_tokens=12)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
    ns, nf = C.c_int(0), C.c_int(0)
    codes = np.zeros((12, 16), dtype=np.int32)
    ptr = cuda.qwen_tts_cuda_generate(ctx, CSV, b"aiden", b"english", C.byref(ns), B.iptr(codes), 12, C.byref(nf))
    assert ptr, cuda.qwen_tts_cuda_last_error()
    pcm = B.take_floats(cuda.qwen_tts_cuda_free_host, ptr, ns.value)
    got = np.zeros(ns.value, dtype=np.int16)
    assert cuda.qwen_tts_cuda_last_pcm_s16(ctx, got.ctypes.data, ns.value) == ns.value, cuda.qwen_tts_cuda_last_error()
    want = np.trunc(np.clip(pcm, np.float32(-1.0), np.float32(1.0)) * np.float32(32767.0)).astype(np.int16)   # fp32 product, then (int16_t)
    assert np.array_equal(got, want)
    assert np.abs(got).max() > 0
    # a shorter buffer takes a prefix; a streamed decode keeps no device copy and says so
    part = np.zeros(1000, dtype=np.int16)
    assert cuda.qwen_tts_cuda_last_pcm_s16(ctx, part.ctypes.data, 1000) == 1000 and np.array_equal(part, want[:1000])
    cuda.qwen_tts_cuda_set_streaming(ctx, 4, 4)
    ptr = cuda.qwen_tts_cuda_generate(ctx, CSV, b"aiden", b"english", C.byref(ns), B.iptr(codes), 12, C.byref(nf))
    assert ptr
    B.take_floats(cuda.qwen_tts_cuda_free_host, ptr, ns.value)
    assert cuda.qwen_tts_cuda_last_pcm_s16(ctx, got.ctypes.data, ns.value) == -1
    assert b"streamed" in cuda.qwen_tts_cuda_last_error()
    cuda.qwen_tts_cuda_release(ctx)
