"""Streaming / chunked codec decode: every layer of the decoder is causal, so a chunked decode that
carries the transformer's K/V by absolute position and re-runs the conv stack over its (< 10 frame) receptive field is EXACT.
  * stage level: any chunking of a code sequence gives the one-shot PCM bit for bit (tiny codec: window 8; full-size codec:
    window 72, 1536-channel vocoder), and the one-shot PCM is the oracle's within 80 dB;
  * whole utterance: qwen_tts_cuda_generate with the decode overlapped on a second stream returns the codes and PCM of the
    decode-at-the-end order, reports time-to-first-audio, and the callback sees every sample once, in order;
  * EOS mode (no fixed budget): only frames the device confirmed are decoded; same PCM."""
import ctypes as C

import numpy as np
import pytest

from conftest import get_model
from harness import binding as B
from test_gpu_parity import CSV, bits, snr_db

pytestmark = pytest.mark.gpu


def _oneshot(cuda, ctx, codes):
    ns = C.c_int(0)
    ptr = cuda.qwen_tts_cuda_codec_decode(ctx, B.iptr(codes), len(codes), C.byref(ns))
    assert ptr and ns.value == len(codes) * 1920, cuda.qwen_tts_cuda_last_error()
    return B.take_floats(cuda.qwen_tts_cuda_free_host, ptr, ns.value)


def _streamed(cuda, ctx, codes, chunks):
    assert cuda.qwen_tts_cuda_codec_stream_begin(ctx, len(codes)) == 0, cuda.qwen_tts_cuda_last_error()
    out, t = [], 0
    i = 0
    while t < len(codes):
        n = min(chunks[i % len(chunks)], len(codes) - t)
        buf = np.zeros(n * 1920, dtype=np.float32)
        part = np.ascontiguousarray(codes[t:t + n])
        assert cuda.qwen_tts_cuda_codec_stream_push(ctx, B.iptr(part), n, B.fptr(buf)) == n * 1920, cuda.qwen_tts_cuda_last_error()
        out.append(buf)
        t += n
        i += 1
    return np.concatenate(out)


@pytest.mark.parametrize("preset,T,chunks", [("tiny", 100, [1, 7, 30, 2, 13]), ("tiny", 40, [40]), ("1.7b", 120, [8, 40, 1, 33]), ("1.7b", 30, [3])],
                         ids=["tiny-ragged", "tiny-one", "full-ragged", "full-small"])
def test_codec_stream_equals_oneshot(cuda, orc, preset, T, chunks):
    hm = get_model(orc, preset)
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    rng = np.random.default_rng(T)
    codes = rng.integers(0, 2048, size=(T, 16), dtype=np.int32)
    codes[3, 2] = -5                                    # out-of-range ids dequantise to zero in every chunking
    full = _oneshot(cuda, ctx, codes)
    got = _streamed(cuda, ctx, codes, chunks)
    assert np.array_equal(bits(got), bits(full)), f"first differing sample {np.argwhere(bits(got) != bits(full))[0]}"
    # a session may be cut short and a new one started
    again = _streamed(cuda, ctx, codes[: T // 2], [5])
    assert np.array_equal(bits(again), bits(full[: (T // 2) * 1920]))
    if preset == "tiny":
        s = orc.orc_create(hm.desc)
        ns = C.c_int(0)
        want = B.take_floats(orc.orc_free, orc.orc_codec_decode(s, B.iptr(codes), T, C.byref(ns)), T * 1920)
        orc.orc_destroy(s)
        assert snr_db(want, full) >= 80.0
    # pushing past the session's frame count is an error, not a crash
    extra = np.zeros((T, 16), dtype=np.int32)
    buf = np.zeros(T * 1920, dtype=np.float32)
    assert cuda.qwen_tts_cuda_codec_stream_push(ctx, B.iptr(extra), T, B.fptr(buf)) == -1
    cuda.qwen_tts_cuda_release(ctx)


def _gen(cuda, ctx, params, n):
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(params))
    ns, nf = C.c_int(0), C.c_int(0)
    codes = np.zeros((n, 16), dtype=np.int32)
    ptr = cuda.qwen_tts_cuda_generate(ctx, CSV, b"aiden", b"english", C.byref(ns), B.iptr(codes), n, C.byref(nf))
    assert ptr, cuda.qwen_tts_cuda_last_error()
    return codes[: nf.value], B.take_floats(cuda.qwen_tts_cuda_free_host, ptr, ns.value)


@pytest.mark.parametrize("preset,frames", [("tiny", 60), ("1.7b", 100)])
def test_generate_streamed_equals_decode_at_end(cuda, orc, preset, frames):
    hm = get_model(orc, preset)
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    p = B.default_params(fixed_codec_tokens=frames)
    assert cuda.qwen_tts_cuda_set_streaming(ctx, 0, 0) == 0
    codes0, pcm0 = _gen(cuda, ctx, p, frames)
    perf = B.Perf()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    assert perf.ttfa_ms == 0.0 and perf.ttff_ms > 0.0
    seen = []

    @B.PcmCallback
    def cb(pcm, n, first_frame, user):
        seen.append((first_frame, np.ctypeslib.as_array(pcm, shape=(n,)).copy()))

    cuda.qwen_tts_cuda_set_stream_callback(ctx, cb, None)
    assert cuda.qwen_tts_cuda_set_streaming(ctx, 8, 24) == 0
    codes1, pcm1 = _gen(cuda, ctx, p, frames)
    cuda.qwen_tts_cuda_set_stream_callback(ctx, B.PcmCallback(), None)
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    assert np.array_equal(codes0, codes1)
    assert pcm1.size == frames * 1920 and np.array_equal(bits(pcm0), bits(pcm1))
    assert 0.0 < perf.ttff_ms < perf.ttfa_ms < perf.device_ms, (perf.ttff_ms, perf.ttfa_ms, perf.device_ms)
    assert 0.0 < perf.frame_ms_p50 <= perf.frame_ms_p99
    assert [f for f, _ in seen] == [0] + list(range(8, frames, 24)), [f for f, _ in seen]
    assert np.array_equal(bits(np.concatenate([a for _, a in seen])), bits(pcm1))
    cuda.qwen_tts_cuda_release(ctx)


def test_generate_streamed_eos_mode(cuda, orc, tiny):
    """no fixed budget: the host polls the stop flag every 8 frames and decodes confirmed frames only"""
    ctx = cuda.qwen_tts_cuda_upload(C.byref(tiny.desc), 1)
    assert ctx
    for p in (B.default_params(max_new_tokens=40, temperature=3.0, top_k=3000, sample_seed=11), B.default_params(max_new_tokens=21)):
        cuda.qwen_tts_cuda_set_streaming(ctx, 0, 0)
        codes0, pcm0 = _gen(cuda, ctx, p, 40)
        cuda.qwen_tts_cuda_set_streaming(ctx, 4, 8)
        codes1, pcm1 = _gen(cuda, ctx, p, 40)
        assert np.array_equal(codes0, codes1) and np.array_equal(bits(pcm0), bits(pcm1))
        assert pcm1.size == len(codes1) * 1920
    cuda.qwen_tts_cuda_release(ctx)
