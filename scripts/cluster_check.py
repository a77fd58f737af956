"""Mode 2 (cluster kernel) against mode 1 (148-CTA fused kernel) on the same model: logits, sub-talker codes, run-to-run
bit equality, and the step timings of both.  Usage (GPU box): python scripts/cluster_check.py [preset] [timing preset]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
from harness import binding as B, modelgen as mg  # noqa: E402


def status(cuda, ctx):
    cap, err = C.c_int(-2), C.c_int(-2)
    cuda.qwen_tts_cuda_cluster_status.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    cuda.qwen_tts_cuda_cluster_status(ctx, C.byref(cap), C.byref(err))
    return cap.value, err.value


def run(cuda, ctx, spec, mode, emb, n_pre, n_steps):
    H, V = spec["talker"]["hidden"], spec["talker"]["vocab"]
    assert cuda.qwen_tts_cuda_set_mode(ctx, mode) == 0
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.greedy_params()))
    cuda.qwen_tts_cuda_reset(ctx)
    assert cuda.qwen_tts_cuda_talker_prefill(ctx, B.fptr(emb), n_pre) == 0
    logits, hids = np.zeros((n_steps, V), dtype=np.float32), np.zeros((n_steps, H), dtype=np.float32)
    for i in range(n_steps):
        assert cuda.qwen_tts_cuda_talker_forward(ctx, B.fptr(emb[n_pre + i]), B.fptr(logits[i])) == 0, cuda.qwen_tts_cuda_last_error()
        assert cuda.qwen_tts_cuda_get_hidden(ctx, B.fptr(hids[i])) == 0
    subs = []
    for i, (pp, fc) in enumerate(((B.greedy_params(), 17), (B.default_params(), 901), (B.default_params(sample_seed=5, subtalker_top_k=4), 2047))):
        cuda.qwen_tts_cuda_set_params(ctx, C.byref(pp))
        out = np.zeros(32, dtype=np.int32)
        assert cuda.qwen_tts_cuda_subtalker_generate(ctx, B.fptr(np.ascontiguousarray(hids[i])), fc, B.iptr(out)) == 0
        subs.append(out[:16].copy())
    return logits, hids, np.array(subs)


def main():
    preset = sys.argv[1] if len(sys.argv) > 1 else "small"
    tpreset = sys.argv[2] if len(sys.argv) > 2 else "1.7b"
    cuda = B.load_cuda()
    assert cuda.qwen_tts_cuda_init(0) == 0
    spec = mg.PRESETS[preset]
    main_t, cod = mg.build_tensors(spec, seed=0, with_codec=False)
    hm = B.HostModel(spec, main_t, {})
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    rng = np.random.default_rng(0)
    n_pre, n_steps = 20, int(os.environ.get("STEPS", "70"))
    emb = (rng.standard_normal((n_pre + n_steps, spec["talker"]["hidden"])) * 0.5).astype(np.float32)
    l1, h1, s1 = run(cuda, ctx, spec, 1, emb, n_pre, n_steps)
    l2, h2, s2 = run(cuda, ctx, spec, 2, emb, n_pre, n_steps)
    print("cluster status (capacity, gave_up):", status(cuda, ctx))
    l3, h3, s3 = run(cuda, ctx, spec, 2, emb, n_pre, n_steps)
    rel = np.abs(l2 - l1).max(axis=1) / np.abs(l1).max(axis=1)
    print(f"{preset}: logits mode 2 vs mode 1: worst rel {rel.max():.3e} (step {rel.argmax()}), first steps {rel[:4]}, argmax equal {np.array_equal(l1.argmax(1), l2.argmax(1))}")
    print(f"  hidden rel {np.abs(h2 - h1).max() / np.abs(h1).max():.3e}; sub-talker codes equal: {np.array_equal(s1, s2)}")
    if not np.array_equal(s1, s2):
        print(s1); print(s2)
    print(f"  mode 2 twice bit-identical: logits {np.array_equal(l2.view(np.uint32), l3.view(np.uint32))}, codes {np.array_equal(s2, s3)}")
    cuda.qwen_tts_cuda_release(ctx)
    if tpreset == "none":
        return
    spec = mg.PRESETS[tpreset]
    main_t, cod = mg.build_tensors(spec, seed=0, with_codec=False)
    hm = B.HostModel(spec, main_t, {})
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx, cuda.qwen_tts_cuda_last_error()
    for mode in (1, 2):
        cuda.qwen_tts_cuda_set_mode(ctx, mode)
        t = cuda.qwen_tts_cuda_bench_talker_step(ctx, 190, 20, 1)
        f = cuda.qwen_tts_cuda_bench_subtalker(ctx, 20, 1)
        by = cuda.qwen_tts_cuda_talker_step_bytes(ctx, 190)
        print(f"{tpreset} mode {mode}: talker step {t * 1e3:.1f} us ({by / t / 1e6:.0f} GB/s), sub-talker frame {f * 1e3:.1f} us; status {status(cuda, ctx)}")
    cuda.qwen_tts_cuda_release(ctx)


if __name__ == "__main__":
    main()
