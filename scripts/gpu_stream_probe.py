"""where the time of a streamed utterance goes: perf counters of qwen_tts_cuda_generate with streaming off / on"""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
from harness import binding as B, modelgen as mg
cuda = B.load_cuda(); assert cuda.qwen_tts_cuda_init(0) == 0
spec = mg.PRESETS["1.7b"]; main, cod = mg.build_tensors(spec, seed=0); hm = B.HostModel(spec, main, cod)
ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
CSV = b"151644,77091,198,2354,2244,2355,4041,2244,11752,13,151645,198,151644,77091,198"
p = B.default_params(fixed_codec_tokens=360); cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
perf = B.Perf()
for first, chunk in ((0, 0), (8, 160), (8, 352), (8, 40), (0, 0)):
    cuda.qwen_tts_cuda_set_streaming(ctx, first, chunk)
    for rep in range(3):
        ns, nf = C.c_int(0), C.c_int(0); codes = np.zeros((360, 16), dtype=np.int32)
        t0 = time.perf_counter()
        ptr = cuda.qwen_tts_cuda_generate(ctx, CSV, b"aiden", b"english", C.byref(ns), B.iptr(codes), 360, C.byref(nf))
        wall = (time.perf_counter() - t0) * 1e3
        cuda.qwen_tts_cuda_free_host(ptr); cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    print(f"stream({first},{chunk}): wall {wall:7.1f} ms total {perf.total_ms:7.1f} prefill {perf.prefill_ms:5.1f} loop {perf.talker_ms:7.1f} codec_tail {perf.codec_ms:6.1f} device {perf.device_ms:7.1f} ttfa {perf.ttfa_ms:6.1f} p99 {perf.frame_ms_p99:5.2f} launches {perf.kernel_launches}")
