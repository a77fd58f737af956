/*
 * qtc_kernel_api.cu - kernel-level test surface: the signatures of
 * c/qwen_tts_kernels.h:27-131 with host pointers in and out (one H2D / D2H per
 * call: parity tests, not speed), plus the measurement hooks of the header.
 * A private scratch context carries the stream and arithmetic mode.
 */
#include "qtc_internal.h"

#include <math.h>

int qtc_conv(qtc_ctx *c, float *out, const float *in, const float *w, const float *bias, const float *res, const float *alpha,
             const float *inv_beta, int cin, int rows, int K, int dil, int Lin, int up);
int qtc_sgemm(qtc_ctx *c, float *C, const float *A, const float *B, int M, int N, int K, const float *bias, const float *scale,
              const float *res, int gelu, int transposed, int ldc);
int qtc_conv_grouped(qtc_ctx *c, float *out, const float *in, const float *w, const float *bias, int cin, int cout, int K, int L, int dil, int groups);
int qtc_snake(qtc_ctx *c, float *out, const float *x, const float *alpha, const float *inv_beta, int C, int L);
int qtc_matmul_bf16(qtc_ctx *c, float *C, const float *A, const uint16_t *B, int M, int N, int K);

static qtc_ctx *g_k = NULL;
static int g_k_strict = 0;

static qtc_ctx *kctx() {
    if (!qwen_tts_cuda_is_available()) { qtc_set_error("kernel API before qwen_tts_cuda_init"); return NULL; }
    if (!g_k) {
        g_k = new qtc_ctx();
        g_k->launches = 0; g_k->arith = g_k_strict; g_k->mode = 0;
        if (cudaStreamCreateWithFlags(&g_k->stream, cudaStreamNonBlocking) != cudaSuccess) { delete g_k; g_k = NULL; return NULL; }
        g_k->cstream = g_k->stream; g_k->stream2 = NULL; g_k->codec_tc = 0; g_k->pf = NULL;
        g_k->sample_scratch = NULL; g_k->sample_scratch_f = NULL;
    }
    g_k->arith = g_k_strict;
    return g_k;
}
/* arithmetic policy of the kernel-level entry points: 0 fast, 1 strict (reference order) */
extern "C" int qwen_tts_cuda_kernel_arith(int strict) { g_k_strict = strict ? 1 : 0; return 0; }

struct DevBuf {
    void *p = NULL;
    DevBuf(size_t bytes) { if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) p = NULL; }
    ~DevBuf() { cudaFree(p); }
    template <class T> T *as() { return (T *)p; }
};
#define H2D(dst, src, bytes) QTC_CUDA(cudaMemcpy((dst), (src), (bytes), cudaMemcpyHostToDevice))
#define D2H(dst, src, bytes) QTC_CUDA(cudaMemcpy((dst), (src), (bytes), cudaMemcpyDeviceToHost))

extern "C" int qwen_tts_cuda_matvec_bf16(float *out, const uint16_t *A, const float *x, int rows, int cols) {
    qtc_ctx *c = kctx();
    if (!c || !out || !A || !x || rows <= 0 || cols <= 0) return -1;
    DevBuf dA((size_t)rows * cols * 2), dx((size_t)cols * 4), dy((size_t)rows * 4);
    if (!dA.p || !dx.p || !dy.p) return -1;
    H2D(dA.p, A, (size_t)rows * cols * 2); H2D(dx.p, x, (size_t)cols * 4);
    GemvSeg s = {dA.as<uint16_t>(), dy.as<float>(), rows};
    if (qtc_gemv(c, &s, 1, dx.as<float>(), cols, EPI_STORE, NULL)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(out, dy.p, (size_t)rows * 4);
    return 0;
}

extern "C" int qwen_tts_cuda_swiglu_matvec_bf16(float *out, const uint16_t *gate_up, const float *x, int inter, int hidden) {
    qtc_ctx *c = kctx();
    if (!c || !out || !gate_up || !x || inter <= 0 || hidden <= 0) return -1;
    DevBuf dA((size_t)2 * inter * hidden * 2), dx((size_t)hidden * 4), dy((size_t)inter * 4);
    if (!dA.p || !dx.p || !dy.p) return -1;
    H2D(dA.p, gate_up, (size_t)2 * inter * hidden * 2); H2D(dx.p, x, (size_t)hidden * 4);
    if (qtc_gemv_swiglu(c, dy.as<float>(), dA.as<uint16_t>(), dA.as<uint16_t>() + (size_t)inter * hidden, dx.as<float>(), inter, hidden)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(out, dy.p, (size_t)inter * 4);
    return 0;
}

extern "C" int qwen_tts_cuda_rms_norm(float *out, const float *x, const float *w, int dim, float eps) {
    qtc_ctx *c = kctx();
    if (!c || !out || !x || !w || dim <= 0) return -1;
    DevBuf dx((size_t)dim * 4), dw((size_t)dim * 4), dy((size_t)dim * 4);
    if (!dx.p || !dw.p || !dy.p) return -1;
    H2D(dx.p, x, (size_t)dim * 4); H2D(dw.p, w, (size_t)dim * 4);
    if (qtc_rmsnorm(c, dy.as<float>(), dx.as<float>(), dw.as<float>(), dim, eps)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(out, dy.p, (size_t)dim * 4);
    return 0;
}

/* returns the token, or -1 on failure; *rng_state is advanced like rand_uniform does */
extern "C" int qwen_tts_cuda_sample_top_k(const float *logits, int vocab, int top_k, float top_p, float temperature, float *rng_state) {
    qtc_ctx *c = kctx();
    if (!c || !logits || !rng_state || vocab <= 0 || vocab > 16384) return -1;
    DevBuf dl((size_t)vocab * 4), dr(16), dt(16), sf((size_t)vocab * 4), si((size_t)vocab * 4);
    if (!dl.p || !dr.p || !dt.p || !sf.p || !si.p) return -1;
    c->sample_scratch_f = sf.as<float>(); c->sample_scratch = si.as<int>();
    H2D(dl.p, logits, (size_t)vocab * 4); H2D(dr.p, rng_state, 4);
    int rc = qtc_sample_bare(c, dl.as<float>(), vocab, top_k, top_p, temperature, dr.as<uint32_t>(), dt.as<int>());
    if (rc == 0 && cudaStreamSynchronize(c->stream) != cudaSuccess) rc = -1;
    c->sample_scratch_f = NULL; c->sample_scratch = NULL;
    if (rc) return -1;
    int tok = -1;
    D2H(&tok, dt.p, 4); D2H(rng_state, dr.p, 4);
    return tok;
}

extern "C" int qwen_tts_cuda_causal_conv1d(float *out, const float *in, const float *w, const float *b, int cin, int cout, int k,
                                           int length, int dilation, int groups) {
    qtc_ctx *c = kctx();
    if (!c || !out || !in || !w || cin <= 0 || cout <= 0 || k <= 0 || length <= 0 || groups <= 0 || cin % groups || cout % groups) return -1;
    const size_t wn = (size_t)cout * (cin / groups) * k;
    DevBuf di((size_t)cin * length * 4), dw(wn * 4), db((size_t)cout * 4), dy((size_t)cout * length * 4);
    if (!di.p || !dw.p || !db.p || !dy.p) return -1;
    H2D(di.p, in, (size_t)cin * length * 4); H2D(dw.p, w, wn * 4);
    if (b) H2D(db.p, b, (size_t)cout * 4);
    int rc;
    if (groups == 1) rc = qtc_conv(c, dy.as<float>(), di.as<float>(), dw.as<float>(), b ? db.as<float>() : NULL, NULL, NULL, NULL, cin, cout, k, dilation, length, 1);
    else rc = qtc_conv_grouped(c, dy.as<float>(), di.as<float>(), dw.as<float>(), b ? db.as<float>() : NULL, cin, cout, k, length, dilation, groups);
    if (rc) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(out, dy.p, (size_t)cout * length * 4);
    return 0;
}

extern "C" int qwen_tts_cuda_transposed_conv1d(float *out, const float *in, const float *w, const float *b, int cin, int cout, int k,
                                               int stride, int length, int *out_length) {
    qtc_ctx *c = kctx();
    if (!c || !out || !in || !w || cin <= 0 || cout <= 0 || k <= 0 || stride <= 0 || length <= 0) return -1;
    int final_len = (length - 1) * stride + k - (k - stride);
    if (final_len < 0) final_len = 0;
    if (out_length) *out_length = final_len;
    if (k < stride) { qtc_set_error("transposed conv: kernel < stride is outside the codec's use"); return -1; }
    const int J = (k + stride - 1) / stride;
    std::vector<float> r((size_t)cout * stride * cin * J, 0.0f);
    for (int oc = 0; oc < cout; oc++)
        for (int p = 0; p < stride; p++)
            for (int ic = 0; ic < cin; ic++)
                for (int j = 0; j < J; j++) {
                    const int tap = p + (J - 1 - j) * stride;
                    if (tap < k) r[((size_t)(oc * stride + p) * cin + ic) * J + j] = w[((size_t)ic * cout + oc) * k + tap];
                }
    DevBuf di((size_t)cin * length * 4), dw(r.size() * 4), db((size_t)cout * 4), dy((size_t)cout * final_len * 4);
    if (!di.p || !dw.p || !db.p || !dy.p) return -1;
    H2D(di.p, in, (size_t)cin * length * 4); H2D(dw.p, r.data(), r.size() * 4);
    if (b) H2D(db.p, b, (size_t)cout * 4);
    if (qtc_conv(c, dy.as<float>(), di.as<float>(), dw.as<float>(), b ? db.as<float>() : NULL, NULL, NULL, NULL, cin, cout * stride, J, 1, length, stride)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(out, dy.p, (size_t)cout * final_len * 4);
    return 0;
}

extern "C" int qwen_tts_cuda_snake_beta(float *out, const float *x, const float *alpha, const float *beta, int channels, int length) {
    qtc_ctx *c = kctx();
    if (!c || !out || !x || !alpha || !beta || channels <= 0 || length <= 0) return -1;
    const size_t n = (size_t)channels * length;
    DevBuf dx(n * 4), da((size_t)channels * 4), db((size_t)channels * 4), dy(n * 4);
    if (!dx.p || !da.p || !db.p || !dy.p) return -1;
    H2D(dx.p, x, n * 4); H2D(da.p, alpha, (size_t)channels * 4); H2D(db.p, beta, (size_t)channels * 4);
    if (qtc_snake(c, dy.as<float>(), dx.as<float>(), da.as<float>(), db.as<float>(), channels, length)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(out, dy.p, n * 4);
    return 0;
}

extern "C" int qwen_tts_cuda_matmul_f32(float *C, const float *A, const float *B, int M, int N, int K) {
    qtc_ctx *c = kctx();
    if (!c || !C || !A || !B || M <= 0 || N <= 0 || K <= 0) return -1;
    DevBuf dA((size_t)M * K * 4), dB((size_t)N * K * 4), dC((size_t)M * N * 4);
    if (!dA.p || !dB.p || !dC.p) return -1;
    H2D(dA.p, A, (size_t)M * K * 4); H2D(dB.p, B, (size_t)N * K * 4);
    if (qtc_sgemm(c, dC.as<float>(), dA.as<float>(), dB.as<float>(), M, N, K, NULL, NULL, NULL, 0, 0, N)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(C, dC.p, (size_t)M * N * 4);
    return 0;
}

/* kernel_matmul_bf16 (prefill projections, c/qwen_tts_kernels.c:185-207): M token rows through the GEMV path */
extern "C" int qwen_tts_cuda_matmul_bf16(float *C, const float *A, const uint16_t *B, int M, int N, int K) {
    qtc_ctx *c = kctx();
    if (!c || !C || !A || !B || M <= 0 || N <= 0 || K <= 0) return -1;
    DevBuf dA((size_t)M * K * 4), dB((size_t)N * K * 2), dC((size_t)M * N * 4);
    if (!dA.p || !dB.p || !dC.p) return -1;
    H2D(dA.p, A, (size_t)M * K * 4); H2D(dB.p, B, (size_t)N * K * 2);
    for (int m = 0; m < M; m++) {
        GemvSeg s = {dB.as<uint16_t>(), dC.as<float>() + (size_t)m * N, N};
        if (qtc_gemv(c, &s, 1, dA.as<float>() + (size_t)m * K, K, EPI_STORE, NULL)) return -1;
    }
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    D2H(C, dC.p, (size_t)M * N * 4);
    return 0;
}

/* ---- measurement hooks ------------------------------------------------------------ */
static float *g_flush = NULL;
static const size_t FLUSH_BYTES = 512u << 20;       /* > 126 MB L2 */
__global__ void k_flush_fill(float *p, size_t n, float v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
static int flush_l2(cudaStream_t s) {
    if (!g_flush) QTC_CUDA(cudaMalloc(&g_flush, FLUSH_BYTES));
    k_flush_fill<<<148 * 8, 256, 0, s>>>(g_flush, FLUSH_BYTES / 4, 1.0f);
    return 0;
}

extern "C" double qwen_tts_cuda_talker_step_bytes(const qtc_ctx_t *c, int pos) {
    if (!c) return 0;
    const LmDev *lm = &c->talker;
    const double H = lm->hidden, qd = (double)lm->heads * lm->head_dim, kvd = (double)lm->kv_heads * lm->head_dim, I = lm->inter;
    const double per_layer = 2.0 * (qd * H + 2 * kvd * H + H * qd + 2 * I * H + H * I);
    return lm->layers * (per_layer + 2.0 * 4.0 * kvd * pos) + 2.0 * lm->vocab * H;   /* bf16 weights + fp32 K and V rows read */
}
extern "C" double qwen_tts_cuda_subtalker_frame_bytes(const qtc_ctx_t *c) {
    if (!c) return 0;
    const LmDev *lm = &c->sub;
    const double h = lm->hidden, qd = (double)lm->heads * lm->head_dim, kvd = (double)lm->kv_heads * lm->head_dim, I = lm->inter;
    const double per_fwd = lm->layers * 2.0 * (qd * h + 2 * kvd * h + h * qd + 2 * I * h + h * I);
    const double proj = c->sub_in_w ? 2.0 * h * c->H : 0.0;
    return c->G * (per_fwd + proj) + (c->G - 1) * 2.0 * lm->vocab * h;                 /* streamed figure, SURVEY 8d */
}

/* mean ms per talker decode step at KV position `pos`, CUDA events on the launch stream */
extern "C" double qwen_tts_cuda_bench_talker_step(qtc_ctx_t *c, int pos, int iters, int do_flush) {
    if (!c || c->talker.layers <= 0 || iters <= 0) return -1.0;
    if (qtc_ensure_kv(c, &c->tk, &c->talker, pos + 2)) return -1.0;
    cudaMemsetAsync(c->next_embed, 0, (size_t)c->H * 4, c->stream);
    double total = 0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < iters + 3; i++) {
        if (qtc_set_int(c, &c->state->pos, pos) || qtc_set_int(c, &c->state->done, 0)) return -1.0;
        if (do_flush && flush_l2(c->stream)) return -1.0;
        cudaEventRecord(e0, c->stream);
        c->tk_reach_hint = ((pos + 1 + 127) / 128) * 128;      /* the K/V stage sized as the frame loop sizes it at this position (qtc_generate.cu: talker_graph_for) */
        const int frc = qtc_talker_forward_dev(c, c->next_embed);
        c->tk_reach_hint = 0;
        if (frc) return -1.0;
        cudaEventRecord(e1, c->stream);
        if (cudaEventSynchronize(e1) != cudaSuccess) { qtc_set_error("bench: %s", cudaGetErrorString(cudaGetLastError())); return -1.0; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (i >= 3) total += ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return total / iters;
}

/* debug / profiles: SM-clock stamps of the fused talker step's phases in layer `layer`, 16 slots per CTA
 * (slot meaning: qtc_fused_kernel.cuh PROF(i)); out receives num_sms*16 values, returns num_sms */
extern "C" int qwen_tts_cuda_debug_phase_clocks(qtc_ctx_t *c, int pos, int layer, long long *out) {
    if (!c || !out || c->talker.layers <= 0) return -1;
    const size_t n = (size_t)c->num_sms * 32;
    if (!c->prof) { c->prof = (long long *)qtc_dmalloc(c, n * 8); if (!c->prof) return -1; }
    cudaMemsetAsync(c->prof, 0, n * 8, c->stream);
    c->prof_layer = layer;
    c->prof_sub = pos < 0;            /* pos < 0: profile the sub-talker frame (last code group wins) */
    if ((pos < 0 ? qwen_tts_cuda_bench_subtalker(c, 2, 1) : qwen_tts_cuda_bench_talker_step(c, pos, 2, 1)) < 0) return -1;
    c->prof_sub = 0;
    QTC_CUDA(cudaMemcpy(out, c->prof, n * 8, cudaMemcpyDeviceToHost));
    c->prof_layer = -1;
    return c->num_sms;
}

extern "C" double qwen_tts_cuda_bench_subtalker(qtc_ctx_t *c, int iters, int do_flush) {
    if (!c || c->sub.layers <= 0 || iters <= 0) return -1.0;
    cudaMemsetAsync(c->st_proj, 0, (size_t)c->H * 4, c->stream);
    double total = 0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < iters + 3; i++) {
        if (qtc_set_int(c, c->frame_codes, 1)) return -1.0;
        if (do_flush && flush_l2(c->stream)) return -1.0;
        cudaEventRecord(e0, c->stream);
        if (qtc_subtalker_dev(c, c->st_proj, c->frame_codes, c->frame_codes)) return -1.0;
        cudaEventRecord(e1, c->stream);
        if (cudaEventSynchronize(e1) != cudaSuccess) { qtc_set_error("bench: %s", cudaGetErrorString(cudaGetLastError())); return -1.0; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (i >= 3) total += ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return total / iters;
}

/* test surface of the tensor-core convolution (qtc_conv_tc.cu): kernel_causal_conv1d (stride <= 1) or kernel_transposed_conv1d
 * (stride > 1; w is then [cin, cout, k]) on channels-first host input [cin, length] with SnakeBeta(alpha, inv_beta) applied to
 * the input first (NULL: none) and an optional residual; the result is TIME-MAJOR [out_length, cout] fp32, as the vocoder keeps it */
extern "C" int qwen_tts_cuda_conv1d_tc(float *out_tm, const float *in, const float *w, const float *b, int cin, int cout, int k, int length,
                                       int dilation, int stride, const float *alpha, const float *inv_beta, const float *res_tm) {
    qtc_ctx *c = kctx();
    if (!c || !out_tm || !in || !w || cin <= 0 || cout <= 0 || k <= 0 || length <= 0) return -1;
    if (!qtc_conv_tc_ok(cin, cout)) { qtc_set_error("conv1d_tc: cin %% 8 / cout %% 32"); return -1; }
    const int s = stride > 1 ? stride : 1;
    const size_t Lout = (size_t)length * s;
    ConvTcW W;
    const size_t allocs0 = c->allocs.size();
    if (s > 1 ? qtc_tconv_tc_prep(c, &W, w, cin, cout, k, s) : qtc_conv_tc_prep(c, &W, w, cout, cin, k)) return -1;
    DevBuf di((size_t)cin * length * 4), dh((size_t)cin * length * 2), dl((size_t)cin * length * 2), db((size_t)cout * 4), dy(Lout * cout * 4);
    DevBuf da((size_t)cin * 4), dib((size_t)cin * 4);
    int rc = -1;
    if (di.p && dh.p && dl.p && db.p && dy.p && da.p && dib.p) {
        H2D(di.p, in, (size_t)cin * length * 4);
        if (b) H2D(db.p, b, (size_t)cout * 4);
        if (alpha) { H2D(da.p, alpha, (size_t)cin * 4); H2D(dib.p, inv_beta, (size_t)cin * 4); }
        if (res_tm) H2D(dy.p, res_tm, Lout * cout * 4);
        rc = qtc_cf_to_planes(c, dh.as<uint16_t>(), dl.as<uint16_t>(), di.as<float>(), cin, length, alpha ? da.as<float>() : NULL, alpha ? dib.as<float>() : NULL);
        if (rc == 0) rc = qtc_conv_tc(c, &W, dh.as<uint16_t>(), dl.as<uint16_t>(), length, dilation, b ? db.as<float>() : NULL, res_tm ? dy.as<float>() : NULL,
                                      dy.as<float>(), NULL, NULL, NULL, NULL);
        if (rc == 0 && cudaStreamSynchronize(c->stream) != cudaSuccess) { qtc_set_error("conv1d_tc: %s", cudaGetErrorString(cudaGetLastError())); rc = -1; }
        if (rc == 0) D2H(out_tm, dy.p, Lout * cout * 4);
    }
    qtc_gemm_tc_forget(dh.p); qtc_gemm_tc_forget(dl.p);
    while (c->allocs.size() > allocs0) { qtc_gemm_tc_forget(c->allocs.back()); cudaFree(c->allocs.back()); c->allocs.pop_back(); }
    return rc;
}
