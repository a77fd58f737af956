/*
 * qtc_prefill.cuh - qwen_tts_talker_prefill (c/qwen_tts_talker.c:254-472) as ONE batched pass: all prompt tokens (of one
 * utterance, or of every utterance of a lock-step batch: R rows in total) go through each projection as a single tensor-core
 * GEMM (the reference does one kernel_matmul_bf16 per projection, :330-332, 425, 436-437, 448; round 1 ran seq_len full decode
 * steps, each streaming the 2.8 GB of weights again).
 *
 * Per layer: RMSNorm rows -> bf16 planes | Q, K, V GEMMs | per-head QK-norm + RoPE, K/V rows into the paged cache | causal
 * attention of every row over its own sequence's cache -> planes | O GEMM (+ residual) | RMSNorm -> planes | gate, up GEMMs |
 * SwiGLU -> planes | down GEMM (+ residual).  The GEMMs are qtc_conv_tc launches (k = 1 "convolutions" over the R rows) with
 * the talker's bf16 weights used in place as the single weight plane and the activations as the EXACT split of their fp32
 * values into three bf16 pieces (x . w = x_hi . w + x_mid . w + x_lo . w, every product exact, fp32 accumulation in TMEM): the
 * result differs from the per-token path by summation order only, like every fast-arithmetic kernel of this library (hidden
 * states within ~1e-5; test bar 2e-4).  Strict arithmetic keeps the per-token path.
 *
 * (included at the end of qtc_lm.cu)
 */
#pragma once

struct PfBuf {                       /* scratch of the batched prefill, grown on demand (counted in device_mallocs) */
    int cap = 0;                     /* rows */
    float *x = NULL, *q = NULL, *k = NULL, *v = NULL, *g = NULL, *u = NULL;
    uint16_t *ph = NULL, *pl = NULL, *px = NULL;   /* hi, mid, lo: x = hi + mid + lo exactly */
    int *slot = NULL, *pos = NULL;   /* [cap] sequence (KV pool slot) and position of every row */
};

/* the exact split of an fp32 value into three bf16 pieces (8 + 8 + 8 mantissa bits): every bf16 x bf16 product is then exact in
 * fp32 and the GEMM differs from the per-token path by summation order only - the class of difference fast arithmetic has anyway */
__device__ __forceinline__ uint32_t pf_rn(float v) { const uint32_t u = __float_as_uint(v); return (u + 0x7fffu + ((u >> 16) & 1u)) >> 16; }
__device__ __forceinline__ void pf_planes(uint16_t *ph, uint16_t *pl, uint16_t *px, size_t i, float v) {
    const uint32_t h = pf_rn(v);
    const float r1 = v - __uint_as_float(h << 16);
    const uint32_t m = pf_rn(r1);
    const float r2 = r1 - __uint_as_float(m << 16);
    ph[i] = (uint16_t)h; pl[i] = (uint16_t)m; px[i] = (uint16_t)pf_rn(r2);
}
/* planes of rmsnorm(x[r,:]) * w; one CTA per row */
__global__ void __launch_bounds__(256) k_pf_norm_planes(uint16_t *ph, uint16_t *pl, uint16_t *px, const float *x, const float *w, int n, float eps) {
    __shared__ float red[32];
    const float *xi = x + (size_t)blockIdx.x * n;
    float ss = 0.0f;
    for (int i = threadIdx.x; i < n; i += 256) ss = fmaf(xi[i], xi[i], ss);
    ss = block_sum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)n + eps);
    for (int i = threadIdx.x; i < n; i += 256) pf_planes(ph, pl, px, (size_t)blockIdx.x * n + i, xi[i] * inv * w[i]);
}
/* grid (R, heads + kv_heads), 128 threads (head_dim <= 256): per-head RMSNorm + rotate-half RoPE of q (in place) and k; k and v
 * rows of the token go into the paged cache of its sequence (c/qwen_tts_talker.c:150-196) */
struct PfQkArgs {
    float *q; const float *k, *v;           /* [R, qd], [R, kvd], [R, kvd] raw projections */
    const float *q_norm, *k_norm, *cos, *sin;
    float *kc, *vc;                         /* this layer's pool slice */
    const int *tables; int table_stride;
    const int *slot, *pos;
    int heads, kv_heads, hd;
    float eps;
};
__global__ void __launch_bounds__(128) k_pf_qk_rope_kv(const PfQkArgs a) {
    __shared__ float red[32];
    __shared__ float buf[256];
    const int r = blockIdx.x, h = blockIdx.y, hd = a.hd, half = hd >> 1, kvd = a.kv_heads * hd;
    const bool isq = h < a.heads;
    const int g = isq ? h : h - a.heads;
    const int pos = a.pos[r];
    const float *src = isq ? a.q + ((size_t)r * a.heads + g) * hd : a.k + ((size_t)r * a.kv_heads + g) * hd;
    const float *nw = isq ? a.q_norm : a.k_norm;
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;
    float ss = 0.0f;
    for (int i = threadIdx.x; i < hd; i += 128) { const float v = src[i]; ss = fmaf(v, v, ss); }
    ss = block_sum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)hd + a.eps);
    for (int i = threadIdx.x; i < hd; i += 128) buf[i] = src[i] * inv * nw[i];
    __syncthreads();
    const int *tab = a.tables + (size_t)a.slot[r] * a.table_stride;
    float *dst = isq ? a.q + ((size_t)r * a.heads + g) * hd : a.kc + kv_off(tab, pos, kvd) + g * hd;
    for (int i = threadIdx.x; i < half; i += 128) {
        const float xa = buf[i], xb = buf[i + half];
        dst[i] = xa * cs[i] - xb * sn[i];
        dst[i + half] = xb * cs[i + half] + xa * sn[i + half];
    }
    if (!isq) {
        const float *vr = a.v + ((size_t)r * a.kv_heads + g) * hd;
        float *vd = a.vc + kv_off(tab, pos, kvd) + g * hd;
        for (int i = threadIdx.x; i < hd; i += 128) vd[i] = vr[i];
    }
}
/* grid (R, heads), 128 threads: softmax(q . K[0..pos]) V over the row's own sequence -> planes [R, qd]; dynamic smem pos_max + 1 floats */
struct PfAttnArgs {
    const float *q; const float *kc, *vc;
    const int *tables; int table_stride;
    const int *slot, *pos;
    uint16_t *ph, *pl, *px;
    int heads, kv_heads, hd;
    float scale;
};
__global__ void __launch_bounds__(128) k_pf_attn(const PfAttnArgs a) {
    extern __shared__ float sc[];
    __shared__ float red[32];
    const int r = blockIdx.x, h = blockIdx.y, hd = a.hd, g = h / (a.heads / a.kv_heads), kvd = a.kv_heads * hd, qd = a.heads * hd;
    const int pos = a.pos[r], n = pos + 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int *tab = a.tables + (size_t)a.slot[r] * a.table_stride;
    const float *qh = a.q + ((size_t)r * a.heads + h) * hd;
    for (int t = warp; t < n; t += 4) {
        const float *kt = a.kc + kv_off(tab, t, kvd) + g * hd;
        float d = 0.0f;
        for (int i = lane; i < hd; i += 32) d = fmaf(qh[i], kt[i], d);
        d = warp_sum(d);
        if (lane == 0) sc[t] = d * a.scale;
    }
    __syncthreads();
    float mx = -3.402823466e+38f;
    for (int t = threadIdx.x; t < n; t += 128) mx = fmaxf(mx, sc[t]);
    mx = block_max(mx, red);
    float tot = 0.0f;
    for (int t = threadIdx.x; t < n; t += 128) { const float e = __expf(sc[t] - mx); sc[t] = e; tot += e; }
    tot = block_sum(tot, red);
    const float inv = 1.0f / tot;
    __syncthreads();
    for (int i = threadIdx.x; i < hd; i += 128) {
        float o = 0.0f;
        for (int t = 0; t < n; t++) o = fmaf(sc[t], a.vc[kv_off(tab, t, kvd) + g * hd + i], o);
        pf_planes(a.ph, a.pl, a.px, (size_t)r * qd + h * hd + i, o * inv);
    }
}
__global__ void k_pf_swiglu_planes(uint16_t *ph, uint16_t *pl, uint16_t *px, const float *g, const float *u, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pf_planes(ph, pl, px, i, swiglu_act<QTC_FAST>(g[i], u[i]));
}

static void pf_free(PfBuf &p) {
    cudaFree(p.x); cudaFree(p.q); cudaFree(p.k); cudaFree(p.v); cudaFree(p.g); cudaFree(p.u); cudaFree(p.ph); cudaFree(p.pl); cudaFree(p.px); cudaFree(p.slot); cudaFree(p.pos);
    for (void *q : {(void *)p.ph, (void *)p.pl, (void *)p.px}) qtc_gemm_tc_forget(q);
    p = PfBuf();
}
void qtc_prefill_release(qtc_ctx *c) { if (c->pf) { pf_free(*c->pf); delete c->pf; c->pf = NULL; } }
static int pf_reserve(qtc_ctx *c, int rows) {
    if (!c->pf) c->pf = new PfBuf();
    PfBuf &p = *c->pf;
    if (rows <= p.cap) return 0;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    pf_free(p);
    const LmDev *lm = &c->talker;
    const size_t n = (size_t)((rows + 127) / 128) * 128, H = lm->hidden, qd = (size_t)lm->heads * lm->head_dim, kvd = (size_t)lm->kv_heads * lm->head_dim, I = lm->inter;
    size_t wide = H > I ? H : I;
    if (qd > wide) wide = qd;
    QTC_CUDA(cudaMalloc(&p.x, n * H * 4)); QTC_CUDA(cudaMalloc(&p.q, n * qd * 4)); QTC_CUDA(cudaMalloc(&p.k, n * kvd * 4)); QTC_CUDA(cudaMalloc(&p.v, n * kvd * 4));
    QTC_CUDA(cudaMalloc(&p.g, n * I * 4)); QTC_CUDA(cudaMalloc(&p.u, n * I * 4));
    QTC_CUDA(cudaMalloc(&p.ph, n * wide * 2)); QTC_CUDA(cudaMalloc(&p.pl, n * wide * 2)); QTC_CUDA(cudaMalloc(&p.px, n * wide * 2));
    QTC_CUDA(cudaMalloc(&p.slot, n * 4)); QTC_CUDA(cudaMalloc(&p.pos, n * 4));
    c->dev_mallocs += 11;
    p.cap = (int)n;
    return 0;
}
static bool pf_ok(const qtc_ctx *c) {
    const LmDev *lm = &c->talker;
    const int qd = lm->heads * lm->head_dim, kvd = lm->kv_heads * lm->head_dim;
    return c->mode >= 1 && c->arith == QTC_FAST && lm->head_dim <= 256 && lm->head_dim % 2 == 0 && qtc_conv_tc_ok(lm->hidden, qd) && qtc_conv_tc_ok(lm->hidden, kvd) &&
           qtc_conv_tc_ok(qd, lm->hidden) && qtc_conv_tc_ok(lm->hidden, lm->inter) && qtc_conv_tc_ok(lm->inter, lm->hidden);
}

/* R rows already in c->pf->x ([R, H] fp32) with c->pf->slot / c->pf->pos filled (device).  pos_max: the largest position among them.
 * Leaves every row's post-layer hidden (before the final norm) in c->pf->x and the K/V rows in the pool. */
static int pf_run(qtc_ctx *c, int R, int pos_max) {
    const LmDev *lm = &c->talker;
    PfBuf &p = *c->pf;
    const int H = lm->hidden, hd = lm->head_dim, qd = lm->heads * hd, kvd = lm->kv_heads * hd, I = lm->inter;
    cudaStream_t st = c->stream;
    for (int l = 0; l < lm->layers; l++) {
        const LmLayerDev *w = &lm->layer[l];
        const ConvTcW wq = {w->wq, NULL, H, qd, 1, 1}, wk = {w->wk, NULL, H, kvd, 1, 1}, wv = {w->wv, NULL, H, kvd, 1, 1}, wo = {w->wo, NULL, qd, H, 1, 1};
        const ConvTcW wg = {w->gate, NULL, H, I, 1, 1}, wu = {w->up, NULL, H, I, 1, 1}, wd = {w->down, NULL, I, H, 1, 1};
        k_pf_norm_planes<<<R, 256, 0, st>>>(p.ph, p.pl, p.px, p.x, w->in_norm, H, lm->eps);
        LAUNCH_CHECK(c);
        if (qtc_conv_tc(c, &wq, p.ph, p.pl, R, 1, NULL, NULL, p.q, NULL, NULL, NULL, NULL, NULL, 0, p.px) || qtc_conv_tc(c, &wk, p.ph, p.pl, R, 1, NULL, NULL, p.k, NULL, NULL, NULL, NULL, NULL, 0, p.px) ||
            qtc_conv_tc(c, &wv, p.ph, p.pl, R, 1, NULL, NULL, p.v, NULL, NULL, NULL, NULL, NULL, 0, p.px)) return -1;
        PfQkArgs qa;
        qa.q = p.q; qa.k = p.k; qa.v = p.v; qa.q_norm = w->q_norm; qa.k_norm = w->k_norm; qa.cos = c->tk.cos; qa.sin = c->tk.sin;
        qa.kc = c->tk.k + (size_t)l * c->tk.kv_max * kvd; qa.vc = c->tk.v + (size_t)l * c->tk.kv_max * kvd;
        qa.tables = c->kv.tables; qa.table_stride = c->kv.max_seq_pages; qa.slot = p.slot; qa.pos = p.pos;
        qa.heads = lm->heads; qa.kv_heads = lm->kv_heads; qa.hd = hd; qa.eps = lm->eps;
        k_pf_qk_rope_kv<<<dim3(R, lm->heads + lm->kv_heads), 128, 0, st>>>(qa);
        LAUNCH_CHECK(c);
        PfAttnArgs aa;
        aa.q = p.q; aa.kc = qa.kc; aa.vc = qa.vc; aa.tables = qa.tables; aa.table_stride = qa.table_stride; aa.slot = p.slot; aa.pos = p.pos;
        aa.ph = p.ph; aa.pl = p.pl; aa.px = p.px; aa.heads = lm->heads; aa.kv_heads = lm->kv_heads; aa.hd = hd; aa.scale = 1.0f / sqrtf((float)hd);
        k_pf_attn<<<dim3(R, lm->heads), 128, (size_t)(pos_max + 2) * 4, st>>>(aa);
        LAUNCH_CHECK(c);
        if (qtc_conv_tc(c, &wo, p.ph, p.pl, R, 1, NULL, p.x, p.x, NULL, NULL, NULL, NULL, NULL, 0, p.px)) return -1;
        k_pf_norm_planes<<<R, 256, 0, st>>>(p.ph, p.pl, p.px, p.x, w->post_norm, H, lm->eps);
        LAUNCH_CHECK(c);
        if (qtc_conv_tc(c, &wg, p.ph, p.pl, R, 1, NULL, NULL, p.g, NULL, NULL, NULL, NULL, NULL, 0, p.px) || qtc_conv_tc(c, &wu, p.ph, p.pl, R, 1, NULL, NULL, p.u, NULL, NULL, NULL, NULL, NULL, 0, p.px)) return -1;
        k_pf_swiglu_planes<<<(unsigned)(((size_t)R * I + 255) / 256), 256, 0, st>>>(p.ph, p.pl, p.px, p.g, p.u, (size_t)R * I);
        LAUNCH_CHECK(c);
        if (qtc_conv_tc(c, &wd, p.ph, p.pl, R, 1, NULL, p.x, p.x, NULL, NULL, NULL, NULL, NULL, 0, p.px)) return -1;
    }
    return 0;
}

/* batch: x[r, :] = prefill[slot[r]][pos[r], :] (prompt rows of every utterance, back to back) */
__global__ void k_pf_gather_rows(float *x, const float *prefill, int rows_cap, const int *slot, const int *pos, int H) {
    const int r = blockIdx.y;
    const float *src = prefill + ((size_t)slot[r] * rows_cap + pos[r]) * H;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H; i += gridDim.x * blockDim.x) x[(size_t)r * H + i] = src[i];
}
/* out[b, :] = x[last[b], :]: every utterance's last prompt row */
__global__ void k_pf_take_last(float *out, const float *x, const int *last, int H) {
    const int b = blockIdx.y;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H; i += gridDim.x * blockDim.x) out[(size_t)b * H + i] = x[(size_t)last[b] * H + i];
}
static bool prefill_gemm_on() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("QTC_PREFILL_GEMM"); on = e ? atoi(e) : 1; }
    return on != 0;
}
__global__ void k_pf_fill_rows(int *slot, int *pos, int R, int s, const int *base_ptr) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < R) { slot[r] = s; pos[r] = *base_ptr + r; }
}
__global__ void k_pf_add_pos(GenState *st, int n) { st->pos += n; }

/* batch 1: seq_len rows of embeds_dev [seq_len, H] at positions state.pos .. ; leaves tk.x = final-normed hidden of the last
 * token (ctx->tk_x) and state.pos += seq_len.  `base` = the host's view of state.pos (pool pages are reserved by the caller). */
int qtc_prefill_gemm(qtc_ctx *c, const float *embeds_dev, int seq_len, int base) {
    if (pf_reserve(c, seq_len)) return -1;
    const int H = c->H;
    QTC_CUDA(cudaMemcpyAsync(c->pf->x, embeds_dev, (size_t)seq_len * H * 4, cudaMemcpyDeviceToDevice, c->stream));
    k_pf_fill_rows<<<(seq_len + 127) / 128, 128, 0, c->stream>>>(c->pf->slot, c->pf->pos, seq_len, 0, &c->state->pos);
    LAUNCH_CHECK(c);
    if (pf_run(c, seq_len, base + seq_len - 1)) return -1;
    if (qtc_rmsnorm(c, c->tk.x, c->pf->x + (size_t)(seq_len - 1) * H, c->talker.final_norm, H, c->talker.eps)) return -1;
    k_pf_add_pos<<<1, 1, 0, c->stream>>>(c->state, seq_len);
    LAUNCH_CHECK(c);
    return 0;
}
bool qtc_prefill_gemm_ok(const qtc_ctx *c) { return pf_ok(c) && prefill_gemm_on(); }
