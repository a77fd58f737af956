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
 * R <= 64 rows (every batch-1 prompt): the projections run as the skinny GEMM of the batched decode path instead
 * (qtc_gemm_tc.cu: the WEIGHTS are the 128-row MMA operand streamed by TMA, the R activation rows - the same three pieces,
 * stacked as the N operand - stay put; split-K over ~148 CTAs), because the op is then weight-bandwidth bound: the convolution
 * kernel at M = 15 keeps ~1 MB of weights in flight and took 7.2 ms for a 15-token prompt of the 1.7B talker.
 *
 * (included at the end of qtc_lm.cu)
 */
#pragma once

struct PfBuf {                       /* scratch of the batched prefill, grown on demand (counted in device_mallocs) */
    int cap = 0;                     /* rows */
    float *x = NULL, *qkvgu = NULL;  /* [cap, H]; one block the q | k | v | gate | up outputs of a pass are carved from (R rows each) */
    uint16_t *planes = NULL;         /* hi, mid, lo pieces (x = hi + mid + lo exactly): three [cap, wide] planes, or - skinny path - piece p of row r at row 64 p + r of a [192, K] matrix */
    size_t wide = 0;
    float *parts = NULL;             /* skinny path: K-split partial sums of a layer's seven projections (64 rows each) */
    int ns[7] = {0}; size_t poff[7] = {0};
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
/* A projection output as the skinny GEMM leaves it: n K-split partial sums `stride` floats apart, added in split order wherever the
 * value is consumed (a fixed order: the prefill stays bit-reproducible run to run, which fp32 atomics into one array would not
 * be).  n == 0: p is the finished array (the convolution-kernel path). */
struct PfParts { const float *p; int n, stride; };
__device__ __forceinline__ float pf_get(const PfParts &a, size_t i) {
    float s = a.p[i];
    for (int k = 1; k < a.n; k++) s += a.p[(size_t)k * a.stride + i];
    return s;
}
/* planes of rmsnorm(x[r,:]) * w; one CTA per row.  add.n > 0: first x[r,:] += the partial sums of the projection before it (O or
 * down: the residual add), written back */
__global__ void __launch_bounds__(256) k_pf_norm_planes(uint16_t *ph, uint16_t *pl, uint16_t *px, float *x, const float *w, int n, float eps, const PfParts add) {
    __shared__ float red[32];
    float *xi = x + (size_t)blockIdx.x * n;
    float ss = 0.0f;
    if (add.n > 0) for (int i = threadIdx.x; i < n; i += 256) xi[i] += pf_get(add, (size_t)blockIdx.x * n + i);     /* the same thread reads i back below */
    for (int i = threadIdx.x; i < n; i += 256) ss = fmaf(xi[i], xi[i], ss);
    ss = block_sum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)n + eps);
    for (int i = threadIdx.x; i < n; i += 256) pf_planes(ph, pl, px, (size_t)blockIdx.x * n + i, xi[i] * inv * w[i]);
}
/* x[r,:] += partial sums (the last layer's down projection) */
__global__ void __launch_bounds__(256) k_pf_fold(float *x, int n, const PfParts add) {
    for (int i = threadIdx.x; i < n; i += 256) x[(size_t)blockIdx.x * n + i] += pf_get(add, (size_t)blockIdx.x * n + i);
}
/* grid (R, heads + kv_heads), 128 threads (head_dim <= 256): per-head RMSNorm + rotate-half RoPE of q (in place) and k; k and v
 * rows of the token go into the paged cache of its sequence (c/qwen_tts_talker.c:150-196) */
struct PfQkArgs {
    float *q; PfParts qp, kp, vp;           /* raw projections [R, qd], [R, kvd], [R, kvd] (q is rotated in place: q == qp.p) */
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
    const PfParts &sp = isq ? a.qp : a.kp;
    const size_t s0 = isq ? ((size_t)r * a.heads + g) * hd : ((size_t)r * a.kv_heads + g) * hd;
    const float *nw = isq ? a.q_norm : a.k_norm;
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;
    float ss = 0.0f;
    for (int i = threadIdx.x; i < hd; i += 128) { const float v = pf_get(sp, s0 + i); buf[i] = v; ss = fmaf(v, v, ss); }    /* the same thread reads buf[i] back */
    ss = block_sum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)hd + a.eps);
    for (int i = threadIdx.x; i < hd; i += 128) buf[i] = buf[i] * inv * nw[i];
    __syncthreads();
    const int *tab = a.tables + (size_t)a.slot[r] * a.table_stride;
    float *dst = isq ? a.q + ((size_t)r * a.heads + g) * hd : a.kc + kv_off(tab, pos, kvd) + g * hd;
    for (int i = threadIdx.x; i < half; i += 128) {
        const float xa = buf[i], xb = buf[i + half];
        dst[i] = xa * cs[i] - xb * sn[i];
        dst[i + half] = xb * cs[i + half] + xa * sn[i + half];
    }
    if (!isq) {
        float *vd = a.vc + kv_off(tab, pos, kvd) + g * hd;
        for (int i = threadIdx.x; i < hd; i += 128) vd[i] = pf_get(a.vp, ((size_t)r * a.kv_heads + g) * hd + i);
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
__global__ void k_pf_swiglu_planes(uint16_t *ph, uint16_t *pl, uint16_t *px, const PfParts g, const PfParts u, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pf_planes(ph, pl, px, i, swiglu_act<QTC_FAST>(pf_get(g, i), pf_get(u, i)));
}

int qtc_gemm_tc_splits(const qtc_ctx *c, int N, int K);
int qtc_gemm_tc_parts(qtc_ctx *c, float *parts, int ldy, int split_stride, const uint16_t *X, const uint16_t *W, int nbatch, int N, int K);
void qtc_gemm_tc_forget(const void *ptr);
#define PF_SKINNY_ROWS 64            /* TC_BN of qtc_gemm_tc.cu */

static void pf_free(PfBuf &p) {
    cudaFree(p.x); cudaFree(p.qkvgu); cudaFree(p.planes); cudaFree(p.parts); cudaFree(p.slot); cudaFree(p.pos);
    for (int k = 0; k < 3 && p.planes; k++) qtc_gemm_tc_forget(p.planes + (size_t)k * p.cap * p.wide);   /* tensor maps are cached by base pointer: the three planes (the skinny matrix starts at plane 0) */
    p = PfBuf();
}
void qtc_prefill_release(qtc_ctx *c) { if (c->pf) { pf_free(*c->pf); delete c->pf; c->pf = NULL; } }
static bool pf_skinny_on() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("QTC_PREFILL_SKINNY"); on = e ? atoi(e) : 1; }
    return on != 0;
}
static bool pf_skinny_ok(const qtc_ctx *c, int R) {
    const LmDev *lm = &c->talker;
    return R <= PF_SKINNY_ROWS && pf_skinny_on() && lm->hidden % 64 == 0 && (lm->heads * lm->head_dim) % 64 == 0 && lm->inter % 64 == 0;
}
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
    QTC_CUDA(cudaMalloc(&p.x, n * H * 4)); QTC_CUDA(cudaMalloc(&p.qkvgu, n * (qd + 2 * kvd + 2 * I) * 4));
    QTC_CUDA(cudaMalloc(&p.planes, 3 * n * wide * 2));                    /* n >= 128: also holds the skinny path's [192, wide] */
    QTC_CUDA(cudaMemsetAsync(p.planes, 0, 3 * n * wide * 2, c->stream));  /* the skinny GEMM reads all 64 rows of a piece; rows past R only feed dropped columns, but keep them finite */
    QTC_CUDA(cudaMalloc(&p.slot, n * 4)); QTC_CUDA(cudaMalloc(&p.pos, n * 4));
    c->dev_mallocs += 5;
    if (pf_skinny_ok(c, 1)) {
        /* K-split partial sums of the seven projections of a layer, 64 rows each: [q | k | v | gate | up | o | down] */
        const int N[7] = {(int)qd, (int)kvd, (int)kvd, (int)I, (int)I, (int)H, (int)H}, K[7] = {(int)H, (int)H, (int)H, (int)H, (int)H, (int)qd, (int)I};
        size_t tot = 0;
        for (int i = 0; i < 7; i++) { p.ns[i] = qtc_gemm_tc_splits(c, N[i], K[i]); p.poff[i] = tot; tot += (size_t)p.ns[i] * PF_SKINNY_ROWS * N[i]; }
        QTC_CUDA(cudaMalloc(&p.parts, tot * 4));
        c->dev_mallocs += 1;
    }
    p.cap = (int)n; p.wide = wide;
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
    const bool skinny = pf_skinny_ok(c, R) && p.parts;
    enum { PQ, PK, PV, PG, PU, PO, PD };
    /* projection outputs, R rows each.  Convolution-kernel path: finished arrays carved from qkvgu (O and down add into x themselves).
     * Skinny path: every projection leaves ns[i] K-split partial arrays, R * N apart, which its consumer adds up in split order. */
    const int Ns[7] = {qd, kvd, kvd, I, I, H, H};
    float *out[7]; PfParts part[7];
    for (int i = 0, off = 0; i < 7; i++) {
        if (skinny) { out[i] = p.parts + p.poff[i]; part[i] = {out[i], p.ns[i], R * Ns[i]}; }
        else { out[i] = i < PO ? p.qkvgu + (size_t)R * off : p.x; part[i] = {out[i], 0, 0}; off += Ns[i]; }
    }
    const PfParts none = {NULL, 0, 0};
    /* piece planes of an [R, K] activation matrix: three [R, K] planes `cap * wide` apart, or (skinny) rows 0.., 64.., 128.. of one [192, K] matrix */
    auto plane = [&](int k, int K) { return p.planes + (skinny ? (size_t)k * PF_SKINNY_ROWS * K : (size_t)k * p.cap * p.wide); };
    auto gemm = [&](int i, const uint16_t *W, int K) -> int {
        const int N = Ns[i];
        if (skinny) return qtc_gemm_tc_parts(c, out[i], N, R * N, p.planes, W, R, N, K);
        const ConvTcW cw = {W, NULL, K, N, 1, 1};
        return qtc_conv_tc(c, &cw, plane(0, K), plane(1, K), R, 1, NULL, i >= PO ? p.x : NULL, out[i], NULL, NULL, NULL, NULL, NULL, 0, plane(2, K));
    };
    for (int l = 0; l < lm->layers; l++) {
        const LmLayerDev *w = &lm->layer[l];
        k_pf_norm_planes<<<R, 256, 0, st>>>(plane(0, H), plane(1, H), plane(2, H), p.x, w->in_norm, H, lm->eps, skinny && l > 0 ? part[PD] : none);    /* x += the layer before's down partials */
        LAUNCH_CHECK(c);
        if (gemm(PQ, w->wq, H) || gemm(PK, w->wk, H) || gemm(PV, w->wv, H)) return -1;
        PfQkArgs qa;
        qa.q = out[PQ]; qa.qp = part[PQ]; qa.kp = part[PK]; qa.vp = part[PV]; qa.q_norm = w->q_norm; qa.k_norm = w->k_norm; qa.cos = c->tk.cos; qa.sin = c->tk.sin;
        qa.kc = c->tk.k + (size_t)l * c->tk.kv_max * kvd; qa.vc = c->tk.v + (size_t)l * c->tk.kv_max * kvd;
        qa.tables = c->kv.tables; qa.table_stride = c->kv.max_seq_pages; qa.slot = p.slot; qa.pos = p.pos;
        qa.heads = lm->heads; qa.kv_heads = lm->kv_heads; qa.hd = hd; qa.eps = lm->eps;
        k_pf_qk_rope_kv<<<dim3(R, lm->heads + lm->kv_heads), 128, 0, st>>>(qa);
        LAUNCH_CHECK(c);
        PfAttnArgs aa;
        aa.q = out[PQ]; aa.kc = qa.kc; aa.vc = qa.vc; aa.tables = qa.tables; aa.table_stride = qa.table_stride; aa.slot = p.slot; aa.pos = p.pos;
        aa.ph = plane(0, qd); aa.pl = plane(1, qd); aa.px = plane(2, qd); aa.heads = lm->heads; aa.kv_heads = lm->kv_heads; aa.hd = hd; aa.scale = 1.0f / sqrtf((float)hd);
        k_pf_attn<<<dim3(R, lm->heads), 128, (size_t)(pos_max + 2) * 4, st>>>(aa);
        LAUNCH_CHECK(c);
        if (gemm(PO, w->wo, qd)) return -1;                                /* x += Wo . att (skinny: by the norm kernel that follows) */
        k_pf_norm_planes<<<R, 256, 0, st>>>(plane(0, H), plane(1, H), plane(2, H), p.x, w->post_norm, H, lm->eps, skinny ? part[PO] : none);
        LAUNCH_CHECK(c);
        if (gemm(PG, w->gate, H) || gemm(PU, w->up, H)) return -1;
        k_pf_swiglu_planes<<<(unsigned)(((size_t)R * I + 255) / 256), 256, 0, st>>>(plane(0, I), plane(1, I), plane(2, I), part[PG], part[PU], (size_t)R * I);
        LAUNCH_CHECK(c);
        if (gemm(PD, w->down, I)) return -1;                               /* x += Wdown . h (skinny: by the next layer's norm kernel, or the fold below) */
    }
    if (skinny) { k_pf_fold<<<R, 256, 0, st>>>(p.x, H, part[PD]); LAUNCH_CHECK(c); }
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
