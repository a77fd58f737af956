/*
 * qtc_lm.cu - launch logic of the per-op decode path (mode 0): one decoder
 * stack step (talker_attention_single x layers, c/qwen_tts_talker.c:119-248,
 * 520-527; ST_FORWARD :638-691), the talker forward with codec head (:478-533),
 * the sub-talker code-group loop (:539-736) and the device-resident front half
 * of a frame (c/qwen_tts.c:1302-1363).  All work is enqueued on ctx->stream and
 * reads positions / tokens from device memory, so every sequence here can be
 * captured into a CUDA graph and replayed.
 */
#include "qtc_fused_main.cuh"
#include "qtc_cluster.cuh"
#include "qtc_gemv.cuh"
#include "qtc_internal.h"
#include "qtc_lm_kernels.cuh"
#include "qtc_sample.cuh"

#include <math.h>

#define LAUNCH_CHECK(c)                                                             \
    do {                                                                            \
        (c)->launches++;                                                            \
        cudaError_t e_ = cudaGetLastError();                                        \
        if (e_ != cudaSuccess) { qtc_set_error("%s:%d launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); return -1; } \
    } while (0)

static int g_attr_done = 0;
static int set_smem_attrs() {
    if (g_attr_done) return 0;
    QTC_CUDA(cudaFuncSetAttribute(k_gemv_strict, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    QTC_CUDA(cudaFuncSetAttribute(k_gemv_swiglu_strict, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    QTC_CUDA(cudaFuncSetAttribute(k_gemv_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    QTC_CUDA(cudaFuncSetAttribute(k_gemv_swiglu_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    QTC_CUDA(cudaFuncSetAttribute(k_attn_decode<QTC_FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    QTC_CUDA(cudaFuncSetAttribute(k_attn_decode<QTC_STRICT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    QTC_CUDA(cudaFuncSetAttribute(k_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
    g_attr_done = 1;
    return 0;
}

int qtc_gemv(qtc_ctx *c, const GemvSeg *segs, int nseg, const float *x, int cols, int epi, const float *aux) {
    if (set_smem_attrs()) return -1;
    GemvArgs a;
    memset(&a, 0, sizeof(a));
    a.nseg = nseg; a.cols = cols; a.epi = epi; a.x = x; a.aux = aux;
    int total = 0, groups = 0;
    for (int i = 0; i < nseg; i++) { a.w[i] = segs[i].w; a.out[i] = segs[i].out; a.rows[i] = segs[i].rows; total += segs[i].rows; groups += (segs[i].rows + 31) / 32; }
    a.total_rows = total;
    if (epi != EPI_STORE && nseg != 1) { qtc_set_error("gemv: epilogue needs a single segment"); return -1; }
    if (cols % 8 != 0 && c->arith == QTC_FAST) { qtc_set_error("gemv: cols %% 8 != 0"); return -1; }
    if (c->arith == QTC_STRICT) {
        if (cols % STRICT_TILE_K == 0) {
            const size_t smem = (size_t)cols * 4 + STRICT_WARPS * STRICT_WARP_SMEM;
            k_gemv_strict<<<(groups + STRICT_WARPS - 1) / STRICT_WARPS, STRICT_WARPS * 32, smem, c->stream>>>(a);
        } else {
            k_gemv_strict_generic<<<(total + 127) / 128, 128, 0, c->stream>>>(a);
        }
    } else {
        k_gemv_fast<<<(total + GEMV_FAST_ROWS - 1) / GEMV_FAST_ROWS, GEMV_FAST_THREADS, (size_t)cols * 4, c->stream>>>(a);
    }
    LAUNCH_CHECK(c);
    return 0;
}

int qtc_gemv_swiglu(qtc_ctx *c, float *out, const uint16_t *gate, const uint16_t *up, const float *x, int inter, int cols) {
    if (set_smem_attrs()) return -1;
    if (c->arith == QTC_STRICT) {
        if (cols % STRICT_TILE_K == 0) {
            const size_t smem = (size_t)cols * 4 + STRICT_WARPS * STRICT_WARP_SMEM;
            k_gemv_swiglu_strict<<<(inter + 63) / 64, STRICT_WARPS * 32, smem, c->stream>>>(out, gate, up, x, inter, cols);
            LAUNCH_CHECK(c);
            return 0;
        }
        qtc_set_error("swiglu strict: cols %% 64 != 0");
        return -1;
    }
    int grid = (inter + 7) / 8;
    if (grid > 148 * 4) grid = 148 * 4;
    k_gemv_swiglu_fast<<<grid, GEMV_FAST_THREADS, (size_t)cols * 4, c->stream>>>(out, gate, up, x, inter, cols);
    LAUNCH_CHECK(c);
    return 0;
}

int qtc_rmsnorm(qtc_ctx *c, float *out, const float *x, const float *w, int n, float eps) {
    if (c->arith == QTC_STRICT) k_rmsnorm<QTC_STRICT><<<1, 256, 0, c->stream>>>(out, x, w, n, eps);
    else k_rmsnorm<QTC_FAST><<<1, 256, 0, c->stream>>>(out, x, w, n, eps);
    LAUNCH_CHECK(c);
    return 0;
}

int qtc_embed_row(qtc_ctx *c, float *out, const uint16_t *table, const int *idx_ptr, int idx_imm, int n) {
    k_embed_row<<<(n + 255) / 256, 256, 0, c->stream>>>(out, table, idx_ptr, idx_imm, n);
    LAUNCH_CHECK(c);
    return 0;
}

/* RoPE rows exactly as the oracle computes them (c/qwen_tts_talker.c:76-113): host libm, uploaded */
static void rope_rows_host(float *cs, float *sn, int pos0, int npos, int hd, float theta) {
    const int half = hd / 2;
    for (int p = 0; p < npos; p++)
        for (int i = 0; i < half; i++) {
            const float freq = 1.0f / powf(theta, (float)(2 * i) / (float)hd);
            const float ang = (float)(pos0 + p) * freq;
            const float cv = cosf(ang), sv = sinf(ang);
            cs[(size_t)p * hd + i] = cv; cs[(size_t)p * hd + i + half] = cv;
            sn[(size_t)p * hd + i] = sv; sn[(size_t)p * hd + i + half] = sv;
        }
}

/* ---- paged KV pool (north_star: "GQA attention over an HBM-resident paged KV cache") -----------------------------------
 * One allocation per K and V at upload: [layers][pages][64][kvd] fp32; RoPE rows for every position a sequence can reach are
 * computed then too.  After that, growing a sequence is a host-side pop from the free list plus a 4-byte table update: no
 * cudaMalloc, no device-to-device copy of the cache, no stale graph (the reference reallocs and copies the whole cache,
 * c/qwen_tts_talker.c:489-496; round 1 did the same on the device). */
int qtc_kv_pool_init(qtc_ctx *c, int max_batch) {
    const LmDev *lm = &c->talker;
    const size_t kvd = (size_t)lm->kv_heads * lm->head_dim, hd = lm->head_dim;
    KvPool &p = c->kv;
    p.slots = max_batch > 1 ? (max_batch > 64 ? 64 : max_batch) : 1;
    /* a single utterance may run to max_new_tokens = 4096 frames after a prompt; a batch row gets 576 positions on average */
    int positions = 4096 + 192;
    if (p.slots * 576 > positions) positions = p.slots * 576;
    p.pages = (positions + QTC_KV_PAGE_POS - 1) / QTC_KV_PAGE_POS;
    p.max_seq_pages = p.pages;
    const size_t floats = (size_t)lm->layers * p.pages * QTC_KV_PAGE_POS * kvd;
    QTC_CUDA(cudaMalloc(&c->tk.k, floats * 4 + 16));
    QTC_CUDA(cudaMalloc(&c->tk.v, floats * 4 + 16));
    QTC_CUDA(cudaMalloc(&p.tables, (size_t)p.slots * p.max_seq_pages * 4));
    QTC_CUDA(cudaMallocHost(&p.tables_host, (size_t)p.slots * p.max_seq_pages * 4));
    memset(p.tables_host, 0, (size_t)p.slots * p.max_seq_pages * 4);
    QTC_CUDA(cudaMemset(p.tables, 0, (size_t)p.slots * p.max_seq_pages * 4));
    p.free_pages.clear();
    for (int i = p.pages - 1; i >= 0; i--) p.free_pages.push_back(i);      /* page 0 is handed out first */
    p.used.assign(p.slots, 0);
    c->tk.kv_max = p.pages * QTC_KV_PAGE_POS;
    c->tk.table = p.tables;
    const int npos = c->tk.kv_max;
    std::vector<float> hc((size_t)npos * hd), hs((size_t)npos * hd);
    rope_rows_host(hc.data(), hs.data(), 0, npos, (int)hd, lm->theta);
    QTC_CUDA(cudaMalloc(&c->tk.cos, hc.size() * 4));
    QTC_CUDA(cudaMalloc(&c->tk.sin, hs.size() * 4));
    QTC_CUDA(cudaMemcpy(c->tk.cos, hc.data(), hc.size() * 4, cudaMemcpyHostToDevice));
    QTC_CUDA(cudaMemcpy(c->tk.sin, hs.data(), hs.size() * 4, cudaMemcpyHostToDevice));
    return 0;
}
void qtc_kv_release_all(qtc_ctx *c) {
    KvPool &p = c->kv;
    for (int s = 0; s < p.slots; s++) {
        for (int i = p.used[s] - 1; i >= 0; i--) p.free_pages.push_back(p.tables_host[(size_t)s * p.max_seq_pages + i]);
        p.used[s] = 0;
    }
}
int qtc_kv_reserve(qtc_ctx *c, int slot, int positions) {
    KvPool &p = c->kv;
    if (slot < 0 || slot >= p.slots) { qtc_set_error("kv: slot %d outside the %d slots this context was uploaded with (max_batch)", slot, p.slots); return -1; }
    const int need = (positions + QTC_KV_PAGE_POS - 1) / QTC_KV_PAGE_POS, have = p.used[slot];
    if (need <= have) return 0;
    if (need > p.max_seq_pages || need - have > (int)p.free_pages.size()) {
        qtc_set_error("kv: pool exhausted (%d pages of 64 positions, %zu free, sequence %d needs %d more): upload with a larger max_batch",
                      p.pages, p.free_pages.size(), slot, need - have);
        return -1;
    }
    int *row = p.tables_host + (size_t)slot * p.max_seq_pages;
    for (int i = have; i < need; i++) { row[i] = p.free_pages.back(); p.free_pages.pop_back(); p.page_allocs++; }
    p.used[slot] = need;
    /* the new entries only; kernels already enqueued never read them (their positions are lower) */
    QTC_CUDA(cudaMemcpyAsync(p.tables + (size_t)slot * p.max_seq_pages + have, row + have, (size_t)(need - have) * 4, cudaMemcpyHostToDevice, c->stream));
    return 0;
}

/* talker: pages for `need` positions of slot 0.  Sub-talker: its contiguous cache of G + 2 positions, allocated once at upload. */
int qtc_ensure_kv(qtc_ctx *c, LmBuffers *b, const LmDev *lm, int need) {
    if (b == &c->tk) {
        if (need > b->kv_max) { qtc_set_error("kv: %d positions asked, the pool holds %d per sequence", need, b->kv_max); return -1; }
        return qtc_kv_reserve(c, 0, need);
    }
    if (need <= b->kv_max) return 0;
    if (b->kv_max > 0) { qtc_set_error("kv: the sub-talker cache holds %d positions, %d asked", b->kv_max, need); return -1; }
    const int nmax = ((need + 63) / 64) * 64;
    const size_t kvd = (size_t)lm->kv_heads * lm->head_dim, hd = lm->head_dim;
    QTC_CUDA(cudaMalloc(&b->k, (size_t)lm->layers * nmax * kvd * 4 + 16));
    QTC_CUDA(cudaMalloc(&b->v, (size_t)lm->layers * nmax * kvd * 4 + 16));
    QTC_CUDA(cudaMalloc(&b->cos, (size_t)nmax * hd * 4));
    QTC_CUDA(cudaMalloc(&b->sin, (size_t)nmax * hd * 4));
    std::vector<float> hc((size_t)nmax * hd), hs((size_t)nmax * hd);
    rope_rows_host(hc.data(), hs.data(), 0, nmax, (int)hd, lm->theta);
    QTC_CUDA(cudaMemcpy(b->cos, hc.data(), hc.size() * 4, cudaMemcpyHostToDevice));
    QTC_CUDA(cudaMemcpy(b->sin, hs.data(), hs.size() * 4, cudaMemcpyHostToDevice));
    b->kv_max = nmax; b->table = NULL;
    return 0;
}


/* ---- mode 2: the cluster kernel (qtc_cluster.cuh) ------------------------------------------------------------------ */
#define CL_MAX_CLUSTERS 15            /* 8-CTA clusters a B200 holds at one CTA per SM (scripts/probe/cluster_capacity.cu) */
static bool cluster_geometry(const LmDev *lm) {
    return lm->head_dim == 128 && lm->kv_heads > 0 && lm->kv_heads <= CL_MAX_CLUSTERS && lm->heads == 2 * lm->kv_heads &&
           lm->hidden % 1024 == 0 && lm->hidden <= 2048 && lm->inter % 8 == 0 && lm->inter / 8 >= 2 * CL_MAX_CLUSTERS * CL_SIZE &&
           lm->inter / 8 <= 8 * CL_MAX_CLUSTERS * CL_SIZE && lm->layers > 0;
}
__global__ void k_relayout_wo(uint16_t *dst, const uint16_t *wo, int H, int qd) {            /* dst[g][r][c] = wo[r][256 g + c] */
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)H * qd) return;
    const int c = (int)(i % 256), r = (int)((i / 256) % H), g = (int)(i / ((size_t)256 * H));
    dst[i] = wo[(size_t)r * qd + g * 256 + c];
}
__global__ void k_relayout_down(uint16_t *dst, const uint16_t *down, int H, int I) {         /* dst[c8][r][j] = down[r][8 c8 + j] */
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)H * I) return;
    const int j = (int)(i % 8), r = (int)((i / 8) % H), c8 = (int)(i / ((size_t)8 * H));
    dst[i] = down[(size_t)r * I + c8 * 8 + j];
}
int qtc_cluster_prepare(qtc_ctx *c, LmDev *d) {
    for (int l = 0; l < d->layers; l++) { d->layer[l].wo_t = NULL; d->layer[l].down_t = NULL; }
    if (!cluster_geometry(d)) return 0;
    const size_t H = d->hidden, qd = (size_t)d->heads * d->head_dim, I = d->inter;
    for (int l = 0; l < d->layers; l++) {
        uint16_t *wo_t = (uint16_t *)qtc_dmalloc(c, H * qd * 2), *down_t = (uint16_t *)qtc_dmalloc(c, H * I * 2);
        if (!wo_t || !down_t) return -1;
        k_relayout_wo<<<(unsigned)((H * qd + 255) / 256), 256>>>(wo_t, d->layer[l].wo, (int)H, (int)qd);
        k_relayout_down<<<(unsigned)((H * I + 255) / 256), 256>>>(down_t, d->layer[l].down, (int)H, (int)I);
        QTC_CUDA(cudaGetLastError());
        d->layer[l].wo_t = wo_t; d->layer[l].down_t = down_t;
    }
    QTC_CUDA(cudaDeviceSynchronize());
    return 0;
}
/* ptab[g-1][v][r] = sum_c in_w[r][c] * emb_g[v][c] + in_b[r] (or the embedding row itself, zero-padded, without an input projection):
 * what the sub-talker's position g >= 1 feeds its first layer when code v was sampled at position g - 1.  A block takes 8 codes: their
 * embedding rows sit in shared memory as fp32, each warp streams weight rows once for all 8; the sum order is the decode kernels'
 * (lane's 16-byte pieces lane + 32 i in turn, then the butterfly), so the table holds what the in-kernel projection would produce. */
__global__ void __launch_bounds__(256) k_proj_table(float *out, const uint16_t *emb, const uint16_t *in_w, const float *in_b, int vocab, int Hin, int h) {
    extern __shared__ __align__(16) float xs8[];
    const int v0 = blockIdx.x * 8, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 8 * Hin; i += 256) { const int k = i / Hin, c = i - k * Hin; xs8[i] = v0 + k < vocab ? bf16_to_f32(emb[(size_t)(v0 + k) * Hin + c]) : 0.0f; }
    __syncthreads();
    if (!in_w) {
        for (int i = tid; i < 8 * h; i += 256) { const int k = i / h, r = i - k * h; if (v0 + k < vocab) out[(size_t)(v0 + k) * h + r] = r < Hin ? xs8[k * Hin + r] : 0.0f; }
        return;
    }
    const int n8 = Hin >> 3;
    for (int r = warp; r < h; r += 8) {
        const uint4 *w4 = reinterpret_cast<const uint4 *>(in_w + (size_t)r * Hin);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = lane; i < n8; i += 32) {
            const uint4 wv = ldg_stream(w4 + i);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const float4 *x4 = reinterpret_cast<const float4 *>(xs8 + k * Hin);
                acc[k] = dot8f(wv, x4[2 * i], x4[2 * i + 1], acc[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < 8; k++) acc[k] = warp_sum(acc[k]);
        if (lane < 8 && v0 + lane < vocab) {
            float vsel = acc[0];
#pragma unroll
            for (int k = 1; k < 8; k++) if (lane == k) vsel = acc[k];
            out[(size_t)(v0 + lane) * h + r] = (in_b ? in_b[r] : 0.0f) + vsel;
        }
    }
}
int qtc_build_proj_tables(qtc_ctx *c) {
    for (int g = 0; g < QTC_MAX_GROUPS - 1; g++) c->sub_ptab[g] = NULL;
    const LmDev *lm = &c->sub;
    if (!cluster_geometry(lm) || c->G < 2 || c->H % 8) return 0;
    const int Hin = c->H, h = lm->hidden;
    if ((size_t)8 * Hin * 4 > 160 * 1024) return 0;
    QTC_CUDA(cudaFuncSetAttribute(k_proj_table, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * Hin * 4));
    for (int g = 1; g < c->G; g++) {
        const int vocab = g == 1 ? c->talker.vocab : lm->vocab;
        const uint16_t *emb = g == 1 ? c->codec_embedding : c->sub_emb[g - 2];
        float *t = (float *)qtc_dmalloc(c, (size_t)vocab * h * 4);
        if (!t) return -1;
        k_proj_table<<<(vocab + 7) / 8, 256, (size_t)8 * Hin * 4>>>(t, emb, c->sub_in_w, c->sub_in_b, vocab, Hin, h);
        QTC_CUDA(cudaGetLastError());
        c->sub_ptab[g - 1] = t;
    }
    QTC_CUDA(cudaDeviceSynchronize());
    return 0;
}

static int g_cluster_attr = 0, g_cluster_max_smem = 0, g_cluster_fit = -1;
static int cluster_setup(qtc_ctx *c) {
    if (g_cluster_attr) return 0;
    cudaFuncAttributes fa;
    QTC_CUDA(cudaFuncGetAttributes(&fa, k_lm_cluster));
    g_cluster_max_smem = 227 * 1024 - (int)fa.sharedSizeBytes;
    QTC_CUDA(cudaFuncSetAttribute(k_lm_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, g_cluster_max_smem));
    QTC_CUDA(cudaFuncSetAttribute(k_lm_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    /* how many 8-CTA clusters the device holds at once (B200, one CTA per SM: 15) */
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CL_MAX_CLUSTERS * CL_SIZE); cfg.blockDim = dim3(CL_THREADS); cfg.dynamicSmemBytes = (size_t)cl_smem_floats(4096, 2048, CL_ATT_CAP) * 4;
    cudaLaunchAttribute at[1];
    memset(at, 0, sizeof(at));
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CL_SIZE; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int ncl = 0;
    if (cudaOccupancyMaxActiveClusters(&ncl, k_lm_cluster, &cfg) != cudaSuccess) { cudaGetLastError(); ncl = 0; }
    g_cluster_fit = ncl;
    { const char *e = getenv("QTC_CLUSTERS"); if (e && atoi(e) > 0 && atoi(e) < g_cluster_fit) g_cluster_fit = atoi(e); }   /* experiment: fewer clusters */
    g_cluster_attr = 1;
    (void)c;
    return 0;
}
static bool cluster_fits(qtc_ctx *c, const LmDev *lm) {
    if (c->mode < 2 || c->arith != QTC_FAST || !cluster_geometry(lm) || !lm->layer[0].wo_t) return false;
    if (cluster_setup(c)) return false;
    /* clusters needed: one per KV head in phase A, and at most 8 chunks of the MLP's inner dimension per CTA in phase B */
    const int need = (lm->inter / 8 + 8 * CL_SIZE - 1) / (8 * CL_SIZE);
    return g_cluster_fit >= lm->kv_heads && g_cluster_fit >= need;
}
/* debug: how many 16-CTA clusters fit the device (-1 before the first mode-2 step) and whether a mode-2 barrier wait ever gave up */
extern "C" int qwen_tts_cuda_cluster_status(qtc_ctx_t *c, int *capacity, int *gave_up) {
    if (!c) return -1;
    if (capacity) *capacity = g_cluster_fit;
    int e = 0;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    if (c->cl_err) QTC_CUDA(cudaMemcpy(&e, c->cl_err, 4, cudaMemcpyDeviceToHost));
    if (gave_up) *gave_up = e;
    return 0;
}
static int cluster_step(qtc_ctx *c, const LmDev *lm, LmBuffers *b, const int *pos_ptr, int pos_add,
                        const uint16_t *head_w, int head_rows, float *logits, GenState *st, const SubProg *prog) {
    ClusterArgs a;
    memset(&a, 0, sizeof(a));
    a.lm = lm == &c->talker ? c->talker_dev : c->sub_dev;
    a.x = b->x; a.xacc = b->xacc; a.phase = b->cl_phase;
    a.kc = b->k; a.vc = b->v; a.table = b->table; a.cos = b->cos; a.sin = b->sin; a.kv_max = b->kv_max;
    a.pos_ptr = pos_ptr; a.pos_add = pos_add;
    a.head_w = head_w; a.head_rows = head_rows; a.logits = logits;
    a.bar = c->bar + 4;                              /* its own counter: this grid has 120 CTAs, k_lm_fused 148 */
    a.st = st; a.err = c->cl_err;
    a.prof = (c->prof_sub ? lm == &c->sub : lm == &c->talker) ? c->prof : NULL; a.prof_layer = c->prof_layer;
    { static int fl = -1; if (fl < 0) { const char *e = getenv("QTC_FUSED_FLAGS"); fl = e ? atoi(e) : 0; } a.flags = fl; }
    {   /* idle-cluster L2 prefetch of the MLP weights (measured on B200, scripts/probe/pf_sweep.sh): the sub-talker takes all of it
         * unpaced (1837 -> 1784 us per frame); the talker's 75 MB per layer must not crowd out the attention block's own loads:
         * three quarters of every slice, 2 us between a warp's requests (re-swept on the closing kernel, scripts/gpu_r2_pftk.sh, us per step:
         * 40 % 924, 50 % 900, 60 % 886, 70 % 882, 80 % 880, 90 % 900, 100 % 953; 80 % at 3 us: 1174; round 2's first sweep had 50 % ahead) */
        static int fr = -2, gp = -1, tfr = -1, tgp = -1, sfr = -1, sgp = -1;
        if (fr == -2) {
            const char *e = getenv("QTC_PF_FRAC"); fr = e ? atoi(e) : -1; e = getenv("QTC_PF_GAP"); gp = e ? atoi(e) : -1;
            e = getenv("QTC_PF_TK_FRAC"); tfr = e ? atoi(e) : -1; e = getenv("QTC_PF_TK_GAP"); tgp = e ? atoi(e) : -1;     /* the talker's alone */
            e = getenv("QTC_PF_ST_FRAC"); sfr = e ? atoi(e) : -1; e = getenv("QTC_PF_ST_GAP"); sgp = e ? atoi(e) : -1;     /* the sub-talker's alone */
        }
        const bool tk = lm == &c->talker;
        a.pf_frac = tk && tfr >= 0 ? tfr : !tk && sfr >= 0 ? sfr : fr >= 0 ? fr : (tk ? 75 : 100);
        a.pf_gap = tk && tgp >= 0 ? tgp : !tk && sgp >= 0 ? sgp : gp >= 0 ? gp : (tk ? 2000 : 0);
    }
    int vec = lm->hidden;
    SubProg sp;
    memset(&sp, 0, sizeof(sp));
    if (prog) {
        sp = *prog;
        if (2 * sp.vocab > vec) vec = 2 * sp.vocab;
        if (sp.Hin > vec) vec = sp.Hin;
    }
    a.xs_floats = (vec + 31) & ~31;
    /* K/V stage: positions this sequence can reach split 8 ways (the sub-talker: G + 2; the talker: the pages it holds), in steps of 16 */
    int reach = b == &c->tk ? c->kv.used[0] * QTC_KV_PAGE_POS : b->kv_max;
    if (b == &c->tk && c->tk_reach_hint > 0 && c->tk_reach_hint < reach) reach = c->tk_reach_hint;
    a.att_cap = (((reach + CL_SIZE - 1) / CL_SIZE + 15) / 16) * 16;
    if (a.att_cap > CL_ATT_CAP) a.att_cap = CL_ATT_CAP;
    { static int cap = -1; if (cap < 0) { const char *e = getenv("QTC_ATT_CAP"); cap = e ? atoi(e) : 0; } if (cap > 0 && b == &c->tk) a.att_cap = cap; }   /* A/B: the stage's size is shared memory taken from L1 */
    if (a.att_cap < 16) a.att_cap = 16;
    if (b != &c->tk && reach <= 64 && a.att_cap < 32) a.att_cap = 32;      /* the sub-talker: ranks 0 / 1 stage every earlier row (one warp per head) */
    while (prog && (size_t)(512 + 2 * lm->hidden + 2 * a.att_cap * 128) * 4 < sizeof(SampleSmem)) a.att_cap += 16;   /* the sampler's workspace overlays aux.. */
    const size_t smem = (size_t)cl_smem_floats(a.xs_floats, lm->hidden, a.att_cap) * 4;
    if (smem > (size_t)g_cluster_max_smem) { qtc_set_error("cluster step: %zu bytes of shared memory needed, %d available", smem, g_cluster_max_smem); return -1; }
    cudaLaunchConfig_t cfg = {};
    const int ncl = g_cluster_fit < CL_MAX_CLUSTERS ? g_cluster_fit : CL_MAX_CLUSTERS;
    cfg.gridDim = dim3(ncl * CL_SIZE); cfg.blockDim = dim3(CL_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = c->stream;
    cudaLaunchAttribute at[2];
    memset(at, 0, sizeof(at));
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CL_SIZE; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeCooperative; at[1].val.cooperative = 1;
    /* cooperative as well where the driver takes both attributes (co-residency then is checked, not assumed); the grid is
     * 128 CTAs at one per SM on an otherwise idle stream-ordered device either way */
    static int coop = -1;
    if (coop < 0) { const char *e = getenv("QTC_CL_COOP"); coop = e ? atoi(e) : 1; }      /* 0: cluster attribute only (ncu could not replay the cooperative + cluster launch) */
    cfg.attrs = at; cfg.numAttrs = coop ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_lm_cluster, a, sp);
    if (e != cudaSuccess && coop) {
        cudaGetLastError();
        coop = 0; cfg.numAttrs = 1;
        e = cudaLaunchKernelEx(&cfg, k_lm_cluster, a, sp);
    }
    if (e != cudaSuccess) { qtc_set_error("cluster step launch -> %s", cudaGetErrorString(e)); return -1; }
    c->launches++;
    return 0;
}

/* mode 1: the whole step (+ optional head GEMV, + pos bump) as one persistent cooperative kernel */
static int g_fused_attr = 0, g_fused_max_smem = 0;
static bool fused_ok(const qtc_ctx *c, const LmDev *lm) {
    return c->mode >= 1 && c->arith == QTC_FAST && lm->head_dim <= FUSED_THREADS && lm->head_dim % 2 == 0 &&
           lm->hidden % 8 == 0 && lm->inter % 8 == 0 && (lm->heads * lm->head_dim) % 8 == 0 && lm->heads <= c->num_sms &&
           lm->heads * (lm->head_dim + 4) <= 2 * ATT_STAGE_MAX * lm->head_dim;
}
static int fused_step(qtc_ctx *c, const LmDev *lm, LmBuffers *b, const int *pos_ptr, int pos_add,
                      const uint16_t *head_w, int head_rows, float *logits, GenState *st, const SubProg *prog = NULL) {
    if (cluster_fits(c, lm)) return cluster_step(c, lm, b, pos_ptr, pos_add, head_w, head_rows, logits, st, prog);
    if (!g_fused_attr) {
        cudaFuncAttributes fa;
        QTC_CUDA(cudaFuncGetAttributes(&fa, k_lm_fused));
        g_fused_max_smem = 227 * 1024 - (int)fa.sharedSizeBytes;         /* opt-in limit includes the static part */
        QTC_CUDA(cudaFuncSetAttribute(k_lm_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, g_fused_max_smem));
        g_fused_attr = 1;
    }
    FusedArgs a;
    memset(&a, 0, sizeof(a));
    a.lm = lm == &c->talker ? c->talker_dev : c->sub_dev;
    a.x = b->x; a.q = b->q; a.kk = b->kk; a.vv = b->vv; a.att = b->att; a.g = b->g;
    a.kc = b->k; a.vc = b->v; a.table = b->table; a.cos = b->cos; a.sin = b->sin; a.kv_max = b->kv_max;
    a.pos_ptr = pos_ptr; a.pos_add = pos_add;
    a.head_w = head_w; a.head_rows = head_rows; a.logits = logits;
    a.bar = c->bar; a.head_cnt = c->head_cnt; a.part = c->attn_part; a.st = st;
    a.prof = (c->prof_sub ? lm == &c->sub : lm == &c->talker) ? c->prof : NULL; a.prof_layer = c->prof_layer;
    { static int fl = -1; if (fl < 0) { const char *e = getenv("QTC_FUSED_FLAGS"); fl = e ? atoi(e) : 0; } a.flags = fl; }
    const int qd = lm->heads * lm->head_dim;
    int vec = lm->hidden > lm->inter ? lm->hidden : lm->inter;
    if (qd > vec) vec = qd;
    SubProg sp;
    memset(&sp, 0, sizeof(sp));
    if (prog) {
        sp = *prog;
        if (2 * sp.vocab > vec) vec = 2 * sp.vocab;       /* the in-kernel sampler works in the staged-vector region */
        if (sp.Hin > vec) vec = sp.Hin;
    }
    /* score buffer: a staged chunk (<= 64) or, for long contexts, ceil(n / splits) positions */
    int sc = (b->kv_max + c->num_sms / lm->heads - 1) / (c->num_sms / lm->heads > 0 ? c->num_sms / lm->heads : 1);
    if (sc < 64) sc = 64;
    int attn = 2 * lm->head_dim + FUSED_THREADS + sc + 64 + 3 * lm->heads * FUSED_MAX_SPLIT;
    if (attn < 2 * FUSED_WARPS + FUSED_WARPS * lm->head_dim) attn = 2 * FUSED_WARPS + FUSED_WARPS * lm->head_dim;   /* cta_attention_warps: 16 (max, sum, out[hd]) triples */
    a.xs_floats = (vec + 3) & ~3;
    a.att_floats = (attn + 3) & ~3;
    const size_t smem = (size_t)(a.xs_floats + a.att_floats + 2 * ATT_STAGE_MAX * lm->head_dim) * 4;
    if (smem > (size_t)g_fused_max_smem) { qtc_set_error("fused step: %zu bytes of shared memory needed, %d available", smem, g_fused_max_smem); return -1; }
    void *args[] = {&a, &sp};
    static int grid_env = -1;            /* experiment: QTC_FUSED_CTAS = CTAs of the mode-1 grid (default: one per SM) */
    if (grid_env < 0) { const char *e = getenv("QTC_FUSED_CTAS"); grid_env = e ? atoi(e) : 0; }
    const int grid = grid_env > 0 && grid_env <= c->num_sms ? grid_env : c->num_sms;
    QTC_CUDA(cudaLaunchCooperativeKernel((void *)k_lm_fused, dim3(grid), dim3(FUSED_BLOCK), args, smem, c->stream));
    c->launches++;
    return 0;
}

/* One token through every layer + final norm.  b->x holds the input on entry and the post-final-norm
 * hidden on exit (ctx->tk_x semantics).  Position = *pos_ptr + pos_add. */
int qtc_lm_step(qtc_ctx *c, const LmDev *lm, LmBuffers *b, const int *pos_ptr, int pos_add) {
    const int H = lm->hidden, hd = lm->head_dim, qd = lm->heads * hd, kvd = lm->kv_heads * hd;
    if (fused_ok(c, lm)) return fused_step(c, lm, b, pos_ptr, pos_add, NULL, 0, NULL, NULL);
    if (set_smem_attrs()) return -1;
    for (int l = 0; l < lm->layers; l++) {
        const LmLayerDev *w = &lm->layer[l];
        if (qtc_rmsnorm(c, b->xn, b->x, w->in_norm, H, lm->eps)) return -1;
        GemvSeg qkv[3] = {{w->wq, b->q, qd}, {w->wk, b->kk, kvd}, {w->wv, b->vv, kvd}};
        if (qtc_gemv(c, qkv, 3, b->xn, H, EPI_STORE, NULL)) return -1;
        AttnArgs a;
        a.q = b->q; a.k = b->kk; a.v = b->vv; a.q_norm = w->q_norm; a.k_norm = w->k_norm;
        a.cos = b->cos; a.sin = b->sin;
        a.kc = b->k + (size_t)l * b->kv_max * kvd; a.vc = b->v + (size_t)l * b->kv_max * kvd; a.table = b->table;
        a.att = b->att; a.pos_ptr = pos_ptr; a.pos_add = pos_add;
        a.heads = lm->heads; a.kv_heads = lm->kv_heads; a.hd = hd; a.eps = lm->eps;
        a.scale = 1.0f / sqrtf((float)hd);
        const int cap = b == &c->tk ? c->kv.used[0] * QTC_KV_PAGE_POS : b->kv_max;      /* positions this sequence can reach */
        const size_t smem = (size_t)(2 * hd + cap + 32) * 4;
        if (smem > 160 * 1024) { qtc_set_error("per-op attention: %d positions do not fit its score buffer (use the fused mode)", cap); return -1; }
        if (c->arith == QTC_STRICT) k_attn_decode<QTC_STRICT><<<lm->heads, 128, smem, c->stream>>>(a);
        else k_attn_decode<QTC_FAST><<<lm->heads, 128, smem, c->stream>>>(a);
        LAUNCH_CHECK(c);
        GemvSeg o = {w->wo, b->x, H};
        if (qtc_gemv(c, &o, 1, b->att, qd, EPI_RESID, b->x)) return -1;
        if (qtc_rmsnorm(c, b->xn, b->x, w->post_norm, H, lm->eps)) return -1;
        if (qtc_gemv_swiglu(c, b->g, w->gate, w->up, b->xn, lm->inter, H)) return -1;
        GemvSeg d = {w->down, b->x, H};
        if (qtc_gemv(c, &d, 1, b->g, lm->inter, EPI_RESID, b->x)) return -1;
    }
    return qtc_rmsnorm(c, b->x, b->x, lm->final_norm, H, lm->eps);
}

__global__ void k_copy_f32(float *dst, const float *src, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}
__global__ void k_bump_pos(GenState *st) { if (!st->done) st->pos += 1; }

/* qwen_tts_talker_forward on device: embed -> all layers -> final norm (tk.x) -> codec head (tk.logits); pos++ */
int qtc_talker_forward_dev(qtc_ctx *c, const float *embed_dev) {
    const LmDev *lm = &c->talker;
    k_copy_f32<<<(lm->hidden + 255) / 256, 256, 0, c->stream>>>(c->tk.x, embed_dev, lm->hidden);
    LAUNCH_CHECK(c);
    if (fused_ok(c, lm)) return fused_step(c, lm, &c->tk, &c->state->pos, 0, c->codec_head, lm->vocab, c->tk.logits, c->state);
    if (qtc_lm_step(c, lm, &c->tk, &c->state->pos, 0)) return -1;
    GemvSeg h = {c->codec_head, c->tk.logits, lm->vocab};
    if (qtc_gemv(c, &h, 1, c->tk.x, lm->hidden, EPI_STORE, NULL)) return -1;
    k_bump_pos<<<1, 1, 0, c->stream>>>(c->state);
    LAUNCH_CHECK(c);
    return 0;
}

/* ST_PROJECT_INPUT, c/qwen_tts_talker.c:693-702: st.x = input_proj(src) (+bias), or copy / zero-pad */
static int sub_project(qtc_ctx *c, const float *src) {
    const int h = c->sub.hidden, H = c->H;
    if (c->sub_in_w) {
        GemvSeg s = {c->sub_in_w, c->st.x, h};
        return qtc_gemv(c, &s, 1, src, H, c->sub_in_b ? EPI_BIAS : EPI_STORE, c->sub_in_b);
    }
    k_copy_pad<<<(h + 255) / 256, 256, 0, c->stream>>>(c->st.x, src, H < h ? H : h, h);
    LAUNCH_CHECK(c);
    return 0;
}

static int sub_sample(qtc_ctx *c, int g, int *codes_out_dev, int reseed) {
    SampleArgs a;
    memset(&a, 0, sizeof(a));
    a.logits = c->st.logits;
    a.cfg.vocab = c->sub.vocab; a.cfg.top_k = c->params.subtalker_top_k; a.cfg.top_p = c->params.subtalker_top_p;
    a.cfg.temperature = c->params.subtalker_temperature; a.cfg.suppress_from = -1; a.cfg.eos_id = -1; a.cfg.rep_penalty = 1.0f;
    a.rng = c->st_rng; a.out_token = codes_out_dev + g; a.role = 1;
    a.reseed = reseed;
    { float f = (float)c->params.sample_seed; memcpy(&a.seed_bits, &f, 4); }   /* re-seeded every frame, :718 */
    a.scratch_f = c->sample_scratch_f; a.scratch_i = c->sample_scratch;
    k_sample<<<1, SAMPLE_THREADS, (size_t)a.cfg.vocab * 8, c->stream>>>(a);
    LAUNCH_CHECK(c);
    return 0;
}


/* the parallel sampler paths only (greedy, or top-k <= 1024 with top_p >= 1): the slow path uses one global scratch */
static bool sampler_fusable(const qtc_ctx *c) {
    const int k = c->params.subtalker_top_k;
    return c->params.subtalker_top_p >= 1.0f && k > 0 && k < c->sub.vocab && k <= SAMPLE_MAX_K;
}

/* qwen_tts_subtalker_generate on device.  first_code_ptr: device int (codes_out_dev[0] already holds it). */
int qtc_subtalker_dev(qtc_ctx *c, const float *hidden_dev, const int *first_code_ptr, int *codes_out_dev) {
    const LmDev *lm = &c->sub;
    const int G = c->G, H = c->H;
    if (fused_ok(c, lm) && sampler_fusable(c) && first_code_ptr == codes_out_dev && H % 4 == 0) {
        /* mode 1: the whole 15-codebook loop is ONE persistent kernel launch */
        SubProg sp;
        memset(&sp, 0, sizeof(sp));
        sp.G = G; sp.Hin = H; sp.hidden = hidden_dev; sp.emb0 = c->codec_embedding;
        for (int g = 0; g < G - 1; g++) { sp.emb[g] = c->sub_emb[g]; sp.heads[g] = c->sub_head[g]; }
        sp.in_w = c->sub_in_w; sp.in_b = c->sub_in_b; sp.vocab = lm->vocab; sp.codes = codes_out_dev;
        sp.cfg.vocab = lm->vocab; sp.cfg.top_k = c->params.subtalker_top_k; sp.cfg.top_p = c->params.subtalker_top_p;
        sp.cfg.temperature = c->params.subtalker_temperature; sp.cfg.suppress_from = -1; sp.cfg.eos_id = -1; sp.cfg.rep_penalty = 1.0f;
        { float f = (float)c->params.sample_seed; memcpy(&sp.seed_bits, &f, 4); }
        sp.scratch_f = c->sample_scratch_f; sp.scratch_i = c->sample_scratch; sp.vec = c->st_embed;
        for (int g = 0; g < G - 1; g++) sp.ptab[g] = c->sub_ptab[g];
        return fused_step(c, lm, &c->st, c->zero_pos, 0, NULL, lm->vocab, c->st.logits, NULL, &sp);
    }
    const int *g_zero_pos = c->zero_pos;          /* device int 0: the sub-talker's position base */
    if (sub_project(c, hidden_dev)) return -1;                                 /* pos 0: no logits */
    if (qtc_lm_step(c, lm, &c->st, g_zero_pos, 0)) return -1;
    for (int g = 1; g < G; g++) {
        const uint16_t *table = g == 1 ? c->codec_embedding : c->sub_emb[g - 2];
        const int *idx = g == 1 ? first_code_ptr : codes_out_dev + (g - 1);
        if (qtc_embed_row(c, c->st_embed, table, idx, 0, H)) return -1;
        if (sub_project(c, c->st_embed)) return -1;
        if (fused_ok(c, lm)) {
            if (fused_step(c, lm, &c->st, g_zero_pos, g, c->sub_head[g - 1], lm->vocab, c->st.logits, NULL)) return -1;
        } else {
            if (qtc_lm_step(c, lm, &c->st, g_zero_pos, g)) return -1;
            GemvSeg hd = {c->sub_head[g - 1], c->st.logits, lm->vocab};
            if (qtc_gemv(c, &hd, 1, c->st.x, lm->hidden, EPI_STORE, NULL)) return -1;
        }
        if (sub_sample(c, g, codes_out_dev, g == 1)) return -1;
    }
    return 0;
}

int qtc_sample_talker(qtc_ctx *c) {
    SampleArgs a;
    memset(&a, 0, sizeof(a));
    const LmDev *lm = &c->talker;
    a.logits = c->tk.logits;
    a.cfg.vocab = lm->vocab; a.cfg.top_k = c->params.top_k; a.cfg.top_p = c->params.top_p;
    a.cfg.temperature = c->params.temperature;
    a.cfg.suppress_from = lm->vocab - 1024; a.cfg.eos_id = c->host.codec_eos_id;
    a.cfg.rep_penalty = c->params.repetition_penalty;
    a.rng = &c->state->rng; a.hist_count = c->hist_count; a.st = c->state; a.codes = c->frame_codes; a.G = c->G; a.role = 0;
    a.scratch_f = c->sample_scratch_f; a.scratch_i = c->sample_scratch;
    if (set_smem_attrs()) return -1;
    k_sample<<<1, SAMPLE_THREADS, (size_t)a.cfg.vocab * 8, c->stream>>>(a);
    LAUNCH_CHECK(c);
    return 0;
}

/* ---- small launchers used by the host-side frame loop (qtc_generate.cu) ----- */
__global__ void k_add_row(float *out, const float *a, const uint16_t *table, int idx, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __fadd_rn(a[i], bf16_to_f32(table[(size_t)idx * n + i]));   /* kernel_add_inplace(dst, codec_emb) */
}
__global__ void k_set_int(int *p, int v) { *p = v; }

int qtc_copy(qtc_ctx *c, float *dst, const float *src, int n) {
    k_copy_f32<<<(n + 255) / 256, 256, 0, c->stream>>>(dst, src, n);
    LAUNCH_CHECK(c);
    return 0;
}
int qtc_silu(qtc_ctx *c, float *x, int n) {
    if (c->arith == QTC_STRICT) k_silu<QTC_STRICT><<<(n + 255) / 256, 256, 0, c->stream>>>(x, n);
    else k_silu<QTC_FAST><<<(n + 255) / 256, 256, 0, c->stream>>>(x, n);
    LAUNCH_CHECK(c);
    return 0;
}
int qtc_add_row(qtc_ctx *c, float *out, const float *a, const uint16_t *table, int idx, int n) {
    k_add_row<<<(n + 255) / 256, 256, 0, c->stream>>>(out, a, table, idx, n);
    LAUNCH_CHECK(c);
    return 0;
}
int qtc_set_int(qtc_ctx *c, int *p, int v) {
    k_set_int<<<1, 1, 0, c->stream>>>(p, v);
    LAUNCH_CHECK(c);
    return 0;
}
int qtc_bump_pos(qtc_ctx *c) {
    k_bump_pos<<<1, 1, 0, c->stream>>>(c->state);
    LAUNCH_CHECK(c);
    return 0;
}

/* c/qwen_tts.c:1302-1363 on device: sample -> sub-talker -> next embedding -> n_generated++ */
int qtc_frame_front(qtc_ctx *c) {
    if (qtc_sample_talker(c)) return -1;
    if (qtc_subtalker_dev(c, c->tk.x, c->frame_codes, c->frame_codes)) return -1;
    NextEmbedArgs a;
    memset(&a, 0, sizeof(a));
    a.out = c->next_embed; a.codec_embedding = c->codec_embedding;
    for (int g = 0; g < c->G - 1; g++) a.sub_emb[g] = c->sub_emb[g];
    a.frame_codes = c->frame_codes; a.all_codes = c->codes; a.trailing = c->trailing; a.tts_pad = c->tts_pad;
    a.st = c->state; a.H = c->H; a.G = c->G;
    k_next_embed<<<(c->H + 255) / 256, 256, 0, c->stream>>>(a);
    LAUNCH_CHECK(c);
    k_advance_step<<<1, 1, 0, c->stream>>>(c->state);
    LAUNCH_CHECK(c);
    return 0;
}

/* kernel-level sample for the test surface: logits already on device */
int qtc_sample_bare(qtc_ctx *c, const float *logits_dev, int vocab, int top_k, float top_p, float temperature,
                    uint32_t *rng_dev, int *out_dev) {
    if (set_smem_attrs()) return -1;
    SampleArgs a;
    memset(&a, 0, sizeof(a));
    a.logits = logits_dev;
    a.cfg.vocab = vocab; a.cfg.top_k = top_k; a.cfg.top_p = top_p; a.cfg.temperature = temperature;
    a.cfg.suppress_from = -1; a.cfg.eos_id = -1; a.cfg.rep_penalty = 1.0f;
    a.rng = rng_dev; a.out_token = out_dev; a.role = 1;
    a.scratch_f = c->sample_scratch_f; a.scratch_i = c->sample_scratch;
    k_sample<<<1, SAMPLE_THREADS, (size_t)vocab * 8, c->stream>>>(a);
    LAUNCH_CHECK(c);
    return 0;
}

#include "qtc_prefill.cuh"
#include "qtc_batch.cuh"

/* ---- allocation-free sampler entry point (no counterpart in the reference) ----
 * For callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
 * Same k_sample launch as qtc_sample_bare, but the slow paths' scratch (V floats + V ints: the top_p / literal-insertion
 * branches of block_sample_t) lives in a file-scope device array instead of two cudaMalloc'd buffers.  Every GPU that loads
 * the library holds QTC_NOALLOC_MAX_VOCAB * 8 = 128 KB of it until exit.  Inputs and outputs are DEVICE pointers of the
 * caller; the launch goes to the caller's stream and is not synchronised here.
 * The array is shared by every call and keeps the last call's contents.  Nothing in it has to start at zero: each branch
 * writes the entries it later reads (p[0..V) and ord[0..V), or val/idx[0..k)) before reading them, so no clear is issued.
 * Calls that overlap on the GPU (two streams, two host threads) race on the array: serialise them.
 * In host code g_noalloc_scratch is a stand-in symbol, not a device address: the address comes from cudaGetSymbolAddress,
 * taken once per device with that device current. */
#ifndef QTC_NOALLOC_MAX_VOCAB
#define QTC_NOALLOC_MAX_VOCAB 16384        /* also in include/qwen_tts_cuda.h */
#endif
#define QTC_NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) unsigned char g_noalloc_scratch[(size_t)QTC_NOALLOC_MAX_VOCAB * 8];
static void *g_noalloc_dev_ptr[QTC_NOALLOC_MAX_DEVICES];

extern "C" int qwen_tts_cuda_sample_top_k_noalloc(const float *logits_dev, int vocab, int top_k, float top_p, float temperature,
                                                  float *rng_state_dev, int *token_dev, void *cuda_stream) {
    if (!logits_dev || !rng_state_dev || !token_dev) { qtc_set_error("sample_top_k_noalloc: NULL device pointer"); return -1; }
    if (vocab <= 0 || vocab > QTC_NOALLOC_MAX_VOCAB) {
        qtc_set_error("sample_top_k_noalloc: vocab %d outside (0, %d] - the static scratch holds %d entries", vocab, QTC_NOALLOC_MAX_VOCAB,
                      QTC_NOALLOC_MAX_VOCAB);
        return -1;
    }
    int dev = -1;
    QTC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= QTC_NOALLOC_MAX_DEVICES) { qtc_set_error("sample_top_k_noalloc: device %d out of range", dev); return -1; }
    if (!g_noalloc_dev_ptr[dev]) {                      /* set-up, once per device, with that device current */
        void *p = NULL;
        QTC_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) { qtc_set_error("sample_top_k_noalloc: cudaGetSymbolAddress returned NULL"); return -1; }
        QTC_CUDA(cudaFuncSetAttribute(k_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
        g_noalloc_dev_ptr[dev] = p;
    }
    SampleArgs a;
    memset(&a, 0, sizeof(a));
    a.logits = logits_dev;
    a.cfg.vocab = vocab; a.cfg.top_k = top_k; a.cfg.top_p = top_p; a.cfg.temperature = temperature;
    a.cfg.suppress_from = -1; a.cfg.eos_id = -1; a.cfg.rep_penalty = 1.0f;
    a.rng = (uint32_t *)rng_state_dev; a.out_token = token_dev; a.role = 1;
    a.scratch_f = (float *)g_noalloc_dev_ptr[dev];
    a.scratch_i = (int *)((unsigned char *)g_noalloc_dev_ptr[dev] + (size_t)QTC_NOALLOC_MAX_VOCAB * 4);
    k_sample<<<1, SAMPLE_THREADS, (size_t)vocab * 8, (cudaStream_t)cuda_stream>>>(a);
    QTC_CUDA(cudaGetLastError());
    return 0;
}
