/*
 * qtc_fused.cuh - one decoder-stack step as ONE persistent cooperative kernel
 * (mode 1): talker_attention_single x layers + final norm + head
 * (c/qwen_tts_talker.c:119-248, 478-533; ST_FORWARD :638-691).
 *
 * B200 mapping
 *   - one CTA of 512 threads per SM (148), cooperative launch, grid-wide
 *     barriers on an L2 counter between the five phases of a layer; no kernel
 *     boundary inside the step, so nothing drains and refills 197 times.
 *   - weights are never staged by a later phase's critical path: at the start
 *     of phase p every CTA issues cp.async.bulk.prefetch.L2 for ITS rows of
 *     phase p+2, so HBM keeps streaming through barriers and through the
 *     weight-less attention phase; the phase itself then reads L2-resident rows
 *     with 128-bit loads that bypass L1.
 *   - rows of each GEMV are split contiguously over CTAs (bytes per SM equal to
 *     within one row), a warp owns whole rows (two in flight), fp32 FMA,
 *     warp-shuffle reduction; RMSNorm is recomputed per CTA into shared memory
 *     (8 KB of L2 reads) instead of costing a barrier.
 *   - GQA decode attention is split over heads x up to 8 KV ranges; the last
 *     CTA to finish a head combines the partial (max, sum, output) triples.
 * Activations crossing CTAs are read with ld.global.cg (L1 is not coherent).
 */
#pragma once
#include "qtc_device.cuh"
#include "qtc_internal.h"

#define FUSED_THREADS 512
#define FUSED_WARPS 16
#define FUSED_MAX_SPLIT 16

struct FusedArgs {
    const LmDev *lm;               /* device copy of the stack descriptor */
    float *x;                      /* [H] residual stream: input on entry, post-final-norm hidden on exit */
    float *q, *kk, *vv, *att, *g;  /* cross-CTA scratch in HBM/L2 */
    float *kc, *vc;                /* [layers, kv_max, kvd]: per layer a pool slice (paged, kv_max = pool positions) or a contiguous cache */
    const int *table;              /* page table of the utterance (talker), NULL = contiguous (sub-talker) */
    const float *cos, *sin;        /* [kv_max, hd] */
    int kv_max;
    const int *pos_ptr; int pos_add;
    const uint16_t *head_w; int head_rows; float *logits;   /* optional head GEMV on the normed hidden */
    unsigned long long *bar;       /* grid barrier counter (monotonic) */
    unsigned *head_cnt;            /* [heads] arrivals of the attention splits */
    float *part;                   /* [heads, FUSED_MAX_SPLIT, hd + 2] */
    GenState *st;                  /* talker: pos += 1 at the end (NULL for the sub-talker) */
    int xs_floats;                 /* size of the staged-vector region of shared memory */
    int att_floats;                /* attention scratch after it; the KV stage [2][64][hd] follows */
    int flags;                     /* experiments (env QTC_FUSED_FLAGS) */
    int ring_off;                  /* byte offset of the weight ring in dynamic shared memory (128-aligned) */
    int nst;                       /* ring stages (k_lm_ring) */
    int att_cap;                   /* positions the shared-memory KV stage holds */
    int kv_floats;                 /* size of the KV stage / sampler region */
    long long *prof; int prof_layer; /* optional: clock64 stamps of one layer, 16 slots per CTA (debug / profiles) */
};

/* The sub-talker's whole code-group loop (qwen_tts_subtalker_generate, c/qwen_tts_talker.c:539-736) as a
 * program for the same kernel: G stack steps, each preceded by an embedding gather + input projection and
 * followed (g >= 1) by its lm_head and an in-kernel sample.  G == 0: a single plain step (talker). */
struct SubProg {
    int G;                                  /* num_code_groups, 0 = not a sub-talker launch */
    int Hin;                                /* talker hidden: width of the gathered embedding rows */
    const float *hidden;                    /* talker hidden state [Hin] (position 0 input) */
    const uint16_t *emb0;                   /* talker codec embedding (input of position 1) */
    const uint16_t *emb[QTC_MAX_GROUPS - 1];/* sub-talker embeddings (input of position g uses emb[g-2]) */
    const uint16_t *in_w; const float *in_b;/* small_to_mtp_projection [h, Hin] (+bias) or NULL: copy / zero-pad */
    const uint16_t *heads[QTC_MAX_GROUPS - 1];
    int vocab;
    int *codes;                             /* [G] device: codes[0] = first code on entry, the rest on exit */
    SampleCfg cfg;
    uint32_t seed_bits;                     /* the RNG is re-seeded every frame (c/qwen_tts_talker.c:718) */
    float *scratch_f; int *scratch_i;       /* slow-path sampler scratch */
    float *vec;                             /* [Hin] cross-CTA scratch for the gathered / projected input */
    const float *ptab[QTC_MAX_GROUPS - 1];  /* mode 2: ptab[g-1][code] = the projected input of position g >= 1, fp32 [vocab, h], built at upload (NULL: project in the kernel) */
};

/* the 512 compute threads synchronise on named barrier 1; the producer warp (threads 512..543) never joins */
#define NB_COMPUTE 1
__device__ __forceinline__ void compute_sync() { asm volatile("bar.sync %0, %1;" ::"n"(NB_COMPUTE), "n"(FUSED_THREADS) : "memory"); }
__device__ __forceinline__ float csum(float v, float *red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    compute_sync();
    if (lane == 0) red[w] = v;
    compute_sync();
    return warp_sum(lane < FUSED_WARPS ? red[lane] : 0.0f);
}
__device__ __forceinline__ float cmax(float v, float *red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_max(v);
    compute_sync();
    if (lane == 0) red[w] = v;
    compute_sync();
    return warp_max(lane < FUSED_WARPS ? red[lane] : -3.402823466e+38f);
}

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void prefetch_l2_bulk(const void *p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

/* Grid barrier on a monotonic L2 counter.  Arrival is a release-reduction (orders this CTA's earlier writes,
 * which bar.sync made visible to thread 0, without a full fence and without waiting for loads in flight);
 * the wait is an acquire-load spin.  `target` lives in thread 0. */
__device__ __forceinline__ void red_release_inc(unsigned long long *p) {
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(1ULL) : "memory");
}
__device__ __forceinline__ void grid_barrier(unsigned long long *bar, unsigned long long &target, unsigned nb) {
    compute_sync();
    if (threadIdx.x == 0) {
        target += nb;
        red_release_inc(bar);
        while (ld_acquire_u64(bar) < target) { }
    }
    compute_sync();
}

/* the same barrier in two halves: loads that need no result of this phase (the next phase's first weight
 * trip) are issued after the CTA has announced its arrival and fly while the others are still coming */
__device__ __forceinline__ void barrier_arrive(unsigned long long *bar, unsigned long long &target, unsigned nb) {
    compute_sync();
    if (threadIdx.x == 0) {
        target += nb;
        red_release_inc(bar);
    }
}
__device__ __forceinline__ void barrier_wait(unsigned long long *bar, const unsigned long long &target) {
    if (threadIdx.x == 0) {
        while (ld_acquire_u64(bar) < target) { }
    }
    compute_sync();
}

__device__ __forceinline__ void row_range(int rows, int cta, int ncta, int &r0, int &r1) {
    /* rows * (cta + 1) < 2^32 for every matrix of these models (rows < 2^24, <= 256 CTAs): 32-bit unsigned division.  The s64
     * division subroutine (~100 instructions, 30 call sites after inlining, every warp, every phase) was ~0.5 us of pure
     * instruction issue per phase. */
    if (rows < (1 << 24) && ncta <= 256) {
        r0 = (int)(((unsigned)rows * (unsigned)cta) / (unsigned)ncta);
        r1 = (int)(((unsigned)rows * (unsigned)(cta + 1)) / (unsigned)ncta);
        return;
    }
    r0 = (int)(((long long)rows * cta) / ncta);
    r1 = (int)(((long long)rows * (cta + 1)) / ncta);
}

/* stage a cross-CTA vector into shared memory (L2 reads) */
__device__ __forceinline__ void stage_vec(float *xs, const float *src, int n) {
    for (int i = threadIdx.x * 4; i < n; i += FUSED_THREADS * 4)
        *reinterpret_cast<float4 *>(xs + i) = __ldcg(reinterpret_cast<const float4 *>(src + i));
    compute_sync();
}

/* xs = rmsnorm(src) * w, computed by every CTA for itself */
__device__ __forceinline__ void stage_rmsnorm(float *xs, const float *src, const float *w, int n, float eps, float *red) {
    if (n <= FUSED_THREADS * 8) {
        /* one pass: values and norm weights stay in registers across the reduction */
        const int i0 = threadIdx.x * 4, i1 = i0 + FUSED_THREADS * 4;
        float4 v0 = make_float4(0, 0, 0, 0), v1 = v0, w0 = v0, w1 = v0;
        if (i0 < n) { v0 = __ldcg(reinterpret_cast<const float4 *>(src + i0)); w0 = *reinterpret_cast<const float4 *>(w + i0); }
        if (i1 < n) { v1 = __ldcg(reinterpret_cast<const float4 *>(src + i1)); w1 = *reinterpret_cast<const float4 *>(w + i1); }
        float ss = v0.x * v0.x;
        ss = fmaf(v0.y, v0.y, ss); ss = fmaf(v0.z, v0.z, ss); ss = fmaf(v0.w, v0.w, ss);
        ss = fmaf(v1.x, v1.x, ss); ss = fmaf(v1.y, v1.y, ss); ss = fmaf(v1.z, v1.z, ss); ss = fmaf(v1.w, v1.w, ss);
        ss = csum(ss, red);
        const float inv = 1.0f / sqrtf(ss / (float)n + eps);
        if (i0 < n) *reinterpret_cast<float4 *>(xs + i0) = make_float4(v0.x * inv * w0.x, v0.y * inv * w0.y, v0.z * inv * w0.z, v0.w * inv * w0.w);
        if (i1 < n) *reinterpret_cast<float4 *>(xs + i1) = make_float4(v1.x * inv * w1.x, v1.y * inv * w1.y, v1.z * inv * w1.z, v1.w * inv * w1.w);
        compute_sync();
        return;
    }
    float ss = 0.0f;
    for (int i = threadIdx.x * 4; i < n; i += FUSED_THREADS * 4) {
        const float4 v = __ldcg(reinterpret_cast<const float4 *>(src + i));
        *reinterpret_cast<float4 *>(xs + i) = v;
        ss = fmaf(v.x, v.x, ss); ss = fmaf(v.y, v.y, ss); ss = fmaf(v.z, v.z, ss); ss = fmaf(v.w, v.w, ss);
    }
    ss = csum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)n + eps);
    for (int i = threadIdx.x; i < n; i += FUSED_THREADS) xs[i] = xs[i] * inv * w[i];
    compute_sync();
}

__device__ __forceinline__ float dot8f(const uint4 w, const float4 x0, const float4 x1, float acc) {
    acc = fmaf(bf16_lo(w.x), x0.x, acc); acc = fmaf(bf16_hi(w.x), x0.y, acc);
    acc = fmaf(bf16_lo(w.y), x0.z, acc); acc = fmaf(bf16_hi(w.y), x0.w, acc);
    acc = fmaf(bf16_lo(w.z), x1.x, acc); acc = fmaf(bf16_hi(w.z), x1.y, acc);
    acc = fmaf(bf16_lo(w.w), x1.z, acc); acc = fmaf(bf16_hi(w.w), x1.w, acc);
    return acc;
}

/* two rows at once, lanes stride the 16-byte pieces, 8 loads in flight per lane per trip */
__device__ __forceinline__ void warp_dot2(const uint16_t *ra, const uint16_t *rb, const float *xs, int cols, float &oa, float &ob) {
    const int lane = threadIdx.x & 31, n8 = cols >> 3;
    const uint4 *w0 = reinterpret_cast<const uint4 *>(ra), *w1 = reinterpret_cast<const uint4 *>(rb);
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    float a = 0.0f, b = 0.0f;
    int i = lane;
    for (; i + 96 < n8; i += 128) {
        const uint4 a0 = ldg_stream(w0 + i), a1 = ldg_stream(w0 + i + 32), a2 = ldg_stream(w0 + i + 64), a3 = ldg_stream(w0 + i + 96);
        const uint4 b0 = ldg_stream(w1 + i), b1 = ldg_stream(w1 + i + 32), b2 = ldg_stream(w1 + i + 64), b3 = ldg_stream(w1 + i + 96);
        a = dot8f(a0, x4[2 * i], x4[2 * i + 1], a);               b = dot8f(b0, x4[2 * i], x4[2 * i + 1], b);
        a = dot8f(a1, x4[2 * i + 64], x4[2 * i + 65], a);         b = dot8f(b1, x4[2 * i + 64], x4[2 * i + 65], b);
        a = dot8f(a2, x4[2 * i + 128], x4[2 * i + 129], a);       b = dot8f(b2, x4[2 * i + 128], x4[2 * i + 129], b);
        a = dot8f(a3, x4[2 * i + 192], x4[2 * i + 193], a);       b = dot8f(b3, x4[2 * i + 192], x4[2 * i + 193], b);
    }
    for (; i < n8; i += 32) {
        const uint4 a0 = ldg_stream(w0 + i), b0 = ldg_stream(w1 + i);
        a = dot8f(a0, x4[2 * i], x4[2 * i + 1], a);
        b = dot8f(b0, x4[2 * i], x4[2 * i + 1], b);
    }
    oa = warp_sum(a);
    ob = warp_sum(b);
}

struct Seg3 { const uint16_t *w[3]; float *out[3]; int rows[3]; int n; };

__device__ __forceinline__ const uint16_t *seg_row(const Seg3 &s, int r, int cols, float *&dst) {
    int k = 0;
    while (k + 1 < s.n && r >= s.rows[k]) { r -= s.rows[k]; k++; }
    dst = s.out[k] + r;
    return s.w[k] + (size_t)r * cols;
}

/* this CTA's share of out = W . xs (+ residual read through L2) */
__device__ __forceinline__ void cta_gemv(const Seg3 &s, int cols, const float *xs, const float *resid, int cta, int ncta) {
    const int total = s.rows[0] + (s.n > 1 ? s.rows[1] : 0) + (s.n > 2 ? s.rows[2] : 0);
    int r0, r1;
    row_range(total, cta, ncta, r0, r1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = r0 + warp; r < r1; r += 2 * FUSED_WARPS) {
        const int rb = r + FUSED_WARPS < r1 ? r + FUSED_WARPS : r;
        float *da, *db;
        const uint16_t *pa = seg_row(s, r, cols, da), *pb = seg_row(s, rb, cols, db);
        float va, vb;
        warp_dot2(pa, pb, xs, cols, va, vb);
        if (lane == 0) {
            *da = resid ? __ldcg(resid + r) + va : va;
            if (rb != r) *db = resid ? __ldcg(resid + rb) + vb : vb;
        }
    }
}

__device__ __forceinline__ void prefetch_seg(const Seg3 &s, int cols, int cta, int ncta) {
    const int total = s.rows[0] + (s.n > 1 ? s.rows[1] : 0) + (s.n > 2 ? s.rows[2] : 0);
    int r0, r1;
    row_range(total, cta, ncta, r0, r1);
    /* the last warp issues: rows of one segment are contiguous, so a handful of <= 32 KB bulk prefetches covers them */
    if ((threadIdx.x >> 5) != FUSED_WARPS - 1) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    for (int k = 0; k < s.n; k++) {
        const int lo = r0 > base ? r0 - base : 0, hi = (r1 < base + s.rows[k] ? r1 : base + s.rows[k]) - base;
        if (hi > lo) {
            const char *p = reinterpret_cast<const char *>(s.w[k]) + (size_t)lo * cols * 2;
            const size_t bytes = (size_t)(hi - lo) * cols * 2;
            for (size_t off = (size_t)lane * 32768; off < bytes; off += 32 * 32768) {
                const size_t nb = bytes - off < 32768 ? bytes - off : 32768;
                prefetch_l2_bulk(p + off, (unsigned)nb);
            }
        }
        base += s.rows[k];
    }
}

/* rows [r0, r1) of one matrix -> L2, issued by the last warp in <= 32 KB bulk prefetches */
__device__ __forceinline__ void prefetch_rows(const uint16_t *w, int r0, int r1, int cols) {
    if ((threadIdx.x >> 5) != FUSED_WARPS - 1 || r1 <= r0) return;
    const int lane = threadIdx.x & 31;
    const char *p = reinterpret_cast<const char *>(w) + (size_t)r0 * cols * 2;
    const size_t bytes = (size_t)(r1 - r0) * cols * 2;
    for (size_t off = (size_t)lane * 32768; off < bytes; off += 32 * 32768) {
        const size_t nb = bytes - off < 32768 ? bytes - off : 32768;
        prefetch_l2_bulk(p + off, (unsigned)nb);
    }
}
/* this CTA's gate/up units (rows u0..u1 of both matrices) -> L2 */
__device__ __forceinline__ void prefetch_swiglu(const uint16_t *gate, const uint16_t *up, int inter, int cols, int cta, int ncta) {
    int u0, u1;
    row_range(inter, cta, ncta, u0, u1);
    prefetch_rows(gate, u0, u1, cols);
    prefetch_rows(up, u0, u1, cols);
}

/* weight phase `k` of layer `l`: 0 QKV, 1 O, 2 gate+up (as two segments of I rows), 3 down */
__device__ __forceinline__ bool phase_seg(const FusedArgs &a, const LmDev &lm, int l, int k, Seg3 &s, int &cols) {
    if (l >= lm.layers) {
        if (l == lm.layers && k == 0 && a.head_w) { s.n = 1; s.w[0] = a.head_w; s.out[0] = a.logits; s.rows[0] = a.head_rows; cols = lm.hidden; return true; }
        return false;
    }
    const LmLayerDev &w = lm.layer[l];
    const int qd = lm.heads * lm.head_dim, kvd = lm.kv_heads * lm.head_dim;
    if (k == 0) { s.n = 3; s.w[0] = w.wq; s.w[1] = w.wk; s.w[2] = w.wv; s.out[0] = a.q; s.out[1] = a.kk; s.out[2] = a.vv; s.rows[0] = qd; s.rows[1] = kvd; s.rows[2] = kvd; cols = lm.hidden; }
    else if (k == 1) { s.n = 1; s.w[0] = w.wo; s.out[0] = a.x; s.rows[0] = lm.hidden; cols = qd; }
    else if (k == 2) { s.n = 2; s.w[0] = w.gate; s.w[1] = w.up; s.out[0] = a.g; s.out[1] = a.g; s.rows[0] = lm.inter; s.rows[1] = lm.inter; cols = lm.hidden; }
    else { s.n = 1; s.w[0] = w.down; s.out[0] = a.x; s.rows[0] = lm.hidden; cols = lm.inter; }
    return true;
}
__device__ __forceinline__ void prefetch_phase(const FusedArgs &a, const LmDev &lm, int p, int cta, int ncta) {
    Seg3 s; int cols;
    if (a.flags & 1) return;
    if (phase_seg(a, lm, p >> 2, p & 3, s, cols)) prefetch_seg(s, cols, cta, ncta);
}

/* gate/up: unit u = (gate row u, up row u) -> g[u] = silu(gate.x) * (up.x) */
__device__ __forceinline__ void cta_swiglu(const uint16_t *gate, const uint16_t *up, float *g, int inter, int cols, const float *xs, int cta, int ncta) {
    int u0, u1;
    row_range(inter, cta, ncta, u0, u1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int u = u0 + warp; u < u1; u += FUSED_WARPS) {
        float va, vb;
        warp_dot2(gate + (size_t)u * cols, up + (size_t)u * cols, xs, cols, va, vb);
        if (lane == 0) g[u] = swiglu_act<QTC_FAST>(va, vb);
    }
}
