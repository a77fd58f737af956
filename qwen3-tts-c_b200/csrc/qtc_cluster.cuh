/*
 * qtc_cluster.cuh - k_lm_cluster (mode 2): one decoder-stack step with TWO grid-wide exchanges per layer
 * instead of five (talker_attention_single, c/qwen_tts_talker.c:119-248; ST_FORWARD :638-691).
 *
 * Round 1/2 measurements (profiles/r01_phase_clocks_*_v9.txt): of the 36.0 us a talker layer takes in k_lm_fused,
 * 13.1 us are its five grid barriers; of the sub-talker's 22.0 us, 11.4 us.  Cheaper barriers lost (mode 3); this
 * kernel removes three of them by changing who owns what:
 *
 *   grid = up to 15 clusters x 8 CTAs of 512 threads (a B200 holds 15 clusters of 8 at one CTA per SM, 7 of 16:
 *   scripts/probe/cluster_capacity.cu; k_lm_fused at 64 CTAs is only 14-20 % slower than at 148 - the step is bound by
 *   exchanges, not by SM count).
 *   Phase A (cluster g < kv_heads owns KV head g and its two query heads; the other clusters go straight to the barrier):
 *     RMSNorm -> the 512 rows [q_2g | q_2g+1 | k_g | v_g] of the QKV projection, 64 rows per CTA, results pushed
 *     into every peer's shared memory (st.shared::cluster) -> cluster barrier -> attention over the cached
 *     positions split 8 ways (both query heads share each K/V row), CTA partials (max, sum, out) combined through
 *     DSMEM -> the O projection restricted to this cluster's 256 input columns (weights re-laid out at upload so the
 *     slice is contiguous) = a PARTIAL sum for every row of x.
 *   Phase B (every CTA owns a run of 8-column chunks of the MLP's inner dimension): RMSNorm -> gate/up -> SwiGLU in
 *     shared memory -> the down projection restricted to those input columns (re-laid out [I/8][H][8]) = a partial
 *     for every row of x, which every CTA adds straight into the accumulator (folding the cluster's 8 partials through DSMEM
 *     first saved 7/8 of the atomics but cost a cluster barrier and a fold: 1.4 us of a 17 us sub-talker layer).
 *   The partial sums of both phases meet in HBM/L2 as 64-bit FIXED-POINT atomic adds (scale 2^32): integer addition
 *   is associative, so the result does not depend on arrival order - the step stays bit-reproducible - and the
 *   residual stream is carried exactly.  Three accumulators rotate: phase p reads buf[p % 3], adds into
 *   buf[(p + 1) % 3] (pre-zeroed), zeroes buf[(p + 2) % 3]; one grid barrier per phase.
 *
 * Arithmetic: fast mode only (fp32 FMA, tree reductions); results differ from k_lm_fused by summation order.
 */
#pragma once
#include <cooperative_groups.h>
#include "qtc_fused_main.cuh"

namespace cg = cooperative_groups;

#define CL_SIZE 8
/* build-time variants for A/B measurements (scripts/probe): each defaults to the measured-best setting */
#ifndef CL_NW_LOOP
#define CL_NW_LOOP 1          /* fold only the warps that had a position */
#endif
#ifndef CL_PIPE
#define CL_PIPE 1             /* second QKV pair and second O batch requested before the first is reduced */
#endif
#ifndef CL_STAGE_AUX
#define CL_STAGE_AUX 1        /* RoPE row, head norms, layer norms staged to shared memory across the grid barriers */
#endif
#define CL_THREADS 512
#define CL_WARPS 16
#define CL_ATT_CAP 64                 /* most positions per CTA the shared-memory K/V stage is sized for (contexts up to 8 x 64 = 512; longer ones read K/V through L2) */
#define CL_REC 132                    /* partial record: max, sum, -, -, out[128] */
#define CL_FULL_MAX 512               /* contexts up to this many positions: every CTA of a cluster walks the whole context (no partial records across CTAs) */
#define CL_FX 4294967296.0f           /* 2^32 */
#define CL_FX_INV 2.3283064365386963e-10f

struct ClusterArgs {
    const LmDev *lm;
    float *x;                          /* [H] input on entry, post-final-norm hidden on exit (talker) */
    long long *xacc;                   /* [3][H] fixed-point residual stream */
    int *phase;                        /* device: phase counter mod 3, carried across launches */
    float *kc, *vc; const int *table; const float *cos, *sin; int kv_max;
    const int *pos_ptr; int pos_add;
    const uint16_t *head_w; int head_rows; float *logits;
    unsigned long long *bar;
    GenState *st;
    int xs_floats;                     /* staged-vector / sampler region */
    int att_cap;                       /* positions per CTA the K/V stage holds (<= CL_ATT_CAP): shared memory is sized by it, and the less of it, the more L1 */
    int flags;
    long long *prof; int prof_layer;
    int *err;                          /* set to 1 if a barrier wait gave up */
    int pf_frac, pf_gap;               /* idle-cluster L2 prefetch of the MLP weights: percent of every consumer's slice, and nanoseconds between a warp's requests */
};

__device__ __forceinline__ void red_add_s64(long long *p, long long v) {
    asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ long long to_fx(float v) { return __float2ll_rn(v * CL_FX); }
__device__ __forceinline__ float from_fx(long long v) { return __ll2float_rn(v) * CL_FX_INV; }

/* grid barrier with a bounded spin: a lost CTA must not hang the device */
__device__ __forceinline__ void cl_barrier_wait(unsigned long long *bar, const unsigned long long &target, int *err) {
    if (threadIdx.x == 0) {
        unsigned spins = 0;
        while (ld_acquire_u64(bar) < target) {
            if (++spins > (1u << 26)) { if (err) *err = 1; break; }
        }
    }
    __syncthreads();
}
__device__ __forceinline__ void cl_barrier_arrive(unsigned long long *bar, unsigned long long &target, unsigned nb) {
    __syncthreads();
    if (threadIdx.x == 0) { target += nb; red_release_inc(bar); }
}

/* xs = rmsnorm(x) * w for H <= 4096; x comes as fp32 (the step's input) or from a fixed-point accumulator.  In two halves so
 * that independent work (the K/V staging copies) can be issued while the loads are in flight. */
struct XRegs { float4 v0, v1, w0, w1; };
__device__ __forceinline__ void cl_x_load(XRegs &x, const float *xf, const long long *xq, const float *w, int n) {
    const int i0 = threadIdx.x * 4, i1 = i0 + CL_THREADS * 4;
    x.v0 = make_float4(0, 0, 0, 0); x.v1 = x.v0; x.w0 = x.v0; x.w1 = x.v0;
    if (xf) {
        if (i0 < n) x.v0 = __ldcg(reinterpret_cast<const float4 *>(xf + i0));
        if (i1 < n) x.v1 = __ldcg(reinterpret_cast<const float4 *>(xf + i1));
    } else {
        if (i0 < n) {
            const longlong2 a = __ldcg(reinterpret_cast<const longlong2 *>(xq + i0)), b = __ldcg(reinterpret_cast<const longlong2 *>(xq + i0 + 2));
            x.v0 = make_float4(from_fx(a.x), from_fx(a.y), from_fx(b.x), from_fx(b.y));
        }
        if (i1 < n) {
            const longlong2 a = __ldcg(reinterpret_cast<const longlong2 *>(xq + i1)), b = __ldcg(reinterpret_cast<const longlong2 *>(xq + i1 + 2));
            x.v1 = make_float4(from_fx(a.x), from_fx(a.y), from_fx(b.x), from_fx(b.y));
        }
    }
    if (i0 < n) x.w0 = *reinterpret_cast<const float4 *>(w + i0);
    if (i1 < n) x.w1 = *reinterpret_cast<const float4 *>(w + i1);
}
__device__ __forceinline__ void cl_norm_finish(float *xs, const XRegs &x, int n, float eps, float *red) {
    const int i0 = threadIdx.x * 4, i1 = i0 + CL_THREADS * 4;
    const float4 v0 = x.v0, v1 = x.v1, w0 = x.w0, w1 = x.w1;
    float ss = v0.x * v0.x;
    ss = fmaf(v0.y, v0.y, ss); ss = fmaf(v0.z, v0.z, ss); ss = fmaf(v0.w, v0.w, ss);
    ss = fmaf(v1.x, v1.x, ss); ss = fmaf(v1.y, v1.y, ss); ss = fmaf(v1.z, v1.z, ss); ss = fmaf(v1.w, v1.w, ss);
    ss = csum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)n + eps);
    if (i0 < n) *reinterpret_cast<float4 *>(xs + i0) = make_float4(v0.x * inv * w0.x, v0.y * inv * w0.y, v0.z * inv * w0.z, v0.w * inv * w0.w);
    if (i1 < n) *reinterpret_cast<float4 *>(xs + i1) = make_float4(v1.x * inv * w1.x, v1.y * inv * w1.y, v1.z * inv * w1.z, v1.w * inv * w1.w);
    __syncthreads();
}
__device__ __forceinline__ void cl_stage_norm(float *xs, const float *xf, const long long *xq, const float *w, int n, float eps, float *red) {
    XRegs x;
    cl_x_load(x, xf, xq, w, n);
    cl_norm_finish(xs, x, n, eps, red);
}

/* ---- weight streaming in register stages ------------------------------------------------------------------------------------
 * A Stage holds 1024 columns of two rows: lane's 16-byte pieces lane + 32 j, j < 4, of each row = 8 loads.  Two stages are kept in
 * flight per warp: while one is multiplied out the other's loads (and the next one's, issued right after the multiply) are under
 * way, so a phase is a continuous stream of requests instead of load - wait - multiply round trips; the first two stages of a phase
 * are issued BEFORE the grid barrier in front of it.  Every register of a stage is written on every issue (zeros where there is
 * nothing to load): a conditionally written stage would stay live around the whole layer loop and be spilled. */
struct Stage { uint4 a[4], b[4]; };
__device__ __forceinline__ void st_issue(Stage &s, const uint16_t *ra, const uint16_t *rb, int half) {
    const uint4 *w0 = reinterpret_cast<const uint4 *>(ra) + half * 128 + (threadIdx.x & 31), *w1 = reinterpret_cast<const uint4 *>(rb) + half * 128 + (threadIdx.x & 31);
#pragma unroll
    for (int j = 0; j < 4; j++) { s.a[j] = ldg_stream(w0 + 32 * j); s.b[j] = ldg_stream(w1 + 32 * j); }
}
__device__ __forceinline__ void st_zero(Stage &s) {
#pragma unroll
    for (int j = 0; j < 4; j++) { s.a[j] = make_uint4(0, 0, 0, 0); s.b[j] = make_uint4(0, 0, 0, 0); }
}
__device__ __forceinline__ void st_consume(const Stage &s, const float *xs, int half, float &a, float &b) {
    const float4 *x4 = reinterpret_cast<const float4 *>(xs) + 2 * (half * 128 + (threadIdx.x & 31));
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const float4 x0 = x4[64 * j], x1 = x4[64 * j + 1];
        a = dot8f(s.a[j], x0, x1, a);
        b = dot8f(s.b[j], x0, x1, b);
    }
}
__device__ __forceinline__ void warp_sum2(float &a, float &b) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ta = __shfl_xor_sync(0xffffffffu, a, o), tb = __shfl_xor_sync(0xffffffffu, b, o);
        a += ta; b += tb;
    }
}
/* npairs row pairs of HS x 1024 columns each, streamed through (s0, s1).  On entry s0 holds stage 0 and s1 stage 1 (a stage = one
 * pair at HS == 1, half a pair at HS == 2).  rows(k, ra, rb): the k-th pair's row pointers; out(k, a, b): its two dot products. */
template <class RowFn, class OutFn>
__device__ __forceinline__ void cl_stream(Stage &s0, Stage &s1, int npairs, int HS, const float *xs, RowFn rows, OutFn out) {
    const int nst = npairs * HS;
    float a = 0.0f, b = 0.0f;
    const uint16_t *ra, *rb;
    for (int i = 0; i < nst; i += 2) {
        st_consume(s0, xs, 0, a, b);
        if (HS == 1) { warp_sum2(a, b); out(i, a, b); a = 0.0f; b = 0.0f; }
        if (i + 2 < nst) { rows(HS == 2 ? (i + 2) >> 1 : i + 2, ra, rb); st_issue(s0, ra, rb, 0); }
        if (i + 1 < nst) {
            st_consume(s1, xs, HS == 2 ? 1 : 0, a, b);
            warp_sum2(a, b);
            out(HS == 2 ? i >> 1 : i + 1, a, b);
            a = 0.0f; b = 0.0f;
            if (i + 3 < nst) { rows(HS == 2 ? (i + 3) >> 1 : i + 3, ra, rb); st_issue(s1, ra, rb, HS == 2 ? 1 : 0); }
        }
    }
}
/* stages 0 and 1 of such a stream (issued ahead of a barrier) */
template <class RowFn>
__device__ __forceinline__ void cl_stream_prefetch(Stage &s0, Stage &s1, int npairs, int HS, RowFn rows) {
    const uint16_t *ra, *rb;
    if (npairs > 0) { rows(0, ra, rb); st_issue(s0, ra, rb, 0); } else st_zero(s0);
    if (npairs * HS > 1) { rows(HS == 2 ? 0 : 1, ra, rb); st_issue(s1, ra, rb, HS == 2 ? 1 : 0); } else st_zero(s1);
}

/* row j (0..511) of cluster g's QKV block: [q head 2g | q head 2g+1 | k head g | v head g] */
__device__ __forceinline__ const uint16_t *cl_qkv_row(const LmLayerDev &w, int g, int j, int H) {
    if (j < 256) return w.wq + ((size_t)(g * 256 + j)) * H;
    if (j < 384) return w.wk + ((size_t)(g * 128 + j - 256)) * H;
    return w.wv + ((size_t)(g * 128 + j - 384)) * H;
}

/* per-warp: RMS-normalise a 128-wide head held in shared memory and rotate it (lane i owns dims 4i..4i+3) */
__device__ __forceinline__ float4 cl_norm_rope(const float *src, const float *nw, const float4 c4, const float4 s4, float eps) {
    const int lane = threadIdx.x & 31;
    const float4 v = reinterpret_cast<const float4 *>(src)[lane], n4 = reinterpret_cast<const float4 *>(nw)[lane];
    float ss = v.x * v.x;
    ss = fmaf(v.y, v.y, ss); ss = fmaf(v.z, v.z, ss); ss = fmaf(v.w, v.w, ss);
    ss = warp_sum(ss);
    const float inv = rsqrtf(ss * 0.0078125f + eps);
    const float4 y = make_float4(v.x * inv * n4.x, v.y * inv * n4.y, v.z * inv * n4.z, v.w * inv * n4.w);
    const float px = __shfl_xor_sync(0xffffffffu, y.x, 16), py = __shfl_xor_sync(0xffffffffu, y.y, 16);
    const float pz = __shfl_xor_sync(0xffffffffu, y.z, 16), pw = __shfl_xor_sync(0xffffffffu, y.w, 16);
    return lane < 16 ? make_float4(y.x * c4.x - px * s4.x, y.y * c4.y - py * s4.y, y.z * c4.z - pz * s4.z, y.w * c4.w - pw * s4.w)
                     : make_float4(y.x * c4.x + px * s4.x, y.y * c4.y + py * s4.y, y.z * c4.z + pz * s4.z, y.w * c4.w + pw * s4.w);
}
__device__ __forceinline__ float dot4(const float4 a, const float4 b) {
    float d = a.x * b.x;
    d = fmaf(a.y, b.y, d); d = fmaf(a.z, b.z, d); d = fmaf(a.w, b.w, d);
    return d;
}

/* Attention of cluster g at position pos: positions [t0, t1) of this CTA, both query heads per K/V row; the first nst of them sit in
 * the shared-memory stage, the others are read through L2.  qkv: this CTA's copy of the cluster's 512 projection outputs.
 * att == NULL (contexts split over the cluster's CTAs): the CTA's partial record goes into every peer's table cpart[rank][head];
 * att != NULL (the CTA took the WHOLE context, t0 = 0, t1 = pos + 1): the finished 256 outputs are written there. */
__device__ __forceinline__ void cl_attention(cg::cluster_group &cl, const ClusterArgs &a, const LmDev &lm, int l, int g, int pos, int t0, int t1, int nst, bool append,
                                             const float *qkv, float *wpart, float *cpart, float *att, const float *kvs, const float *aux) {
    const int cap = a.att_cap;
#define APROF(i) do { if (a.prof && threadIdx.x == 0 && l == a.prof_layer) a.prof[blockIdx.x * 32 + (i)] = clock64(); } while (0)
    /* aux (staged with the K/V rows): [cos row | sin row | q_norm | k_norm], 128 floats each */
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, kvd = lm.kv_heads * 128;
    const float *q_norm = aux + 256, *k_norm = aux + 384;
    float *kc = a.kc + (size_t)l * a.kv_max * kvd + g * 128, *vc = a.vc + (size_t)l * a.kv_max * kvd + g * 128;
    const float4 c4 = reinterpret_cast<const float4 *>(aux)[lane], s4 = reinterpret_cast<const float4 *>(aux + 128)[lane];
    const bool mine = warp < t1 - t0;                                  /* this warp has at least one position */
    const bool has_new = t1 == pos + 1 && pos >= t0 && ((pos - t0) % CL_WARPS) == warp;
    float m0 = -3.402823466e+38f, m1 = m0, l0 = 0.0f, l1 = 0.0f;
    float4 o0 = make_float4(0, 0, 0, 0), o1 = o0;
    if (mine) {
        const float4 q0 = cl_norm_rope(qkv, q_norm, c4, s4, lm.eps), q1 = cl_norm_rope(qkv + 128, q_norm, c4, s4, lm.eps);
        float4 k_new = make_float4(0, 0, 0, 0), v_new = k_new;
        if (has_new) {
            k_new = cl_norm_rope(qkv + 256, k_norm, c4, s4, lm.eps);
            v_new = reinterpret_cast<const float4 *>(qkv + 384)[lane];
            if (append) {
                const size_t off = kv_off(a.table, pos, kvd);
                *reinterpret_cast<float4 *>(kc + off + lane * 4) = k_new;
                *reinterpret_cast<float4 *>(vc + off + lane * 4) = v_new;
            }
        }
        APROF(16);
        const float *Ks = kvs, *Vs = kvs + cap * 128;
        const float scale = 0.08838834764831845f;                      /* 1 / sqrt(128) */
        for (int tb = t0 + warp; tb < t1; tb += 4 * CL_WARPS) {
            float d0[4], d1[4];
            float4 vr[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int t = tb + CL_WARPS * j;
                float4 kr = make_float4(0, 0, 0, 0);
                vr[j] = kr;
                if (t < t1) {
                    if (t < pos) {
                        if (t - t0 < nst) {
                            kr = *reinterpret_cast<const float4 *>(Ks + (size_t)(t - t0) * 128 + lane * 4);
                            vr[j] = *reinterpret_cast<const float4 *>(Vs + (size_t)(t - t0) * 128 + lane * 4);
                        } else {
                            const size_t off = kv_off(a.table, t, kvd);
                            kr = __ldcg(reinterpret_cast<const float4 *>(kc + off) + lane);
                            vr[j] = __ldcg(reinterpret_cast<const float4 *>(vc + off) + lane);
                        }
                    } else { kr = k_new; vr[j] = v_new; }
                }
                d0[j] = dot4(q0, kr); d1[j] = dot4(q1, kr);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                float u0[4], u1[4];
#pragma unroll
                for (int j = 0; j < 4; j++) { u0[j] = __shfl_xor_sync(0xffffffffu, d0[j], o); u1[j] = __shfl_xor_sync(0xffffffffu, d1[j], o); }
#pragma unroll
                for (int j = 0; j < 4; j++) { d0[j] += u0[j]; d1[j] += u1[j]; }
            }
            float b0 = m0, b1 = m1;
#pragma unroll
            for (int j = 0; j < 4; j++) if (tb + CL_WARPS * j < t1) { d0[j] *= scale; d1[j] *= scale; b0 = fmaxf(b0, d0[j]); b1 = fmaxf(b1, d1[j]); }
            const float f0 = __expf(m0 - b0), f1 = __expf(m1 - b1);      /* first batch: exp(-huge) = 0 on zeros */
            l0 *= f0; o0.x *= f0; o0.y *= f0; o0.z *= f0; o0.w *= f0;
            l1 *= f1; o1.x *= f1; o1.y *= f1; o1.z *= f1; o1.w *= f1;
            m0 = b0; m1 = b1;
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (tb + CL_WARPS * j < t1) {
                    const float e0 = __expf(d0[j] - m0), e1 = __expf(d1[j] - m1);
                    l0 += e0; l1 += e1;
                    o0.x = fmaf(e0, vr[j].x, o0.x); o0.y = fmaf(e0, vr[j].y, o0.y); o0.z = fmaf(e0, vr[j].z, o0.z); o0.w = fmaf(e0, vr[j].w, o0.w);
                    o1.x = fmaf(e1, vr[j].x, o1.x); o1.y = fmaf(e1, vr[j].y, o1.y); o1.z = fmaf(e1, vr[j].z, o1.z); o1.w = fmaf(e1, vr[j].w, o1.w);
                }
        }
    }
    APROF(17);
    /* warp records: [warp][head][max, sum, -, -, out[128]]; only warps that had a position write and are folded */
#if CL_NW_LOOP
    const int nw = t1 - t0 < CL_WARPS ? t1 - t0 : CL_WARPS;
    if (mine) {
#else
    const int nw = CL_WARPS;
    {
#endif
        float *r0 = wpart + (size_t)(warp * 2) * CL_REC, *r1 = r0 + CL_REC;
        if (lane == 0) { r0[0] = m0; r0[1] = l0; r1[0] = m1; r1[1] = l1; }
        *reinterpret_cast<float4 *>(r0 + 4 + lane * 4) = o0;
        *reinterpret_cast<float4 *>(r1 + 4 + lane * 4) = o1;
    }
    __syncthreads();
    APROF(18);
    if (tid < 256) {
        const int hh = tid >> 7, d = tid & 127;
        float M = -3.402823466e+38f;
#pragma unroll 8
        for (int k = 0; k < nw; k++) M = fmaxf(M, wpart[(size_t)(k * 2 + hh) * CL_REC]);
        float L = 0.0f, o = 0.0f;
#pragma unroll 8
        for (int k = 0; k < nw; k++) {                     /* select, not branch: the loads of all records go out together */
            const float *r = wpart + (size_t)(k * 2 + hh) * CL_REC;
            const float m = r[0], lw = r[1], ov = r[4 + d];
            const float e = __expf(m - M);
            const float f = lw > 0.0f ? e : 0.0f;
            L = fmaf(lw, f, L);
            o = fmaf(ov, f, o);
        }
        if (att) att[hh * 128 + d] = __fdividef(o, L);      /* the whole context was this CTA's */
        else {
            /* the CTA's record goes into every peer's table: after the cluster barrier each CTA folds the 8 records locally */
            float *mine = cpart + (size_t)(cl.block_rank() * 2 + hh) * CL_REC;
#pragma unroll
            for (int k = 0; k < CL_SIZE; k++) {
                float *dst = cl.map_shared_rank(mine, k);
                dst[4 + d] = o;
                if (d == 0) { dst[0] = M; dst[1] = L; }
            }
        }
    }
    APROF(19);
}
#undef APROF

/* Contexts of at most 32 positions (every sub-talker step): the whole head in ONE warp, no partial records, no fold.  EVERY CTA of
 * the cluster runs both heads for itself (warps 0 and 1; it holds the cluster's 512 projection outputs anyway and stages the <= 31
 * cached K/V rows across the grid barrier like the two CTAs that used to do the work and push the 256 outputs to their peers):
 * the cluster barrier between attention and O projection - 0.7 us of a 17 us layer - becomes a block barrier.  Rank 0 appends the
 * new K/V row to the cache.  The warp's latency chain is kept short:
 * the q and k norms run as one interleaved butterfly, scores are position-parallel (lane t takes the whole 128-wide dot product of
 * position t against q in shared memory - K rows are staged with a 132-float stride so that the lanes' rows fall into different
 * banks - instead of 40 shuffles per eight positions), V rows are weighted with lane-owns-4-dims.  kvs: rows [0, pos) of K, then V. */
#define CL_KST_SMALL 132
__device__ __forceinline__ void cl_attention_small(const ClusterArgs &a, const LmDev &lm, int l, int g, int pos, int hh, bool append,
                                                   const float *qkv, float *att, const float *kvs, const float *aux, float *scr) {
#define SPROF(i) do { if (a.prof && threadIdx.x == 0 && l == a.prof_layer) a.prof[blockIdx.x * 32 + (i)] = clock64(); } while (0)
    SPROF(24);
    const int lane = threadIdx.x & 31, kvd = lm.kv_heads * 128, n = pos + 1, cap = a.att_cap;
    float *qs = scr, *ks = scr + 128;
    const float4 c4 = reinterpret_cast<const float4 *>(aux)[lane], s4 = reinterpret_cast<const float4 *>(aux + 128)[lane];
    const float4 qv = reinterpret_cast<const float4 *>(qkv + hh * 128)[lane], kv = reinterpret_cast<const float4 *>(qkv + 256)[lane];
    const float4 qn = reinterpret_cast<const float4 *>(aux + 256)[lane], kn = reinterpret_cast<const float4 *>(aux + 384)[lane];
    const float4 v_new = reinterpret_cast<const float4 *>(qkv + 384)[lane];
    float sq = dot4(qv, qv), sk = dot4(kv, kv);
    warp_sum2(sq, sk);
    const float iq = rsqrtf(sq * 0.0078125f + lm.eps), ik = rsqrtf(sk * 0.0078125f + lm.eps);
    const float4 qy = make_float4(qv.x * iq * qn.x, qv.y * iq * qn.y, qv.z * iq * qn.z, qv.w * iq * qn.w);
    const float4 ky = make_float4(kv.x * ik * kn.x, kv.y * ik * kn.y, kv.z * ik * kn.z, kv.w * ik * kn.w);
    const float sgn = lane < 16 ? -1.0f : 1.0f;          /* rotate-half: first half y cos - partner sin, second half y cos + partner sin */
    float4 q4, k_new;
    {
        const float qx = __shfl_xor_sync(0xffffffffu, qy.x, 16), qyy = __shfl_xor_sync(0xffffffffu, qy.y, 16);
        const float qz = __shfl_xor_sync(0xffffffffu, qy.z, 16), qw = __shfl_xor_sync(0xffffffffu, qy.w, 16);
        const float kx = __shfl_xor_sync(0xffffffffu, ky.x, 16), kyy = __shfl_xor_sync(0xffffffffu, ky.y, 16);
        const float kz = __shfl_xor_sync(0xffffffffu, ky.z, 16), kw = __shfl_xor_sync(0xffffffffu, ky.w, 16);
        q4 = make_float4(fmaf(sgn * qx, s4.x, qy.x * c4.x), fmaf(sgn * qyy, s4.y, qy.y * c4.y), fmaf(sgn * qz, s4.z, qy.z * c4.z), fmaf(sgn * qw, s4.w, qy.w * c4.w));
        k_new = make_float4(fmaf(sgn * kx, s4.x, ky.x * c4.x), fmaf(sgn * kyy, s4.y, ky.y * c4.y), fmaf(sgn * kz, s4.z, ky.z * c4.z), fmaf(sgn * kw, s4.w, ky.w * c4.w));
    }
    reinterpret_cast<float4 *>(qs)[lane] = q4;
    reinterpret_cast<float4 *>(ks)[lane] = k_new;
    if (append) {
        float *kc = a.kc + (size_t)l * a.kv_max * kvd + g * 128, *vc = a.vc + (size_t)l * a.kv_max * kvd + g * 128;
        const size_t off = kv_off(a.table, pos, kvd);
        *reinterpret_cast<float4 *>(kc + off + lane * 4) = k_new;
        *reinterpret_cast<float4 *>(vc + off + lane * 4) = v_new;
    }
    __syncwarp();
    SPROF(25);
    const float *Ks = kvs, *Vs = kvs + cap * 128;
    float sc = -3.402823466e+38f;
    if (lane < n) {
        const float4 *kr = reinterpret_cast<const float4 *>(lane < pos ? Ks + (size_t)lane * CL_KST_SMALL : ks), *q4p = reinterpret_cast<const float4 *>(qs);
        float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll 8
        for (int c = 0; c < 32; c++) {
            const float4 kk = kr[c], qq = q4p[c];
            a0 = fmaf(kk.x, qq.x, a0); a1 = fmaf(kk.y, qq.y, a1); a2 = fmaf(kk.z, qq.z, a2); a3 = fmaf(kk.w, qq.w, a3);
        }
        sc = ((a0 + a1) + (a2 + a3)) * 0.08838834764831845f;
    }
    SPROF(26);
    const float mx = warp_max(sc);
    const float e = lane < n ? __expf(sc - mx) : 0.0f;
    const float tot = warp_sum(e);
    SPROF(27);
    float4 o4 = make_float4(0, 0, 0, 0);
    for (int tb = 0; tb < n; tb += 8) {                  /* eight rows requested before the first is used: the loop was one shuffle -> load -> FMA chain per row */
        float pr[8];
        float4 vr[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int t = tb + u;
            pr[u] = __shfl_sync(0xffffffffu, e, t & 31);
            vr[u] = t < pos ? *reinterpret_cast<const float4 *>(Vs + (size_t)t * 128 + lane * 4) : v_new;
            if (t >= n) pr[u] = 0.0f;
        }
#pragma unroll
        for (int u = 0; u < 8; u++) { o4.x = fmaf(pr[u], vr[u].x, o4.x); o4.y = fmaf(pr[u], vr[u].y, o4.y); o4.z = fmaf(pr[u], vr[u].z, o4.z); o4.w = fmaf(pr[u], vr[u].w, o4.w); }
    }
    SPROF(28);
    const float inv = __fdividef(1.0f, tot);
    const float4 r4 = make_float4(o4.x * inv, o4.y * inv, o4.z * inv, o4.w * inv);
    *reinterpret_cast<float4 *>(att + hh * 128 + lane * 4) = r4;
    SPROF(29);
}
#undef SPROF

/* every CTA folds the 8 CTA records of both heads (pushed into its table cpart[rank][head] by the peers) for itself */
__device__ __forceinline__ void cl_attention_combine(const float *cpart, float *att) {
    const int tid = threadIdx.x;
    if (tid < 256) {
        const int hh = tid >> 7, d = tid & 127;
        float m[CL_SIZE], lsum[CL_SIZE], o[CL_SIZE];
#pragma unroll
        for (int k = 0; k < CL_SIZE; k++) {
            const float *rp = cpart + (size_t)(k * 2 + hh) * CL_REC;
            m[k] = rp[0]; lsum[k] = rp[1]; o[k] = rp[4 + d];
        }
        float M = m[0];
#pragma unroll
        for (int k = 1; k < CL_SIZE; k++) M = fmaxf(M, m[k]);
        float L = 0.0f, acc = 0.0f;
#pragma unroll
        for (int k = 0; k < CL_SIZE; k++) {
            const float f = lsum[k] > 0.0f ? __expf(m[k] - M) : 0.0f;
            L = fmaf(lsum[k], f, L);
            acc = fmaf(o[k], f, acc);
        }
        att[hh * 128 + d] = __fdividef(acc, L);
    }
    __syncthreads();
}

/* L2 prefetch of slice `part` of `nparts` of [base, base + bytes): 32 KB bulk prefetches issued by lane 0 of every warp */
__device__ __forceinline__ void cl_prefetch_slice(const void *base, size_t bytes, int part, int nparts) {
    if ((threadIdx.x & 31) != 0) return;
    const size_t per = ((bytes / nparts) + 127) & ~(size_t)127;
    size_t lo = per * part, hi = lo + per;
    if (part == nparts - 1 || hi > bytes) hi = bytes;
    const char *p = reinterpret_cast<const char *>(base);
    for (size_t off = lo + (size_t)(threadIdx.x >> 5) * 32768; off < hi; off += (size_t)CL_WARPS * 32768) {
        const size_t nb = hi - off < 32768 ? hi - off : 32768;
        prefetch_l2_bulk(p + off, (unsigned)(nb & ~(size_t)15));
    }
}

/* paced L2 prefetch of the first `bytes` of a range: 16 KB requests dealt out over the CTA's warps (lane 0 issues), `gap` ns apart */
__device__ __forceinline__ void cl_prefetch_paced(const void *base, size_t bytes, unsigned gap, int &slot) {
    if ((threadIdx.x & 31) != 0) return;
    const char *p = reinterpret_cast<const char *>(base);
    for (size_t off = 0; off < bytes; off += 16384, slot++) {
        if ((slot & (CL_WARPS - 1)) != (int)(threadIdx.x >> 5)) continue;
        const size_t nb = bytes - off < 16384 ? bytes - off : 16384;
        prefetch_l2_bulk(p + off, (unsigned)(nb & ~(size_t)15));
        if (gap) __nanosleep(gap);
    }
}

/* dynamic shared memory (floats): [ xs | qkv 512 | att 256 | cpart 8 x 2 x 132 | wpart 16 x 2 x 132 | fold factors 64 | g 64 |
 * aux 512 | in_norm H | post_norm H | kv stage 2 x att_cap x 128 ]; the in-kernel sampler's workspace overlays aux.. (sampling happens
 * between steps; only this CTA writes there).  Kept small on purpose: shared memory is carved out of the SM's 228 KB and what is left is L1, whose size bounds
 * the loads an SM keeps in flight - measured: +17 KB of shared memory (a 132 -> 164 KB carve-out) cost the talker step 12 %. */
__host__ __device__ inline int cl_smem_floats(int xs_floats, int H, int att_cap) {
    return xs_floats + 512 + 256 + CL_SIZE * 2 * CL_REC + CL_WARPS * 2 * CL_REC + 64 + 64 + 512 + 2 * H + 2 * att_cap * 128;
}
__device__ __forceinline__ void cl_cp16(float *dst_smem, const float *src) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src));
}
/* n floats (multiple of 4) global -> shared, cp.async, strided over the CTA */
__device__ __forceinline__ void cl_stage_async(float *dst, const float *src, int n) {
    for (int i = threadIdx.x * 4; i < n; i += CL_THREADS * 4) cl_cp16(dst + i, src + i);
}

__global__ void __launch_bounds__(CL_THREADS, 1) k_lm_cluster(const ClusterArgs a0, const SubProg sp) {
    extern __shared__ __align__(128) float cl_smem[];
    __shared__ float red[32];
    __shared__ int s_codes[QTC_MAX_GROUPS];
    __shared__ uint32_t s_rng;
    cg::cluster_group cl = cg::this_cluster();
    const ClusterArgs &a = a0;                          /* stays in the constant bank: no copy, no registers */
    const uint16_t *head_w = a0.head_w;
    const LmDev &lm = *a.lm;
    const int cta = blockIdx.x, ncta = gridDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rank = (int)cl.block_rank(), g = cta / CL_SIZE;
    const int H = lm.hidden, I = lm.inter, L = lm.layers;
    const bool att_cluster = g < lm.kv_heads;           /* this cluster owns a KV head in phase A */
    float *sm = cl_smem, *qkv = sm + a.xs_floats, *att = qkv + 512, *cpart = att + 256, *wpart = cpart + CL_SIZE * 2 * CL_REC;
    float *gs = wpart + CL_WARPS * 2 * CL_REC + 64, *aux = gs + 64, *nin = aux + 512, *npost = nin + H, *kvs = npost + H;
    SampleSmem &S_samp = *reinterpret_cast<SampleSmem *>(aux);      /* 17.2 KB over aux | in_norm | post_norm | kv stage: regions only this CTA writes (peers push into qkv) */
    const int pos_base = *a.pos_ptr + a.pos_add;
    int p = *a.phase;                                   /* phase counter mod 3 */
    unsigned long long target = 0;
    if (tid == 0) { const unsigned long long v = ld_acquire_u64(a.bar); target = v - v % (unsigned long long)ncta; }
    const int nsteps = sp.G > 0 ? sp.G : 1;
    if (sp.G > 0 && tid == 0) s_codes[0] = sp.codes[0];
    const int RPC = H / CL_SIZE;                        /* rows of x a CTA owns inside its cluster: 128 or 256 */
    const int NCH = I / 8;                              /* 8-column chunks of the MLP's inner dimension, dealt out contiguously */
    const int c0 = (int)(((unsigned)NCH * (unsigned)cta) / (unsigned)ncta), c1 = (int)(((unsigned)NCH * (unsigned)(cta + 1)) / (unsigned)ncta);
    const int u0 = c0 * 8, u1 = c1 * 8, NC = c1 - c0;   /* gate/up units of this CTA; NC <= 8 */
    const int zper = (H + ncta - 1) / ncta, z0 = cta * zper;
    Seg3 s; int cols;
    Stage s0, s1;
    const int HS = H >> 10;                                /* 1024-column stages per weight row */
    const int ja = 64 * rank + warp, jb = ja + CL_WARPS;   /* the warp's first pair of QKV rows; the second pair is +32 */
    const int gu_pairs = (u1 - u0 - warp + CL_WARPS - 1) / CL_WARPS;      /* gate/up units of this warp: u0 + warp + 16 k */
    auto qkv_rows = [&](const LmLayerDev &wl) { return [&wl, g, ja, jb, H](int k, const uint16_t *&ra, const uint16_t *&rb) { ra = cl_qkv_row(wl, g, ja + 32 * k, H); rb = cl_qkv_row(wl, g, jb + 32 * k, H); }; };
    auto gu_rows = [&](const LmLayerDev &wl) { return [&wl, u0, warp, H](int k, const uint16_t *&ra, const uint16_t *&rb) { const size_t u = (size_t)(u0 + warp + CL_WARPS * k); ra = wl.gate + u * H; rb = wl.up + u * H; }; };
    /* Everything phase A of layer ll reads that does not depend on this step's activations goes to shared memory by cp.async while
     * the CTA waits at the grid barrier before it: the K/V rows of this CTA's position range, the RoPE row, the head norms and the
     * input norm's weights.  (Read on demand, each of these was a DRAM round trip on the critical path: the 2.8 GB streamed per
     * step evict them from L2 every time.) */
    auto stage_phase_a = [&](int ll, int pos, int t0, int t1, int rows, int kst) {
        const LmLayerDev &wl = lm.layer[ll];
        if (att_cluster) {
            const int kvd = lm.kv_heads * 128;
            const float *kc = a.kc + (size_t)ll * a.kv_max * kvd + g * 128, *vc = a.vc + (size_t)ll * a.kv_max * kvd + g * 128;
            if (rows > 0) {
                for (int i = tid; i < rows * 64; i += CL_THREADS) {
                    const int which = i >= rows * 32, j = which ? i - rows * 32 : i, r = j >> 5, c = j & 31;
                    cl_cp16(kvs + (size_t)which * a.att_cap * 128 + r * (which ? 128 : kst) + c * 4, (which ? vc : kc) + kv_off(a.table, t0 + r, kvd) + c * 4);
                }
            }
            /* whole-context attention: the rows past the stage -> L2, every CTA of the cluster an eighth of them (128-byte lines: 4 per row and matrix) */
            const int past = (t1 < pos ? t1 : pos) - t0 - rows;
            if (rows > 0 && past > 0 && t0 == 0 && t1 == pos + 1) {
                for (int i = tid; i < past * 8; i += CL_THREADS) {
                    const int r = rows + (i >> 3), which = (i >> 2) & 1, c = i & 3;
                    if ((r & (CL_SIZE - 1)) != rank) continue;
                    asm volatile("prefetch.global.L2 [%0];" ::"l"((which ? vc : kc) + kv_off(a.table, r, kvd) + c * 32));
                }
            }
#if CL_STAGE_AUX
            if (tid < 128) {
                const int k = tid >> 5, c = tid & 31;
                const float *src = k == 0 ? a.cos + (size_t)pos * 128 : k == 1 ? a.sin + (size_t)pos * 128 : k == 2 ? wl.q_norm : wl.k_norm;
                cl_cp16(aux + k * 128 + c * 4, src + c * 4);
            }
            cl_stage_async(nin, wl.in_norm, H);
#endif
        }
        asm volatile("cp.async.commit_group;");
    };
    if (sp.G == 0 && att_cluster) cl_stream_prefetch(s0, s1, 2, HS, qkv_rows(lm.layer[0]));

#define PROF(i) do { if (a.prof && tid == 0 && l == a.prof_layer) a.prof[cta * 32 + (i)] = clock64(); } while (0)
    for (int gstep = 0; gstep < nsteps; gstep++) {
        const int pos = pos_base + gstep;
        const int n = pos + 1;
        const bool small_att = n <= 32 && (n - 1) * CL_KST_SMALL <= a.att_cap * 128;      /* every CTA takes both heads, one warp each */
        const int kst = small_att ? CL_KST_SMALL : 128;    /* K row stride in the stage */
        /* Experiment (QTC_FUSED_FLAGS & 512; off): contexts up to CL_FULL_MAX positions walked WHOLE by every CTA of the cluster (a warp
         * per position, the first att_cap rows from the stage, the others prefetched into L2 and read from there) instead of an eighth
         * each - no partial records, no push into the peers, no cluster barrier, no fold of 8 records.  Parity green, but measured
         * SLOWER at 190 positions (talker step 1.019 vs 0.899 ms on one box): 12 positions per warp, two thirds of them through L2,
         * cost more than the exchange they save. */
        const bool full_att = !small_att && n <= CL_FULL_MAX && (a.flags & 512);
        const int chunk = small_att || full_att ? n : (n + CL_SIZE - 1) / CL_SIZE;
        int t0 = small_att || full_att ? 0 : rank * chunk, t1 = small_att || full_att ? n : (t0 + chunk < n ? t0 + chunk : n);
        if (t0 > n) t0 = n;
        if (t1 < t0) t1 = t0;
        const bool staged = small_att || full_att || chunk <= a.att_cap;
        const int cached = (t1 < pos ? t1 : pos) - t0;     /* rows of [t0, t1) that are in the cache (the last position is the new token) */
        const int nst = !staged ? 0 : (small_att || cached < a.att_cap) ? cached : a.att_cap;       /* ... and how many of them wait in the stage */
        if (sp.G == 0) stage_phase_a(0, pos, t0, t1, nst, kst);
        const float *x_in = a.x;                        /* fp32 input of layer 0 */
        if (sp.G > 0 && gstep >= 1 && sp.ptab[gstep - 1]) {
            /* ST_PROJECT_INPUT by table: the projected embedding of every code was computed at upload, and every CTA knows the code it
             * sampled - no gather, no GEMV, no grid barrier between the sample and the first layer */
            __syncthreads();                            /* the sampler's workspace (over aux..) is free */
            x_in = sp.ptab[gstep - 1] + (size_t)s_codes[gstep - 1] * H;
            head_w = sp.heads[gstep - 1];
            if (att_cluster) cl_stream_prefetch(s0, s1, 2, HS, qkv_rows(lm.layer[0]));
            stage_phase_a(0, pos, t0, t1, nst, kst);
        } else if (sp.G > 0) {
            /* ST_PROJECT_INPUT: gather (every CTA for itself) -> input projection rows -> x (fp32, like the step's input) */
            __syncthreads();
            if (gstep == 0) stage_vec(sm, sp.hidden, sp.Hin);
            else {
                const uint16_t *row = (gstep == 1 ? sp.emb0 : sp.emb[gstep - 2]) + (size_t)s_codes[gstep - 1] * sp.Hin;
                for (int i = tid; i < sp.Hin; i += CL_THREADS) sm[i] = bf16_to_f32(row[i]);
                __syncthreads();
            }
            if (sp.in_w) {
                s.n = 1; s.w[0] = sp.in_w; s.out[0] = a.x; s.rows[0] = H; cols = sp.Hin;
                cta_gemv(s, cols, sm, sp.in_b, cta, ncta);
            } else if (cta == 0) {
                for (int i = tid; i < H; i += CL_THREADS) a.x[i] = i < sp.Hin ? sm[i] : 0.0f;
            }
            head_w = gstep >= 1 ? sp.heads[gstep - 1] : NULL;
            cl_barrier_arrive(a.bar, target, ncta);
            if (att_cluster) cl_stream_prefetch(s0, s1, 2, HS, qkv_rows(lm.layer[0]));
            stage_phase_a(0, pos, t0, t1, nst, kst);
            cl_barrier_wait(a.bar, target, a.err);
        }
        for (int l = 0; l < L; l++) {
            const LmLayerDev &w = lm.layer[l];
            long long *xin = a.xacc + (size_t)(p % 3) * H, *xmid = a.xacc + (size_t)((p + 1) % 3) * H, *xout = a.xacc + (size_t)((p + 2) % 3) * H;
            /* ================= phase A: reads x (fp32 input at l == 0, else xin), adds the attention block's output into xmid, zeroes xout */
            PROF(0);
            if (tid < zper && z0 + tid < H) xout[z0 + tid] = 0;
            if (att_cluster) {
                asm volatile("cp.async.wait_all;" ::: "memory");       /* this thread's share of nin (and of the K/V rows, aux) has landed */
                XRegs xr;
#if CL_STAGE_AUX
                cl_x_load(xr, l == 0 ? x_in : NULL, xin, nin, H);
                cl_stage_async(npost, w.post_norm, H);
                asm volatile("cp.async.commit_group;");
#else
                cl_x_load(xr, l == 0 ? x_in : NULL, xin, w.in_norm, H);
                if (tid < 128) {                                        /* read on demand (the variant without staging) */
                    const int k = tid >> 5, c = tid & 31;
                    const float *src = k == 0 ? a.cos + (size_t)pos * 128 : k == 1 ? a.sin + (size_t)pos * 128 : k == 2 ? w.q_norm : w.k_norm;
                    *reinterpret_cast<float4 *>(aux + k * 128 + c * 4) = *reinterpret_cast<const float4 *>(src + c * 4);
                }
#endif
                cl_norm_finish(sm, xr, H, lm.eps, red);
                PROF(1);
                {
                    cl_stream(s0, s1, 2, HS, sm, qkv_rows(w), [&](int k, float va, float vb) {
                        if (lane < CL_SIZE) { float *peer = cl.map_shared_rank(qkv, lane); peer[ja + 32 * k] = va; peer[jb + 32 * k] = vb; }
                    });
                }
                PROF(2);
                asm volatile("cp.async.wait_all;" ::: "memory");
                cl.sync();
                PROF(3);
                if (small_att) { if (warp < 2) cl_attention_small(a, lm, l, g, pos, warp, warp == 0 && rank == 0, qkv, att, kvs, aux, wpart + warp * 256); }
                else cl_attention(cl, a, lm, l, g, pos, t0, t1, nst, !full_att || rank == 0, qkv, wpart, cpart, full_att ? att : NULL, kvs, aux);
                /* this CTA's rows of the O partial: RPC rows x 256 bf16; the first eight per warp are requested before the combine */
                const uint16_t *wo = w.wo_t + ((size_t)g * H + (size_t)rank * RPC) * 256;
                uint4 ow[8];
#pragma unroll
                for (int i = 0; i < 8; i++) ow[i] = ldg_stream(reinterpret_cast<const uint4 *>(wo + (size_t)(warp * 8 + i) * 256) + lane);
                PROF(4);
                if (small_att || full_att) __syncthreads();             /* att[] is this CTA's own: no cluster barrier */
                else cl.sync();
                PROF(5);
                if (!small_att && !full_att) cl_attention_combine(cpart, att);
                if (a.flags & 8) cl.sync();                             /* experiment: the third cluster barrier of the push variant */
                PROF(6);
                const float4 a0v = reinterpret_cast<const float4 *>(att)[2 * lane], a1v = reinterpret_cast<const float4 *>(att)[2 * lane + 1];
                uint4 ow2[8];                                           /* H = 2048: rows 128..255 of this CTA, requested before the first eight are reduced */
#pragma unroll
                for (int i = 0; i < 8; i++)
                    ow2[i] = CL_PIPE && RPC > 8 * CL_WARPS ? ldg_stream(reinterpret_cast<const uint4 *>(wo + (size_t)(8 * CL_WARPS + warp * 8 + i) * 256) + lane) : make_uint4(0, 0, 0, 0);
                auto o_batch = [&](const uint4 (&wv)[8], int rb) {
                    long long resid = 0;
                    if (g == 0 && lane < 8) {
                        const int row = rank * RPC + rb + warp * 8 + lane;
                        resid = l == 0 ? to_fx(__ldcg(x_in + row)) : __ldcg(xin + row);
                    }
                    float acc[8];
#pragma unroll
                    for (int i = 0; i < 8; i++) acc[i] = dot8f(wv[i], a0v, a1v, 0.0f);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        float t8[8];
#pragma unroll
                        for (int i = 0; i < 8; i++) t8[i] = __shfl_xor_sync(0xffffffffu, acc[i], o);
#pragma unroll
                        for (int i = 0; i < 8; i++) acc[i] += t8[i];
                    }
                    if (rb == 0) PROF(23);
#pragma unroll
                    for (int i = 0; i < 8; i++)
                        if (lane == i && !(a.flags & 32)) red_add_s64(xmid + rank * RPC + rb + warp * 8 + i, to_fx(acc[i]) + resid);
                };
                PROF(20);
                o_batch(ow, 0);
                PROF(21);
                if (RPC > 8 * CL_WARPS) {
#if !CL_PIPE
#pragma unroll
                    for (int i = 0; i < 8; i++) ow2[i] = ldg_stream(reinterpret_cast<const uint4 *>(wo + (size_t)(8 * CL_WARPS + warp * 8 + i) * 256) + lane);
#endif
                    o_batch(ow2, 8 * CL_WARPS);
                }
            }
#if CL_STAGE_AUX
            else {
                cl_stage_async(npost, w.post_norm, H);
                asm volatile("cp.async.commit_group;");
            }
#endif
            if (!att_cluster && a.pf_frac > 0) {
                /* clusters without a KV head: pull (the first pf_frac percent of) every CTA's slice of this layer's MLP weights into L2
                 * while the attention block runs; consumer CTAs are dealt out over the idle CTAs */
                const int nidle = ncta - lm.kv_heads * CL_SIZE, me = cta - lm.kv_heads * CL_SIZE;
                int slot = 0;
                for (int cons = me; cons < ncta; cons += nidle) {
                    const int cc0 = (int)(((unsigned)NCH * (unsigned)cons) / (unsigned)ncta), cc1 = (int)(((unsigned)NCH * (unsigned)(cons + 1)) / (unsigned)ncta);
                    const size_t rows = (size_t)((cc1 - cc0) * 8 * a.pf_frac / 100), chunks = (size_t)((cc1 - cc0) * a.pf_frac / 100);
                    cl_prefetch_paced(w.gate + (size_t)cc0 * 8 * H, rows * H * 2, (unsigned)a.pf_gap, slot);
                    cl_prefetch_paced(w.up + (size_t)cc0 * 8 * H, rows * H * 2, (unsigned)a.pf_gap, slot);
                    cl_prefetch_paced(w.down_t + (size_t)cc0 * H * 8, chunks * H * 16, (unsigned)a.pf_gap, slot);
                }
            }
            PROF(7);
            cl_barrier_arrive(a.bar, target, ncta);
            cl_stream_prefetch(s0, s1, gu_pairs, HS, gu_rows(w));
            cl_barrier_wait(a.bar, target, a.err);
            p++;
            PROF(8);
            /* ================= phase B: reads xmid, adds the MLP's output into xout, zeroes xin */
            if (tid < zper && z0 + tid < H) xin[z0 + tid] = 0;
            if (att_cluster && l + 1 < L && !(a.flags & 2)) {
                /* next layer's attention-block weights of this CTA -> L2 while the MLP phase runs (its own weights came in during phase A) */
                const LmLayerDev &wn = lm.layer[l + 1];
                const int j0 = 64 * rank;                               /* rows j0..j0+63 of the cluster's QKV block lie inside one of its four 128-row parts */
                cl_prefetch_slice(cl_qkv_row(wn, g, j0, H), (size_t)64 * H * 2, 0, 1);
                cl_prefetch_slice(wn.wo_t + ((size_t)g * H + (size_t)rank * RPC) * 256, (size_t)RPC * 512, 0, 1);
            }
#if CL_STAGE_AUX
            asm volatile("cp.async.wait_all;" ::: "memory");
            cl_stage_norm(sm, NULL, xmid, npost, H, lm.eps, red);
#else
            cl_stage_norm(sm, NULL, xmid, w.post_norm, H, lm.eps, red);
#endif
            PROF(9);
            cl_stream(s0, s1, gu_pairs, HS, sm, gu_rows(w), [&](int k, float va, float vb) {
                if (lane == 0) gs[warp + CL_WARPS * k] = swiglu_act<QTC_FAST>(va, vb);
            });
            {
                /* down partial of row r = sum over this CTA's columns of W_down[r][.] g[.]; pieces (chunk c, row r) are 16 bytes at
                 * (c H + r).  A batch = rows (r, r + 512) x 4 chunks = 8 loads in a Stage; batches: row pairs x chunk groups, two in flight;
                 * the first two are requested before the SwiGLU values are complete (they do not depend on them). */
                const uint4 *dw = reinterpret_cast<const uint4 *>(w.down_t) + (size_t)c0 * H;
                const int CB = NC > 4 ? 2 : 1, nb = HS * CB;                 /* H = 1024: one row pair per thread; 2048: two */
                auto d_issue = [&](Stage &st, int bi) {
                    const int r = tid + 2 * CL_THREADS * (bi / CB), cb = 4 * (bi % CB);
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        st.a[c] = cb + c < NC ? ldg_stream(dw + (size_t)(cb + c) * H + r) : make_uint4(0, 0, 0, 0);
                        st.b[c] = cb + c < NC ? ldg_stream(dw + (size_t)(cb + c) * H + r + CL_THREADS) : make_uint4(0, 0, 0, 0);
                    }
                };
                float sa = 0.0f, sb = 0.0f;
                /* every CTA adds its partial of every row straight into the fixed-point accumulator: exact integer adds, so the sum does
                 * not depend on the order in which the 120 CTAs arrive */
                long long res0 = 0, res1 = 0, res2 = 0, res3 = 0;       /* the residual rows this CTA contributes (cluster 0: rank r owns rows [r RPC, (r+1) RPC)) */
                if (g == 0) {
                    const int r0 = tid, r1 = tid + CL_THREADS, r2 = tid + 2 * CL_THREADS, r3 = tid + 3 * CL_THREADS;
                    if (r0 / RPC == rank) res0 = __ldcg(xmid + r0);
                    if (r1 / RPC == rank) res1 = __ldcg(xmid + r1);
                    if (r2 < H && r2 / RPC == rank) res2 = __ldcg(xmid + r2);
                    if (r3 < H && r3 / RPC == rank) res3 = __ldcg(xmid + r3);
                }
                auto d_consume = [&](const Stage &st, int bi) {
                    const int r = tid + 2 * CL_THREADS * (bi / CB), cb = 4 * (bi % CB);
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const float4 g0 = reinterpret_cast<const float4 *>(gs)[2 * (cb + c)], g1 = reinterpret_cast<const float4 *>(gs)[2 * (cb + c) + 1];
                        sa = dot8f(st.a[c], g0, g1, sa);
                        sb = dot8f(st.b[c], g0, g1, sb);
                    }
                    if (bi % CB == CB - 1) {
                        red_add_s64(xout + r, to_fx(sa) + (bi / CB == 0 ? res0 : res2));
                        red_add_s64(xout + r + CL_THREADS, to_fx(sb) + (bi / CB == 0 ? res1 : res3));
                        sa = 0.0f; sb = 0.0f;
                    }
                };
                d_issue(s0, 0);
                if (nb > 1) d_issue(s1, 1); else st_zero(s1);
                __syncthreads();                                        /* gs complete */
                PROF(10);
                for (int bi = 0; bi < nb; bi += 2) {
                    d_consume(s0, bi);
                    if (bi + 2 < nb) d_issue(s0, bi + 2);
                    if (bi + 1 < nb) { d_consume(s1, bi + 1); if (bi + 3 < nb) d_issue(s1, bi + 3); }
                }
            }
            PROF(11);
            PROF(13);
            cl_barrier_arrive(a.bar, target, ncta);
            if (l + 1 < L && att_cluster) cl_stream_prefetch(s0, s1, 2, HS, qkv_rows(lm.layer[l + 1]));
            if (l + 1 < L) stage_phase_a(l + 1, pos, t0, t1, nst, kst);
            cl_barrier_wait(a.bar, target, a.err);
            p++;
            PROF(14);
        }
        /* final norm of the stack's output (buf[p % 3]) into shared memory; talker: CTA 0 publishes it as the hidden state */
        if (sp.G == 0 || head_w) cl_stage_norm(sm, NULL, a.xacc + (size_t)(p % 3) * H, lm.final_norm, H, lm.eps, red);
        if (sp.G == 0 && cta == 0) {
            for (int i = tid; i < H; i += CL_THREADS) a.x[i] = sm[i];
            if (tid == 0 && a.st && !a.st->done) a.st->pos += 1;
        }
        if (head_w) {
            s.n = 1; s.w[0] = head_w; s.out[0] = a.logits; s.rows[0] = sp.G > 0 ? sp.vocab : a.head_rows; cols = H;
            cta_gemv(s, cols, sm, NULL, cta, ncta);
        }
        if (sp.G > 0 && gstep >= 1) {
            cl_barrier_arrive(a.bar, target, ncta);
            cl_barrier_wait(a.bar, target, a.err);
            const int V = sp.vocab;
            float *lg = sm, *vv = sm + V;
            for (int i = tid; i < V; i += CL_THREADS) lg[i] = __ldcg(a.logits + i);
            if (gstep == 1 && tid == 0) s_rng = sp.seed_bits;
            __syncthreads();
            /* flags & 64 (diagnostic): a fixed code instead of the sample, to time the sampler's share of the frame */
            const int tok = (a.flags & 64) ? 1 : (a.flags & 128) ? block_sample(lg, vv, V, sp.cfg.top_k, sp.cfg.top_p, sp.cfg.temperature, &s_rng, S_samp, sp.scratch_f, sp.scratch_i) : block_sample_thr(lg, vv, V, sp.cfg.top_k, sp.cfg.top_p, sp.cfg.temperature, &s_rng, S_samp, sp.scratch_f, sp.scratch_i, a.prof ? a.prof + cta * 32 + 16 : NULL);
            if (tid == 0) { s_codes[gstep] = tok; if (cta == 0) sp.codes[gstep] = tok; }
        }
    }
#undef PROF
    if (cta == 0 && tid == 0) *a.phase = p % 3;
}
