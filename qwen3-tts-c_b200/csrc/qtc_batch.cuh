/*
 * qtc_batch.cuh - lock-step decode of B <= 64 independent utterances on one GPU (BASELINE configs 4-5): the
 * reference's frame loop (c/qwen_tts.c:1282-1373) with every per-utterance vector turned into a [B, .] matrix.
 * Projections run as tcgen05 GEMMs (qtc_gemm_tc.cu: one pass over the bf16 weights serves the whole batch),
 * everything else (RMSNorm, per-head QK-norm + RoPE + KV append + GQA attention, SwiGLU, gathers, sampling,
 * next-input embedding) as small batched kernels with the utterance index in blockIdx.y.
 *
 * Activations enter the GEMMs as `pieces` bf16 pieces per value: 3 (default) splits every fp32 value exactly
 * (hi + mid + lo), so the products are exact and the results differ from the batch-1 fast path only by fp32
 * summation order - sampled codes then track the C path like batch 1 does; 1 rounds to bf16 (logits within 1e-2
 * relative, BASELINE north_star), one third of the tensor work.
 *
 * Utterances advance together: prompts are right-aligned (utterance b starts `maxlen - len_b` steps late, its
 * position is negative until then and every kernel skips it), every frame all B rows take one talker step and one
 * 15-codebook sub-talker loop.  State per utterance (position, frame count, RNG, done flag) lives in HBM so that
 * the frame is two CUDA graphs replayed by the host.
 */
/* (included at the end of qtc_lm.cu: it shares that translation unit's kernels - sampler, next-input embedding) */
#pragma once

#define BLAUNCH(c)                                                                  \
    do {                                                                            \
        (c)->launches++;                                                            \
        cudaError_t e_ = cudaGetLastError();                                        \
        if (e_ != cudaSuccess) { qtc_set_error("%s:%d launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); return -1; } \
    } while (0)

#define BROWS 64            /* rows per activation piece in the GEMM staging buffer (TC_BN) */

int qtc_gemm_tc(qtc_ctx *c, float *Y, int ldy, const uint16_t *X, const uint16_t *W, int nbatch, int N, int K, int pieces);
int qtc_gemm_tc_ex(qtc_ctx *c, float *Y, int ldy, const uint16_t *X, const uint16_t *W, int nbatch, int N, int K, int pieces, uint16_t *swiglu_out);
void qtc_gemm_tc_forget(const void *ptr);

/* value -> `pieces` bf16 pieces at xb[(p*64 + b) * K + i] */
__device__ __forceinline__ void store_pieces(uint16_t *xb, int b, int i, int K, int pieces, float v) {
    for (int p = 0; p < pieces; p++) {
        const uint32_t u = __float_as_uint(v), h = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
        xb[((size_t)p * BROWS + b) * K + i] = (uint16_t)h;
        v -= __uint_as_float(h << 16);
    }
}

/* two values -> `pieces` packed bf16x2 words at xb[(p*64 + b) * K + i], i even */
__device__ __forceinline__ void store_pieces2(uint16_t *xb, int b, int i, int K, int pieces, float v0, float v1) {
    for (int p = 0; p < pieces; p++) {
        const uint32_t u0 = __float_as_uint(v0), h0 = (u0 + 0x7fffu + ((u0 >> 16) & 1u)) >> 16;
        const uint32_t u1 = __float_as_uint(v1), h1 = (u1 + 0x7fffu + ((u1 >> 16) & 1u)) >> 16;
        *reinterpret_cast<uint32_t *>(xb + ((size_t)p * BROWS + b) * K + i) = h0 | (h1 << 16);
        v0 -= __uint_as_float(h0 << 16); v1 -= __uint_as_float(h1 << 16);
    }
}

/* RMSNorm of row b of x [B, n] -> bf16 pieces (GEMM input), optionally the fp32 normed row too (out_f32 [B, n]);
 * also clears zero_n floats of row b of `zero` (the next GEMM's output, which accumulates).  pos[b] < 0: inactive.
 * n % 4 == 0; rows up to 256 x 4 x 4 = 4096 values stay in registers between the two passes. */
__global__ void __launch_bounds__(256) k_b_norm(uint16_t *xb, float *out_f32, const float *x, const float *w, int n, float eps, int pieces,
                                                float *zero, int zero_n, const int *pos) {
    pdl_enter();
    __shared__ float red[32];
    const int b = blockIdx.x, tid = threadIdx.x;
    if (zero) {
        float4 *z = reinterpret_cast<float4 *>(zero + (size_t)b * zero_n);
        for (int i = tid; i < (zero_n >> 2); i += 256) z[i] = make_float4(0, 0, 0, 0);
        for (int i = (zero_n & ~3) + tid; i < zero_n; i += 256) zero[(size_t)b * zero_n + i] = 0.0f;
    }
    const float4 *xi = reinterpret_cast<const float4 *>(x + (size_t)b * n);
    const float4 *wi = reinterpret_cast<const float4 *>(w);
    const int n4 = n >> 2;
    float4 v[4], g[4];
    float ss = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int i = tid + 256 * j;
        v[j] = make_float4(0, 0, 0, 0); g[j] = v[j];
        if (i < n4) { v[j] = xi[i]; g[j] = wi[i]; }
        ss = fmaf(v[j].x, v[j].x, ss); ss = fmaf(v[j].y, v[j].y, ss); ss = fmaf(v[j].z, v[j].z, ss); ss = fmaf(v[j].w, v[j].w, ss);
    }
    for (int i = tid + 1024; i < n4; i += 256) { const float4 t = xi[i]; ss = fmaf(t.x, t.x, ss); ss = fmaf(t.y, t.y, ss); ss = fmaf(t.z, t.z, ss); ss = fmaf(t.w, t.w, ss); }
    ss = block_sum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)n + eps);
    const bool live = !pos || pos[b] >= 0;
    auto emit = [&](const float4 t, const float4 gw, int i) {
        float4 y = make_float4(t.x * inv * gw.x, t.y * inv * gw.y, t.z * inv * gw.z, t.w * inv * gw.w);
        if (!live) y = make_float4(0, 0, 0, 0);
        store_pieces2(xb, b, 4 * i, n, pieces, y.x, y.y);
        store_pieces2(xb, b, 4 * i + 2, n, pieces, y.z, y.w);
        if (out_f32) reinterpret_cast<float4 *>(out_f32 + (size_t)b * n)[i] = y;
    };
#pragma unroll
    for (int j = 0; j < 4; j++) { const int i = tid + 256 * j; if (i < n4) emit(v[j], g[j], i); }
    for (int i = tid + 1024; i < n4; i += 256) emit(xi[i], wi[i], i);
}

/* fp32 rows [B, n] -> bf16 pieces; optionally clears the next GEMM's output rows */
__global__ void k_b_pieces(uint16_t *xb, const float *src, int n, int pieces, float *zero, int zero_n) {
    pdl_enter();
    const int b = blockIdx.y;
    if (zero) for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < zero_n; i += gridDim.x * blockDim.x) zero[(size_t)b * zero_n + i] = 0.0f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) store_pieces(xb, b, i, n, pieces, src[(size_t)b * n + i]);
}

/* SwiGLU of gu [B, 2I] (gate/up interleaved: 2u, 2u+1) -> bf16 pieces [.., I] */
__global__ void k_b_swiglu(uint16_t *xb, const float *gu, int I, int pieces) {
    pdl_enter();
    const int b = blockIdx.y;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < I; i += gridDim.x * blockDim.x) {
        const float2 v = *reinterpret_cast<const float2 *>(gu + (size_t)b * 2 * I + 2 * i);
        const float g = v.x, u = v.y;
        store_pieces(xb, b, i, I, pieces, swiglu_act<QTC_FAST>(g, u));
    }
}

struct BAttnArgs {
    const float *qkv;              /* [B, qd + 2 kvd]: q | k | v of the new token */
    const float *q_norm, *k_norm, *cos, *sin;
    float *kc, *vc;                /* this layer's cache: the pool slice (paged: tables != NULL) or [B][..][kv_max, kvd] with utterance stride kv_stride */
    size_t kv_stride;
    const int *tables; int table_stride;   /* page tables [B][table_stride] of the utterances, or NULL */
    uint16_t *xb;                  /* out: attention output as bf16 pieces [.., qd] */
    const int *pos; int pos_add;   /* position of utterance b = pos[b] + pos_add (pos == NULL: pos_add) */
    int heads, kv_heads, hd, pieces;
    float eps, scale;
};
/* talker_attention_single steps 3-6 (c/qwen_tts_talker.c:150-225) for utterance blockIdx.y, Q head blockIdx.x.
 * dynamic smem: 2 hd + (pos + 1) floats */
__global__ void __launch_bounds__(128) k_b_attn(const BAttnArgs a) {
    pdl_enter();
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[32];
    const int hd = a.hd, half = hd >> 1, b = blockIdx.y, h = blockIdx.x;
    const int rep = a.heads / a.kv_heads, g = h / rep, kvd = a.kv_heads * hd, qd = a.heads * hd;
    const int pos = (a.pos ? a.pos[b] : 0) + a.pos_add;
    if (pos < 0) { for (int i = threadIdx.x; i < hd; i += blockDim.x) store_pieces(a.xb, b, h * hd + i, qd, a.pieces, 0.0f); return; }
    float *qs = sm, *ks = sm + hd, *sc = sm + 2 * hd;
    const float *row = a.qkv + (size_t)b * (qd + 2 * kvd), *qr = row + h * hd, *kr = row + qd + g * hd, *vr = row + qd + kvd + g * hd;
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;
    const int *tab = a.tables ? a.tables + (size_t)b * a.table_stride : NULL;
    float *kc = a.kc + (tab ? 0 : (size_t)b * a.kv_stride), *vc = a.vc + (tab ? 0 : (size_t)b * a.kv_stride);
    float s1 = 0.0f, s2 = 0.0f;
    for (int i = threadIdx.x; i < hd; i += blockDim.x) { s1 = fmaf(qr[i], qr[i], s1); s2 = fmaf(kr[i], kr[i], s2); }
    s1 = block_sum(s1, red);
    s2 = block_sum(s2, red);
    const float qinv = 1.0f / sqrtf(s1 / (float)hd + a.eps), kinv = 1.0f / sqrtf(s2 / (float)hd + a.eps);
    for (int i = threadIdx.x; i < hd; i += blockDim.x) { qs[i] = qr[i] * qinv * a.q_norm[i]; ks[i] = kr[i] * kinv * a.k_norm[i]; }
    __syncthreads();
    float rq0 = 0, rq1 = 0, rk0 = 0, rk1 = 0;
    for (int i0 = threadIdx.x; i0 < half; i0 += blockDim.x) {          /* hd <= 256: one trip */
        const float qa = qs[i0], qb = qs[i0 + half], ka = ks[i0], kb = ks[i0 + half];
        rq0 = qa * cs[i0] - qb * sn[i0]; rq1 = qb * cs[i0 + half] + qa * sn[i0 + half];
        rk0 = ka * cs[i0] - kb * sn[i0]; rk1 = kb * cs[i0 + half] + ka * sn[i0 + half];
    }
    __syncthreads();
    if (threadIdx.x < half) { qs[threadIdx.x] = rq0; qs[threadIdx.x + half] = rq1; ks[threadIdx.x] = rk0; ks[threadIdx.x + half] = rk1; }
    __syncthreads();
    if (h % rep == 0)
        for (int i = threadIdx.x; i < hd; i += blockDim.x) { kc[kv_off(tab, pos, kvd) + g * hd + i] = ks[i]; vc[kv_off(tab, pos, kvd) + g * hd + i] = vr[i]; }
    const int n = pos + 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int t = warp; t < n; t += nw) {
        const float *kt = t < pos ? kc + kv_off(tab, t, kvd) + g * hd : ks;
        float d = 0.0f;
        for (int i = lane; i < hd; i += 32) d = fmaf(qs[i], kt[i], d);
        d = warp_sum(d);
        if (lane == 0) sc[t] = d * a.scale;
    }
    __syncthreads();
    float mx = -3.402823466e+38f;
    for (int t = threadIdx.x; t < n; t += blockDim.x) mx = fmaxf(mx, sc[t]);
    mx = block_max(mx, red);
    float tot = 0.0f;
    for (int t = threadIdx.x; t < n; t += blockDim.x) { const float e = __expf(sc[t] - mx); sc[t] = e; tot += e; }
    tot = block_sum(tot, red);
    const float inv = 1.0f / tot;
    __syncthreads();
    for (int i = threadIdx.x; i < hd; i += blockDim.x) {
        float o = 0.0f, o1 = 0.0f;
        const float *vcol = vc + g * hd + i;
        int t = 0;
        for (; t + 1 < pos; t += 2) { o = fmaf(sc[t], vcol[kv_off(tab, t, kvd)], o); o1 = fmaf(sc[t + 1], vcol[kv_off(tab, t + 1, kvd)], o1); }
        for (; t < pos; t++) o = fmaf(sc[t], vcol[kv_off(tab, t, kvd)], o);
        o = fmaf(sc[pos], vr[i], o + o1);
        store_pieces(a.xb, b, h * hd + i, qd, a.pieces, o * inv);
    }
}

/* The same for head_dim 128 and two query heads per KV head (every shipped talker), one CTA of 8 warps per (KV head, utterance):
 * a warp takes positions w, w + 8, ..., lane i owns dims 4i..4i+3, and each K / V row it loads (one 512-byte request per row
 * instead of 128 scalar-strided ones) serves BOTH query heads of the group - k_b_attn reads every row once per query head and
 * walks the V rows one position per dependent FMA.  Scores, running maximum, sum and output are kept per warp (online softmax)
 * and folded over the 8 warps at the end.  Differs from k_b_attn by summation order only. */
#define BA2_WARPS 8
#define BA2_REC 132
__device__ __forceinline__ float4 ba2_norm_rope(const float *src, const float *nw, const float *cs, const float *sn, float eps, int lane) {
    const float4 v = reinterpret_cast<const float4 *>(src)[lane], n4 = reinterpret_cast<const float4 *>(nw)[lane];
    const float4 c4 = reinterpret_cast<const float4 *>(cs)[lane], s4 = reinterpret_cast<const float4 *>(sn)[lane];
    float ss = v.x * v.x;
    ss = fmaf(v.y, v.y, ss); ss = fmaf(v.z, v.z, ss); ss = fmaf(v.w, v.w, ss);
    ss = warp_sum(ss);
    const float inv = 1.0f / sqrtf(ss / 128.0f + eps);
    const float4 y = make_float4(v.x * inv * n4.x, v.y * inv * n4.y, v.z * inv * n4.z, v.w * inv * n4.w);
    const float px = __shfl_xor_sync(0xffffffffu, y.x, 16), py = __shfl_xor_sync(0xffffffffu, y.y, 16);
    const float pz = __shfl_xor_sync(0xffffffffu, y.z, 16), pw = __shfl_xor_sync(0xffffffffu, y.w, 16);
    /* rotate-half: dims < 64 (lanes < 16): y cos - partner sin; dims >= 64: y cos + partner sin */
    return lane < 16 ? make_float4(y.x * c4.x - px * s4.x, y.y * c4.y - py * s4.y, y.z * c4.z - pz * s4.z, y.w * c4.w - pw * s4.w)
                     : make_float4(y.x * c4.x + px * s4.x, y.y * c4.y + py * s4.y, y.z * c4.z + pz * s4.z, y.w * c4.w + pw * s4.w);
}
__global__ void __launch_bounds__(BA2_WARPS * 32) k_b_attn_gqa2(const BAttnArgs a) {
    pdl_enter();
    __shared__ __align__(16) float wpart[BA2_WARPS][2][BA2_REC];
    const int g = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kvd = a.kv_heads * 128, qd = a.heads * 128;
    const int pos = (a.pos ? a.pos[b] : 0) + a.pos_add;
    if (pos < 0) { store_pieces(a.xb, b, g * 256 + tid, qd, a.pieces, 0.0f); return; }
    const float *row = a.qkv + (size_t)b * (qd + 2 * kvd), *cs = a.cos + (size_t)pos * 128, *sn = a.sin + (size_t)pos * 128;
    const int *tab = a.tables ? a.tables + (size_t)b * a.table_stride : NULL;
    float *kc = a.kc + (tab ? 0 : (size_t)b * a.kv_stride) + g * 128, *vc = a.vc + (tab ? 0 : (size_t)b * a.kv_stride) + g * 128;
    const int n = pos + 1;
    float m0 = -3.402823466e+38f, m1 = m0, l0 = 0.0f, l1 = 0.0f;
    float4 o0 = make_float4(0, 0, 0, 0), o1 = o0;
    if (warp < n) {
        const float4 q0 = ba2_norm_rope(row + (2 * g) * 128, a.q_norm, cs, sn, a.eps, lane), q1 = ba2_norm_rope(row + (2 * g + 1) * 128, a.q_norm, cs, sn, a.eps, lane);
        float4 k_new = make_float4(0, 0, 0, 0), v_new = k_new;
        if (((n - 1) % BA2_WARPS) == warp) {             /* the warp that takes the new position also appends it to the cache */
            k_new = ba2_norm_rope(row + qd + g * 128, a.k_norm, cs, sn, a.eps, lane);
            v_new = reinterpret_cast<const float4 *>(row + qd + kvd + g * 128)[lane];
            const size_t off = kv_off(tab, pos, kvd);
            *reinterpret_cast<float4 *>(kc + off + lane * 4) = k_new;
            *reinterpret_cast<float4 *>(vc + off + lane * 4) = v_new;
        }
        for (int tb = warp; tb < n; tb += 4 * BA2_WARPS) {
            float d0[4], d1[4];
            float4 vr[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int t = tb + BA2_WARPS * j;
                float4 kr = make_float4(0, 0, 0, 0);
                vr[j] = kr;
                if (t < pos) {
                    const size_t off = kv_off(tab, t, kvd);
                    kr = *reinterpret_cast<const float4 *>(kc + off + lane * 4);
                    vr[j] = *reinterpret_cast<const float4 *>(vc + off + lane * 4);
                } else if (t == pos) { kr = k_new; vr[j] = v_new; }
                d0[j] = fmaf(q0.w, kr.w, fmaf(q0.z, kr.z, fmaf(q0.y, kr.y, q0.x * kr.x)));
                d1[j] = fmaf(q1.w, kr.w, fmaf(q1.z, kr.z, fmaf(q1.y, kr.y, q1.x * kr.x)));
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
            for (int j = 0; j < 4; j++) if (tb + BA2_WARPS * j < n) { d0[j] *= a.scale; d1[j] *= a.scale; b0 = fmaxf(b0, d0[j]); b1 = fmaxf(b1, d1[j]); }
            const float f0 = __expf(m0 - b0), f1 = __expf(m1 - b1);      /* first batch: exp(-huge) = 0 on zeros */
            l0 *= f0; o0.x *= f0; o0.y *= f0; o0.z *= f0; o0.w *= f0;
            l1 *= f1; o1.x *= f1; o1.y *= f1; o1.z *= f1; o1.w *= f1;
            m0 = b0; m1 = b1;
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (tb + BA2_WARPS * j < n) {
                    const float e0 = __expf(d0[j] - m0), e1 = __expf(d1[j] - m1);
                    l0 += e0; l1 += e1;
                    o0.x = fmaf(e0, vr[j].x, o0.x); o0.y = fmaf(e0, vr[j].y, o0.y); o0.z = fmaf(e0, vr[j].z, o0.z); o0.w = fmaf(e0, vr[j].w, o0.w);
                    o1.x = fmaf(e1, vr[j].x, o1.x); o1.y = fmaf(e1, vr[j].y, o1.y); o1.z = fmaf(e1, vr[j].z, o1.z); o1.w = fmaf(e1, vr[j].w, o1.w);
                }
        }
    }
    if (lane == 0) { wpart[warp][0][0] = m0; wpart[warp][0][1] = l0; wpart[warp][1][0] = m1; wpart[warp][1][1] = l1; }
    *reinterpret_cast<float4 *>(&wpart[warp][0][4 + lane * 4]) = o0;
    *reinterpret_cast<float4 *>(&wpart[warp][1][4 + lane * 4]) = o1;
    __syncthreads();
    {
        const int hh = tid >> 7, d = tid & 127;              /* 256 threads: (query head of the group, dim) */
        float M = -3.402823466e+38f;
#pragma unroll
        for (int k = 0; k < BA2_WARPS; k++) M = fmaxf(M, wpart[k][hh][0]);
        float L = 0.0f, o = 0.0f;
#pragma unroll
        for (int k = 0; k < BA2_WARPS; k++) {
            const float lw = wpart[k][hh][1], f = lw > 0.0f ? __expf(wpart[k][hh][0] - M) : 0.0f;
            L = fmaf(lw, f, L);
            o = fmaf(wpart[k][hh][4 + d], f, o);
        }
        store_pieces(a.xb, b, (2 * g + hh) * 128 + d, qd, a.pieces, o / L);
    }
}

/* the same for short contexts (n = pos + 1 <= 32, head_dim <= 128: every sub-talker step): a warp per (utterance, head),
 * lane i owns dims 4i..4i+3, eight K/V rows in flight per lane, no block barrier.  grid (ceil(heads / 4), B), 128 threads */
__global__ void __launch_bounds__(128) k_b_attn_small(const BAttnArgs a) {
    pdl_enter();
    __shared__ __align__(16) float scr[4][2][128];
    const int hd = a.hd, half = hd >> 1, b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x * 4 + warp;
    if (h >= a.heads) return;
    const int rep = a.heads / a.kv_heads, g = h / rep, kvd = a.kv_heads * hd, qd = a.heads * hd;
    const int pos = (a.pos ? a.pos[b] : 0) + a.pos_add, n = pos + 1;
    const bool act = lane * 4 < hd;
    float *qs = scr[warp][0], *ks = scr[warp][1];
    const float *row = a.qkv + (size_t)b * (qd + 2 * kvd), *qr = row + h * hd, *kr = row + qd + g * hd, *vr = row + qd + kvd + g * hd;
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;
    const int *tab = a.tables ? a.tables + (size_t)b * a.table_stride : NULL;
    float *kc = a.kc + (tab ? 0 : (size_t)b * a.kv_stride), *vc = a.vc + (tab ? 0 : (size_t)b * a.kv_stride);
    /* The cached rows do not depend on this step's q: the first 16 K and V rows (every sub-talker context: <= 17 positions, the last
     * one being the new token) are requested before anything else, together with q, k, v and the norm / RoPE vectors - one round
     * trip to L2 instead of five dependent ones (a 10 us kernel, 80 launches per batched frame, was almost all load latency). */
    float4 kpre[16], vpre[16];
#pragma unroll
    for (int t = 0; t < 16; t++) {
        kpre[t] = make_float4(0, 0, 0, 0); vpre[t] = kpre[t];
        if (t < pos && act) {
            const size_t off = kv_off(tab, t, kvd) + g * hd + lane * 4;
            kpre[t] = *reinterpret_cast<const float4 *>(kc + off);
            vpre[t] = *reinterpret_cast<const float4 *>(vc + off);
        }
    }
    float sq = 0.0f, sk = 0.0f;
    for (int i = lane; i < hd; i += 32) { const float vq = qr[i], vk = kr[i]; qs[i] = vq; ks[i] = vk; sq = fmaf(vq, vq, sq); sk = fmaf(vk, vk, sk); }
    const float4 v_new = act ? *reinterpret_cast<const float4 *>(vr + lane * 4) : make_float4(0, 0, 0, 0);
    sq = warp_sum(sq); sk = warp_sum(sk);
    const float iq = 1.0f / sqrtf(sq / (float)hd + a.eps), ik = 1.0f / sqrtf(sk / (float)hd + a.eps);
    __syncwarp();
    for (int i = lane; i < half; i += 32) {
        const float c0 = cs[i], c1 = cs[i + half], s0 = sn[i], s1 = sn[i + half];
        float xa = qs[i] * iq * a.q_norm[i], xb = qs[i + half] * iq * a.q_norm[i + half];
        qs[i] = xa * c0 - xb * s0; qs[i + half] = xb * c1 + xa * s1;
        xa = ks[i] * ik * a.k_norm[i]; xb = ks[i + half] * ik * a.k_norm[i + half];
        ks[i] = xa * c0 - xb * s0; ks[i + half] = xb * c1 + xa * s1;
    }
    __syncwarp();
    const float4 q4 = act ? *reinterpret_cast<const float4 *>(qs + lane * 4) : make_float4(0, 0, 0, 0);
    const float4 k_new = act ? *reinterpret_cast<const float4 *>(ks + lane * 4) : make_float4(0, 0, 0, 0);
    if (act && h % rep == 0) {
        *reinterpret_cast<float4 *>(kc + kv_off(tab, pos, kvd) + g * hd + lane * 4) = k_new;
        *reinterpret_cast<float4 *>(vc + kv_off(tab, pos, kvd) + g * hd + lane * 4) = v_new;
    }
    float sc = -3.402823466e+38f;
#pragma unroll
    for (int t0 = 0; t0 < 32; t0 += 8) {
        if (t0 < n) {
            float d[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int t = t0 + u;
                float4 kv = make_float4(0, 0, 0, 0);
                if (t < pos) {
                    if (t0 < 16) kv = kpre[t0 < 16 ? t : 0];
                    else if (act) kv = *reinterpret_cast<const float4 *>(kc + kv_off(tab, t, kvd) + g * hd + lane * 4);
                } else if (t == pos) kv = k_new;
                d[u] = fmaf(q4.w, kv.w, fmaf(q4.z, kv.z, fmaf(q4.y, kv.y, q4.x * kv.x)));
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                float t8[8];
#pragma unroll
                for (int u = 0; u < 8; u++) t8[u] = __shfl_xor_sync(0xffffffffu, d[u], o);
#pragma unroll
                for (int u = 0; u < 8; u++) d[u] += t8[u];
            }
#pragma unroll
            for (int u = 0; u < 8; u++) if (lane == t0 + u && t0 + u < n) sc = d[u] * a.scale;
        }
    }
    const float mx = warp_max(sc);
    const float e = lane < n ? __expf(sc - mx) : 0.0f;
    const float tot = warp_sum(e);
    float4 o4 = make_float4(0, 0, 0, 0);
#pragma unroll
    for (int t0 = 0; t0 < 32; t0 += 8) {
        if (t0 < n) {
            float4 vv[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int t = t0 + u;
                vv[u] = make_float4(0, 0, 0, 0);
                if (t < pos) {
                    if (t0 < 16) vv[u] = vpre[t0 < 16 ? t : 0];
                    else if (act) vv[u] = *reinterpret_cast<const float4 *>(vc + kv_off(tab, t, kvd) + g * hd + lane * 4);
                } else if (t == pos) vv[u] = v_new;
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const float p = __shfl_sync(0xffffffffu, e, (t0 + u) & 31);
                if (t0 + u < n) { o4.x = fmaf(p, vv[u].x, o4.x); o4.y = fmaf(p, vv[u].y, o4.y); o4.z = fmaf(p, vv[u].z, o4.z); o4.w = fmaf(p, vv[u].w, o4.w); }
            }
        }
    }
    if (act) {
        const float inv = 1.0f / tot;
        store_pieces2(a.xb, b, h * hd + lane * 4, qd, a.pieces, o4.x * inv, o4.y * inv);
        store_pieces2(a.xb, b, h * hd + lane * 4 + 2, qd, a.pieces, o4.z * inv, o4.w * inv);
    }
}

/* ---- small batched helpers --------------------------------------------------------------------------- */
/* x[b] = prefill[b][pos[b]] while the utterance is inside its prompt window (pos[b] >= 0), else 0 */
__global__ void k_b_fill_prefill(float *x, const float *prefill, int rows_cap, const int *pos, int H) {
    pdl_enter();
    const int b = blockIdx.y, p = pos[b];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H; i += gridDim.x * blockDim.x)
        x[(size_t)b * H + i] = p >= 0 ? prefill[((size_t)b * rows_cap + p) * H + i] : 0.0f;
}
__global__ void k_b_copy_rows(float *dst, const float *src, int n) {
    pdl_enter();
    const int b = blockIdx.y;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[(size_t)b * n + i] = src[(size_t)b * n + i];
}
__global__ void k_b_pos_add(int *pos, int n, int add) { pdl_enter(); const int b = blockIdx.x * blockDim.x + threadIdx.x; if (b < n) pos[b] += add; }
/* talker position of every utterance follows its GenState (captured graphs replay without host patching) */
__global__ void k_b_bump_state_pos(GenState *st, int *pos, int n) {
    pdl_enter();
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < n) { if (!st[b].done) st[b].pos += 1; pos[b] = st[b].pos; }
}
/* sub-talker input of code group g (ST_PROJECT_INPUT, c/qwen_tts_talker.c:693-726): the talker hidden state (g = 0) or the
 * bf16 embedding row of the previous code -> fp32 rows [B, Hin] */
__global__ void k_b_sub_input(float *out, const float *hidden, const uint16_t *table, const int *codes, int code_stride, int code_idx, int Hin) {
    pdl_enter();
    const int b = blockIdx.y;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < Hin; i += gridDim.x * blockDim.x)
        out[(size_t)b * Hin + i] = table ? bf16_to_f32(table[(size_t)codes[(size_t)b * code_stride + code_idx] * Hin + i]) : hidden[(size_t)b * Hin + i];
}
/* sub-talker input of position g >= 1 by table (qtc_build_proj_tables: the projected embedding of every code, computed at upload):
 * x[b, :] = ptab[code of utterance b at position g - 1] - one gather instead of gather + pieces + bias rows + a GEMM */
__global__ void k_b_sub_input_tab(float *x, const float *ptab, const int *codes, int code_stride, int code_idx, int h) {
    pdl_enter();
    const int b = blockIdx.y;
    const float4 *src = reinterpret_cast<const float4 *>(ptab + (size_t)codes[(size_t)b * code_stride + code_idx] * h);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < (h >> 2); i += gridDim.x * blockDim.x) reinterpret_cast<float4 *>(x + (size_t)b * h)[i] = src[i];
}
/* y[b, i] = bias[i] (or 0): initial value of a GEMM output that accumulates */
__global__ void k_b_init_rows(float *y, const float *bias, int n) {
    pdl_enter();
    const int b = blockIdx.y;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) y[(size_t)b * n + i] = bias ? bias[i] : 0.0f;
}
__global__ void k_b_copy_pad_rows(float *dst, const float *src, int n_in, int n_out) {
    pdl_enter();
    const int b = blockIdx.y;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_out; i += gridDim.x * blockDim.x) dst[(size_t)b * n_out + i] = i < n_in ? src[(size_t)b * n_in + i] : 0.0f;
}

/* ---- device-resident batch state ----------------------------------------------------------------------- */
struct BStack {                   /* one decoder stack's batched buffers */
    float *x, *qkv, *gu, *logits; /* [B,H] [B,qd+2kvd] [B,2I] [B,vocab] */
    float *kc, *vc;               /* [B][layers][kv_max][kvd] */
    int kv_max;
    /* batch-path weight layouts (copies made on first use): q|k|v rows concatenated, gate/up rows interleaved */
    uint16_t *wqkv[QTC_MAX_LM_LAYERS], *wgu[QTC_MAX_LM_LAYERS];
    uint16_t *xh;                 /* second GEMM input staging [pieces*64, inter]: SwiGLU output of the gate/up epilogue */
};
struct BatchBuf {
    int B;                        /* capacity (utterances) */
    int pieces;
    uint16_t *xb;                 /* GEMM input staging: [pieces*64, maxK] bf16 */
    int maxK;
    BStack tk, st;
    float *hidden;                /* [B,H] post-final-norm talker hidden (ctx->tk_x of every utterance) */
    float *sub_in;                /* [B,H] gathered sub-talker input */
    int *pos;                     /* [B] talker position of the token in flight (negative: not started) */
    GenState *state; GenState *state_host;
    int *hist;                    /* [B, vocab] */
    int *frame_codes;             /* [B, G] */
    int *codes; int codes_cap;    /* [B, codes_cap, G] */
    float *next_embed;            /* [B,H] */
    float *prefill; int prefill_cap;   /* [B, prefill_cap, H] */
    float *trailing; int trailing_cap; /* [B, trailing_cap, H] */
    uint32_t *sub_rng;            /* [B] */
    uint32_t *seeds;              /* [B] bit pattern of (float)(sample_seed + b): RNG seed of utterance b */
    float *pcm_all = NULL; size_t pcm_cap = 0;   /* PCM of up to 64 decoded utterances, [.., max_tokens * 1920], on its way to the host */
    float *qpcm = NULL; size_t qpcm_cap = 0;     /* generate_queue with overlapped decode: PCM of every request [n_utt, max_tokens * 1920] until the call ends */
    int *qcodes = NULL; size_t qcodes_cap = 0;   /* generate_queue: codes of retired utterances [n_utt, limit, G], kept for the decode at the end */
    float *scratch_f; int *scratch_i;  /* slow-path sampler scratch [B, vocab] */
    cudaGraphExec_t g_front, g_talker;
    int graph_key, graph_B;
    long long n_front, n_talker;  /* kernel launches inside each captured graph */
    std::vector<void *> allocs;
    long long mallocs = 0;
};

static void *balloc(BatchBuf *bb, size_t bytes) {
    void *p = NULL;
    bb->mallocs++;
    if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) { qtc_set_error("batch: cudaMalloc of %.1f MB failed", bytes / 1e6); return NULL; }
    cudaMemset(p, 0, bytes ? bytes : 16);
    bb->allocs.push_back(p);
    return p;
}
void qtc_batch_invalidate_graphs(qtc_ctx *c) { if (c->bb) c->bb->graph_key = -1; }
void qtc_batch_release(qtc_ctx *c) {
    BatchBuf *bb = c->bb;
    if (!bb) return;
    cudaStreamSynchronize(c->stream);
    if (bb->g_front) cudaGraphExecDestroy(bb->g_front);
    if (bb->g_talker) cudaGraphExecDestroy(bb->g_talker);
    for (void *p : bb->allocs) { qtc_gemm_tc_forget(p); cudaFree(p); }
    if (bb->state_host) cudaFreeHost(bb->state_host);
    delete bb;
    c->bb = NULL;
}

/* kv_max > 0: the stack keeps its own contiguous cache [B][layers][kv_max][kvd] (sub-talker); 0: the talker's rows live in the
 * paged pool of the context */
static int bstack_alloc(BatchBuf *bb, BStack *s, const LmDev *lm, int B, int kv_max) {
    const size_t qd = (size_t)lm->heads * lm->head_dim, kvd = (size_t)lm->kv_heads * lm->head_dim;
    s->x = (float *)balloc(bb, (size_t)B * lm->hidden * 4);
    s->qkv = (float *)balloc(bb, (size_t)B * (qd + 2 * kvd) * 4);
    s->gu = (float *)balloc(bb, (size_t)B * 2 * lm->inter * 4);
    s->logits = (float *)balloc(bb, (size_t)B * lm->vocab * 4);
    s->kc = s->vc = NULL;
    if (kv_max > 0) {
        s->kc = (float *)balloc(bb, (size_t)B * lm->layers * kv_max * kvd * 4);
        s->vc = (float *)balloc(bb, (size_t)B * lm->layers * kv_max * kvd * 4);
        if (!s->kc || !s->vc) return -1;
    }
    s->kv_max = kv_max;
    if (!(s->x && s->qkv && s->gu && s->logits)) return -1;
    s->xh = (uint16_t *)balloc(bb, (size_t)bb->pieces * BROWS * lm->inter * 2);
    if (!s->xh) return -1;
    const size_t H = lm->hidden, I = lm->inter;
    for (int l = 0; l < lm->layers; l++) {
        const LmLayerDev *w = &lm->layer[l];
        s->wqkv[l] = (uint16_t *)balloc(bb, (qd + 2 * kvd) * H * 2);
        s->wgu[l] = (uint16_t *)balloc(bb, 2 * I * H * 2);
        if (!s->wqkv[l] || !s->wgu[l]) return -1;
        if (cudaMemcpy(s->wqkv[l], w->wq, qd * H * 2, cudaMemcpyDeviceToDevice) != cudaSuccess ||
            cudaMemcpy(s->wqkv[l] + qd * H, w->wk, kvd * H * 2, cudaMemcpyDeviceToDevice) != cudaSuccess ||
            cudaMemcpy(s->wqkv[l] + (qd + kvd) * H, w->wv, kvd * H * 2, cudaMemcpyDeviceToDevice) != cudaSuccess ||
            cudaMemcpy2D(s->wgu[l], 2 * H * 2, w->gate, H * 2, H * 2, I, cudaMemcpyDeviceToDevice) != cudaSuccess ||
            cudaMemcpy2D(s->wgu[l] + H, 2 * H * 2, w->up, H * 2, H * 2, I, cudaMemcpyDeviceToDevice) != cudaSuccess) {
            qtc_set_error("batch: building the fused weight layouts failed: %s", cudaGetErrorString(cudaGetLastError()));
            return -1;
        }
    }
    return 0;
}

/* (re)build the batch state for B utterances, kv_max talker positions, `frames` frames, prompt / trailing capacities */
static int batch_ensure(qtc_ctx *c, int B, int kv_need, int frames, int pre_rows, int tr_rows) {
    BatchBuf *bb = c->bb;
    const int pieces = c->batch_pieces == 1 ? 1 : 3;
    /* the talker's KV: every row takes pages for kv_need positions from the pool (slots 0..B-1), page by page across the rows, so
     * the tables interleave and no row's cache is contiguous */
    qtc_kv_release_all(c);
    for (int pg = 1; pg <= (kv_need + QTC_KV_PAGE_POS - 1) / QTC_KV_PAGE_POS; pg++)
        for (int b = 0; b < B; b++) if (qtc_kv_reserve(c, b, pg * QTC_KV_PAGE_POS)) return -1;
    if (bb && bb->B >= B && bb->codes_cap >= frames && bb->prefill_cap >= pre_rows && bb->trailing_cap >= tr_rows &&
        bb->pieces == pieces) return 0;
    qtc_batch_release(c);
    bb = new BatchBuf();
    bb->B = B; bb->pieces = pieces; bb->g_front = NULL; bb->g_talker = NULL; bb->graph_key = -1; bb->graph_B = 0; bb->state_host = NULL;
    c->bb = bb;
    const LmDev *tk = &c->talker, *st = &c->sub;
    const int H = c->H, G = c->G;
    int maxK = tk->hidden > tk->inter ? tk->hidden : tk->inter;
    if (tk->heads * tk->head_dim > maxK) maxK = tk->heads * tk->head_dim;
    if (st->inter > maxK) maxK = st->inter;
    if (st->heads * st->head_dim > maxK) maxK = st->heads * st->head_dim;
    bb->maxK = maxK;
    bb->xb = (uint16_t *)balloc(bb, (size_t)pieces * BROWS * maxK * 2);
    if (!bb->xb || bstack_alloc(bb, &bb->tk, tk, B, 0) || bstack_alloc(bb, &bb->st, st, B, G + 2)) return -1;
    bb->hidden = (float *)balloc(bb, (size_t)B * H * 4);
    bb->sub_in = (float *)balloc(bb, (size_t)B * H * 4);
    bb->pos = (int *)balloc(bb, (size_t)B * 4);
    bb->state = (GenState *)balloc(bb, (size_t)B * sizeof(GenState));
    bb->hist = (int *)balloc(bb, (size_t)B * tk->vocab * 4);
    bb->frame_codes = (int *)balloc(bb, (size_t)B * G * 4);
    bb->codes_cap = frames + 8;
    bb->codes = (int *)balloc(bb, (size_t)B * bb->codes_cap * G * 4);
    bb->next_embed = (float *)balloc(bb, (size_t)B * H * 4);
    bb->prefill_cap = pre_rows; bb->trailing_cap = tr_rows;
    bb->prefill = (float *)balloc(bb, (size_t)B * pre_rows * H * 4);
    bb->trailing = (float *)balloc(bb, (size_t)B * tr_rows * H * 4);
    bb->sub_rng = (uint32_t *)balloc(bb, (size_t)B * 4);
    bb->seeds = (uint32_t *)balloc(bb, (size_t)B * 4);
    const int vmax = tk->vocab > st->vocab ? tk->vocab : st->vocab;
    bb->scratch_f = (float *)balloc(bb, (size_t)B * vmax * 4);
    bb->scratch_i = (int *)balloc(bb, (size_t)B * vmax * 4);
    if (!bb->hidden || !bb->sub_in || !bb->pos || !bb->state || !bb->hist || !bb->frame_codes || !bb->codes || !bb->next_embed || !bb->prefill ||
        !bb->trailing || !bb->sub_rng || !bb->seeds || !bb->scratch_f || !bb->scratch_i) return -1;
    if (cudaMallocHost(&bb->state_host, (size_t)B * sizeof(GenState)) != cudaSuccess) { qtc_set_error("batch: pinned state"); return -1; }
    c->dev_mallocs += bb->mallocs;
    /* balloc's memsets and the weight re-layout copies ran on the legacy default stream, which c->stream (non-blocking) does
     * not order against: nothing on c->stream may touch these buffers before they have landed */
    QTC_CUDA(cudaDeviceSynchronize());
    return 0;
}

static bool b_attn_gqa_on() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("QTC_BATCH_ATTN_GQA"); on = e ? atoi(e) : 1; }
    return on != 0;
}
/* ---- one decoder-stack step for the whole batch ------------------------------------------------------------
 * s->x [B,H] holds the inputs on entry.  pos (device, per utterance, may be NULL) + pos_add = KV position.
 * hidden_out (optional) receives the post-final-norm hidden rows; head_w (optional) -> s->logits [B, head_rows]. */
static int b_lm_step(qtc_ctx *c, BatchBuf *bb, const LmDev *lm, BStack *s, const LmBuffers *rope, int B, const int *pos, int pos_add,
                     float *hidden_out, const uint16_t *head_w, int head_rows) {
    const int H = lm->hidden, I = lm->inter, hd = lm->head_dim, qd = lm->heads * hd, kvd = lm->kv_heads * hd, np = bb->pieces;
    const int ldq = qd + 2 * kvd;
    const size_t kv_stride = (size_t)lm->layers * s->kv_max * kvd;
    for (int l = 0; l < lm->layers; l++) {
        const LmLayerDev *w = &lm->layer[l];
        qtc_launch_pdl(c->stream, k_b_norm, dim3(B), dim3(256), 0, bb->xb, (float *)NULL, s->x, w->in_norm, H, lm->eps, np, s->qkv, ldq, pos);
        BLAUNCH(c);
        if (qtc_gemm_tc(c, s->qkv, ldq, bb->xb, s->wqkv[l], B, ldq, H, np)) return -1;
        BAttnArgs a;
        a.qkv = s->qkv; a.q_norm = w->q_norm; a.k_norm = w->k_norm; a.cos = rope->cos; a.sin = rope->sin;
        if (s->kv_max > 0) { a.kc = s->kc + (size_t)l * s->kv_max * kvd; a.vc = s->vc + (size_t)l * s->kv_max * kvd; a.kv_stride = kv_stride; a.tables = NULL; a.table_stride = 0; }
        else { a.kc = c->tk.k + (size_t)l * c->tk.kv_max * kvd; a.vc = c->tk.v + (size_t)l * c->tk.kv_max * kvd; a.kv_stride = 0; a.tables = c->kv.tables; a.table_stride = c->kv.max_seq_pages; }
        a.xb = bb->xb; a.pos = pos; a.pos_add = pos_add; a.heads = lm->heads; a.kv_heads = lm->kv_heads; a.hd = hd; a.pieces = np;
        a.eps = lm->eps; a.scale = 1.0f / sqrtf((float)hd);
        if (!pos && pos_add + 1 <= 32 && hd <= 128 && hd % 4 == 0) qtc_launch_pdl(c->stream, k_b_attn_small, dim3(dim3((lm->heads + 3) / 4, B)), dim3(128), 0, a);
        else if (hd == 128 && lm->heads == 2 * lm->kv_heads && b_attn_gqa_on()) qtc_launch_pdl(c->stream, k_b_attn_gqa2, dim3(dim3(lm->kv_heads, B)), dim3(BA2_WARPS * 32), 0, a);
        else qtc_launch_pdl(c->stream, k_b_attn, dim3(dim3(lm->heads, B)), dim3(128), (size_t)(2 * hd + (s->kv_max > 0 ? s->kv_max : c->kv.used[0] * QTC_KV_PAGE_POS) + 32) * 4, a);
        BLAUNCH(c);
        if (qtc_gemm_tc(c, s->x, H, bb->xb, w->wo, B, H, qd, np)) return -1;                      /* x += Wo . att */
        static int epi = -1;
        if (epi < 0) { const char *e = getenv("QTC_BATCH_SWIGLU_EPI"); epi = e ? atoi(e) : 0; }   /* measured: split-K + kernel 7.6 ms/frame, epilogue 9.7 (B = 64, 1.7B) */
        qtc_launch_pdl(c->stream, k_b_norm, dim3(B), dim3(256), 0, bb->xb, (float *)NULL, s->x, w->post_norm, H, lm->eps, np, epi ? (float *)NULL : s->gu, 2 * I, pos);
        BLAUNCH(c);
        if (epi) {
            if (qtc_gemm_tc_ex(c, NULL, 0, bb->xb, s->wgu[l], B, 2 * I, H, np, s->xh)) return -1; /* gate/up + SwiGLU in the epilogue */
        } else {
            if (qtc_gemm_tc(c, s->gu, 2 * I, bb->xb, s->wgu[l], B, 2 * I, H, np)) return -1;      /* split-K, then SwiGLU as a kernel */
            qtc_launch_pdl(c->stream, k_b_swiglu, dim3(dim3((I + 255) / 256, B)), dim3(256), 0, s->xh, s->gu, I, np);
            BLAUNCH(c);
        }
        if (qtc_gemm_tc(c, s->x, H, s->xh, w->down, B, H, I, np)) return -1;                      /* x += Wdown . h */
    }
    qtc_launch_pdl(c->stream, k_b_norm, dim3(B), dim3(256), 0, bb->xb, hidden_out, s->x, lm->final_norm, H, lm->eps, np, head_w ? s->logits : (float *)NULL, head_rows, pos);
    BLAUNCH(c);
    if (head_w && qtc_gemm_tc(c, s->logits, head_rows, bb->xb, head_w, B, head_rows, H, np)) return -1;
    return 0;
}

/* qwen_tts_talker_forward for every utterance: embeds [B,H] (device) -> hidden, logits; positions advance */
static int b_talker_forward(qtc_ctx *c, BatchBuf *bb, int B, const float *embeds_dev, int with_head) {
    const int H = c->H;
    qtc_launch_pdl(c->stream, k_b_copy_rows, dim3(dim3((H + 255) / 256, B)), dim3(256), 0, bb->tk.x, embeds_dev, H);
    BLAUNCH(c);
    if (b_lm_step(c, bb, &c->talker, &bb->tk, &c->tk, B, bb->pos, 0, bb->hidden, with_head ? c->codec_head : NULL, c->talker.vocab)) return -1;
    qtc_launch_pdl(c->stream, k_b_pos_add, dim3((B + 63) / 64), dim3(64), 0, bb->pos, B, 1);
    BLAUNCH(c);
    return 0;
}

static void b_sub_sample_args(qtc_ctx *c, BatchBuf *bb, SampleArgs *a, int g, const uint32_t *seed_tab) {
    memset(a, 0, sizeof(*a));
    const int G = c->G, v = c->sub.vocab;
    a->logits = bb->st.logits; a->bs_logits = v;
    a->cfg.vocab = v; a->cfg.top_k = c->params.subtalker_top_k; a->cfg.top_p = c->params.subtalker_top_p;
    a->cfg.temperature = c->params.subtalker_temperature; a->cfg.suppress_from = -1; a->cfg.eos_id = -1; a->cfg.rep_penalty = 1.0f;
    a->rng = bb->sub_rng; a->bs_rng = 1; a->out_token = bb->frame_codes + g; a->bs_out = G; a->role = 1;
    a->reseed = g == 1; a->seed_tab = seed_tab;
    a->scratch_f = bb->scratch_f; a->scratch_i = bb->scratch_i; a->bs_scratch = c->talker.vocab > v ? c->talker.vocab : v;
}

/* qwen_tts_subtalker_generate (c/qwen_tts_talker.c:539-736) for every utterance: bb->hidden, bb->frame_codes[:,0] -> frame_codes[:,1..] */
static int b_subtalker(qtc_ctx *c, BatchBuf *bb, int B, const uint32_t *seed_tab) {
    const LmDev *lm = &c->sub;
    const int H = c->H, h = lm->hidden, G = c->G, np = bb->pieces;
    if (set_smem_attrs()) return -1;
    for (int g = 0; g < G; g++) {
        const uint16_t *table = g == 0 ? NULL : (g == 1 ? c->codec_embedding : c->sub_emb[g - 2]);
        static int use_tab = -1;
        if (use_tab < 0) { const char *e = getenv("QTC_BATCH_PTAB"); use_tab = e ? atoi(e) : 1; }
        if (g >= 1 && use_tab && c->sub_ptab[g - 1] && h % 4 == 0) {
            qtc_launch_pdl(c->stream, k_b_sub_input_tab, dim3(dim3((h / 4 + 255) / 256, B)), dim3(256), 0, bb->st.x, c->sub_ptab[g - 1], (const int *)bb->frame_codes, G, g - 1, h);
            BLAUNCH(c);
        } else {
        qtc_launch_pdl(c->stream, k_b_sub_input, dim3(dim3((H + 255) / 256, B)), dim3(256), 0, bb->sub_in, bb->hidden, table, bb->frame_codes, G, g > 0 ? g - 1 : 0, H);
        BLAUNCH(c);
        if (c->sub_in_w) {
            qtc_launch_pdl(c->stream, k_b_pieces, dim3(dim3((H + 255) / 256, B)), dim3(256), 0, bb->xb, bb->sub_in, H, np, (float *)NULL, 0);
            BLAUNCH(c);
            qtc_launch_pdl(c->stream, k_b_init_rows, dim3(dim3((h + 255) / 256, B)), dim3(256), 0, bb->st.x, c->sub_in_b, h);
            BLAUNCH(c);
            if (qtc_gemm_tc(c, bb->st.x, h, bb->xb, c->sub_in_w, B, h, H, np)) return -1;
        } else {
            qtc_launch_pdl(c->stream, k_b_copy_pad_rows, dim3(dim3((h + 255) / 256, B)), dim3(256), 0, bb->st.x, bb->sub_in, H < h ? H : h, h);
            BLAUNCH(c);
        }
        }
        if (b_lm_step(c, bb, lm, &bb->st, &c->st, B, NULL, g, NULL, g >= 1 ? c->sub_head[g - 1] : NULL, lm->vocab)) return -1;
        if (g >= 1) {
            SampleArgs a;
            b_sub_sample_args(c, bb, &a, g, seed_tab);
            qtc_launch_pdl(c->stream, k_sample, dim3(B), dim3(SAMPLE_THREADS), (size_t)a.cfg.vocab * 8, a);
            BLAUNCH(c);
        }
    }
    return 0;
}

/* c/qwen_tts.c:1302-1363 for every utterance: logits policy + sample, sub-talker, next-input embedding, frame counter */
static int b_frame_front(qtc_ctx *c, BatchBuf *bb, int B, const uint32_t *seed_tab) {
    const LmDev *lm = &c->talker;
    const int H = c->H, G = c->G;
    SampleArgs a;
    memset(&a, 0, sizeof(a));
    a.logits = bb->tk.logits; a.bs_logits = lm->vocab;
    a.cfg.vocab = lm->vocab; a.cfg.top_k = c->params.top_k; a.cfg.top_p = c->params.top_p; a.cfg.temperature = c->params.temperature;
    a.cfg.suppress_from = lm->vocab - 1024; a.cfg.eos_id = c->host.codec_eos_id; a.cfg.rep_penalty = c->params.repetition_penalty;
    a.rng = &bb->state->rng; a.bs_rng = (int)(sizeof(GenState) / 4);
    a.hist_count = bb->hist; a.bs_hist = lm->vocab; a.st = bb->state; a.codes = bb->frame_codes; a.bs_codes = G; a.G = G; a.role = 0;
    a.scratch_f = bb->scratch_f; a.scratch_i = bb->scratch_i; a.bs_scratch = lm->vocab > c->sub.vocab ? lm->vocab : c->sub.vocab;
    if (set_smem_attrs()) return -1;
    qtc_launch_pdl(c->stream, k_sample, dim3(B), dim3(SAMPLE_THREADS), (size_t)lm->vocab * 8, a);
    BLAUNCH(c);
    if (b_subtalker(c, bb, B, seed_tab)) return -1;
    NextEmbedArgs n;
    memset(&n, 0, sizeof(n));
    n.out = bb->next_embed; n.codec_embedding = c->codec_embedding;
    for (int g = 0; g < G - 1; g++) n.sub_emb[g] = c->sub_emb[g];
    n.frame_codes = bb->frame_codes; n.bs_fc = G; n.all_codes = bb->codes; n.bs_codes = bb->codes_cap * G;
    n.trailing = bb->trailing; n.bs_tr = bb->trailing_cap * H; n.tts_pad = c->tts_pad; n.st = bb->state; n.H = H; n.G = G;
    qtc_launch_pdl(c->stream, k_next_embed, dim3(dim3((H + 255) / 256, B)), dim3(256), 0, n);
    BLAUNCH(c);
    qtc_launch_pdl(c->stream, k_advance_step, dim3(B), dim3(1), 0, bb->state);
    BLAUNCH(c);
    return 0;
}

/* ---- frame loop as two CUDA graphs ------------------------------------------------------------------------ */
static int b_params_key(const qtc_ctx *c, int B) {
    uint32_t h = 2166136261u;
    const unsigned char *p = (const unsigned char *)&c->params;
    for (size_t i = 0; i < sizeof(c->params); i++) h = (h ^ p[i]) * 16777619u;
    h = (h ^ (uint32_t)B) * 16777619u;
    h = (h ^ (uint32_t)c->kv.used[0]) * 16777619u;      /* k_b_attn's score buffer (dynamic smem) is sized by the rows' capacity */
    return (int)(h & 0x7fffffff);
}
static int b_capture(qtc_ctx *c, cudaGraphExec_t *out, int which, int B, long long *nlaunch) {
    BatchBuf *bb = c->bb;
    cudaGraph_t g = NULL;
    QTC_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const long long before = c->launches;
    const int rc = which == 0 ? b_frame_front(c, bb, B, bb->seeds) : b_talker_forward(c, bb, B, bb->next_embed, 1);
    *nlaunch = c->launches - before;
    c->launches = before;
    const cudaError_t e = cudaStreamEndCapture(c->stream, &g);
    if (rc != 0 || e != cudaSuccess) { if (g) cudaGraphDestroy(g); if (e != cudaSuccess) qtc_set_error("batch graph capture failed: %s", cudaGetErrorString(e)); return -1; }
    if (*out) { cudaGraphExecDestroy(*out); *out = NULL; }
    const cudaError_t ei = cudaGraphInstantiate(out, g, 0);
    cudaGraphDestroy(g);
    if (ei != cudaSuccess) { qtc_set_error("batch graph instantiate failed: %s", cudaGetErrorString(ei)); return -1; }
    return 0;
}

static int b_pull_state(qtc_ctx *c, BatchBuf *bb, int B) {
    QTC_CUDA(cudaMemcpyAsync(bb->state_host, bb->state, (size_t)B * sizeof(GenState), cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

/* codec decode of `count` finished utterances (codes on the device, `codes_stride` ints apart, nfr[u] frames each) into
 * pcm_out [count, max_tokens * 1920] (host): the decodes are enqueued back to back on the codec stream, each one's PCM parked
 * in a device-side stash, and the stash goes to the host once per 64 utterances - no host round trip between two decodes
 * (a decode of 250 frames is ~440 short launches: with a synchronise, a malloc and a pageable copy after every utterance the
 * device idled between them). */
static int b_decode_all(qtc_ctx *c, BatchBuf *bb, int count, const int *codes_base, size_t codes_stride, const int *nfr, float *pcm_out, int max_tokens) {
    const size_t per = (size_t)max_tokens * 1920;
    std::vector<int> ns(count, 0);
    for (int lo = 0; lo < count; lo += BROWS) {
        const int hi = lo + BROWS < count ? lo + BROWS : count;
        const size_t need = (size_t)(hi - lo) * per;
        if (bb->pcm_cap < need) {
            bb->pcm_all = (float *)balloc(bb, need * 4);
            if (!bb->pcm_all) { bb->pcm_cap = 0; return -1; }
            bb->pcm_cap = need; c->dev_mallocs++;
            QTC_CUDA(cudaDeviceSynchronize());           /* balloc clears on the NULL stream */
        }
        for (int u = lo; u < hi; u++) {
            if (nfr[u] <= 0) continue;
            float *wav = NULL;
            if (qtc_codec_stream_begin(c, nfr[u])) return -1;
            const int L = qtc_codec_stream_push(c, codes_base + (size_t)u * codes_stride, nfr[u], &wav);
            if (L <= 0 || (size_t)L > per) { if (L > 0) qtc_set_error("batch decode: %d samples for a slot of %zu", L, per); return -1; }
            QTC_CUDA(cudaMemcpyAsync(bb->pcm_all + (size_t)(u - lo) * per, wav, (size_t)L * 4, cudaMemcpyDeviceToDevice, c->cstream));
            ns[u] = L;
        }
        cudaEventRecord(c->ev3, c->cstream);             /* end of device work; what follows is the D2H of the PCM */
        bool full = true;
        for (int u = lo; u < hi; u++) full = full && (size_t)ns[u] == per;
        if (full) QTC_CUDA(cudaMemcpyAsync(pcm_out + (size_t)lo * per, bb->pcm_all, need * 4, cudaMemcpyDeviceToHost, c->cstream));
        else for (int u = lo; u < hi; u++) if (ns[u] > 0) QTC_CUDA(cudaMemcpyAsync(pcm_out + (size_t)u * per, bb->pcm_all + (size_t)(u - lo) * per, (size_t)ns[u] * 4, cudaMemcpyDeviceToHost, c->cstream));
        QTC_CUDA(cudaStreamSynchronize(c->cstream));
    }
    c->last_pcm_dev = NULL; c->last_pcm_n = 0;
    return 0;
}

/* ---- whole batch: prompts -> codes (-> PCM) ------------------------------------------------------------------
 * qwen_tts_generate (c/qwen_tts.c:1059-1443) for B utterances at once; utterance b draws from seed sample_seed + b. */
extern "C" int qwen_tts_cuda_generate_batch(qtc_ctx_t *c, int batch, const int *token_ids, const int *offsets, const int *speaker_ids,
                                            const int *language_ids, int *out_codes, int *out_n_frames, float *pcm_out, int max_tokens) {
    if (!c || c->talker.layers <= 0) { qtc_set_error("generate_batch: context has no talker weights"); return -1; }
    if (batch < 1 || batch > BROWS || !token_ids || !offsets || !out_codes || max_tokens < 1) { qtc_set_error("generate_batch: batch must be 1..64 and buffers non-null"); return -1; }
    if (c->talker.hidden % 64 || c->talker.inter % 64 || (c->talker.heads * c->talker.head_dim) % 64 || c->sub.hidden % 64 || c->sub.inter % 64 ||
        (c->sub.heads * c->sub.head_dim) % 64) { qtc_set_error("generate_batch: model dimensions must be multiples of 64"); return -1; }
    const int B = batch, H = c->H, G = c->G;
    const int fixed = c->params.fixed_codec_tokens > 0 ? c->params.fixed_codec_tokens : 0;
    int limit = fixed > 0 ? fixed : c->params.max_new_tokens;
    if (limit > max_tokens) limit = max_tokens;
    int max_tr = 1, max_plen = 10;
    for (int b = 0; b < B; b++) {
        const int n = offsets[b + 1] - offsets[b];
        if (n < 8) { qtc_set_error("generate_batch: utterance %d has %d token ids, need >= 8", b, n); return -1; }
        if (n - 8 > max_tr) max_tr = n - 8;
    }
    if (batch_ensure(c, B, max_plen + limit + 2, limit, max_plen + 3, max_tr + 1)) return -1;
    BatchBuf *bb = c->bb;
    cudaEventRecord(c->ev2, c->stream);
    /* prompts: built per utterance with the batch-1 prompt kernels straight into the batch buffers */
    std::vector<int> plen(B), ntr(B), pos0(B);
    std::vector<uint32_t> seeds(B);
    int maxlen = 0;
    for (int b = 0; b < B; b++) {
        const int spk = speaker_ids ? speaker_ids[b] : -1, lang = language_ids ? language_ids[b] : -1;
        if (qtc_build_prompt_dev(c, token_ids + offsets[b], offsets[b + 1] - offsets[b], spk, lang, bb->prefill + (size_t)b * bb->prefill_cap * H, &plen[b],
                                 bb->trailing + (size_t)b * bb->trailing_cap * H, &ntr[b], c->tts_pad)) return -1;
        if (plen[b] > maxlen) maxlen = plen[b];
    }
    for (int b = 0; b < B; b++) {
        GenState s;
        memset(&s, 0, sizeof(s));
        const float f = (float)(c->params.sample_seed + b);
        memcpy(&s.rng, &f, 4); memcpy(&seeds[b], &f, 4);
        s.limit = limit; s.fixed = fixed > 0; s.n_trailing = ntr[b]; s.stop_step = limit;
        bb->state_host[b] = s;
        pos0[b] = plen[b] - maxlen;               /* right-aligned prompts: negative until the utterance starts */
    }
    QTC_CUDA(cudaMemcpyAsync(bb->state, bb->state_host, (size_t)B * sizeof(GenState), cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaMemcpyAsync(bb->pos, pos0.data(), (size_t)B * 4, cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaMemcpyAsync(bb->seeds, seeds.data(), (size_t)B * 4, cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaMemsetAsync(bb->hist, 0, (size_t)B * c->talker.vocab * 4, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));      /* pos0 / seeds are stack vectors */
    if (pf_ok(c) && prefill_gemm_on()) {
        /* prefill of the whole batch as ONE ragged pass (qtc_prefill.cuh): R = sum of the prompt lengths rows through every
         * projection as one tensor-core GEMM; a row knows its sequence (KV pool slot b) and position.  Every utterance then
         * stands at its own position plen[b] - the lock-step frames that follow never needed the prompts aligned. */
        int R = 0;
        for (int b = 0; b < B; b++) R += plen[b];
        if (pf_reserve(c, R)) return -1;
        std::vector<int> hs(R), hp(R), last(B);
        for (int b = 0, r = 0; b < B; b++) { for (int t = 0; t < plen[b]; t++, r++) { hs[r] = b; hp[r] = t; } last[b] = r - 1; }
        QTC_CUDA(cudaMemcpyAsync(c->pf->slot, hs.data(), (size_t)R * 4, cudaMemcpyHostToDevice, c->stream));
        QTC_CUDA(cudaMemcpyAsync(c->pf->pos, hp.data(), (size_t)R * 4, cudaMemcpyHostToDevice, c->stream));
        QTC_CUDA(cudaMemcpyAsync(bb->frame_codes, last.data(), (size_t)B * 4, cudaMemcpyHostToDevice, c->stream));     /* scratch: last row of every utterance */
        QTC_CUDA(cudaStreamSynchronize(c->stream));
        k_pf_gather_rows<<<dim3((H + 255) / 256, R), 256, 0, c->stream>>>(c->pf->x, bb->prefill, bb->prefill_cap, c->pf->slot, c->pf->pos, H);
        BLAUNCH(c);
        if (pf_run(c, R, maxlen - 1)) return -1;
        k_pf_take_last<<<dim3((H + 255) / 256, B), 256, 0, c->stream>>>(bb->tk.x, c->pf->x, bb->frame_codes, H);
        BLAUNCH(c);
        for (int b = 0; b < B; b++) pos0[b] = plen[b] - 1;
        QTC_CUDA(cudaMemcpyAsync(bb->pos, pos0.data(), (size_t)B * 4, cudaMemcpyHostToDevice, c->stream));
        QTC_CUDA(cudaStreamSynchronize(c->stream));
        /* final norm -> hidden of every utterance (ctx->tk_x), codec head -> logits of frame 0 */
        qtc_launch_pdl(c->stream, k_b_norm, dim3(B), dim3(256), 0, bb->xb, bb->hidden, bb->tk.x, c->talker.final_norm, H, c->talker.eps, bb->pieces, bb->tk.logits, c->talker.vocab, bb->pos);
        BLAUNCH(c);
        if (qtc_gemm_tc(c, bb->tk.logits, c->talker.vocab, bb->xb, c->codec_head, B, c->talker.vocab, H, bb->pieces)) return -1;
        qtc_launch_pdl(c->stream, k_b_pos_add, dim3((B + 63) / 64), dim3(64), 0, bb->pos, B, 1);
        BLAUNCH(c);
        QTC_CUDA(cudaMemsetAsync(bb->frame_codes, 0, (size_t)B * G * 4, c->stream));
    } else
    /* lock-step prefill: per token the reference's batched prefill does a decode step's arithmetic */
    for (int s = 0; s < maxlen; s++) {
        qtc_launch_pdl(c->stream, k_b_fill_prefill, dim3(dim3((H + 255) / 256, B)), dim3(256), 0, bb->tk.x, bb->prefill, bb->prefill_cap, bb->pos, H);
        BLAUNCH(c);
        const int last = s + 1 == maxlen;
        if (b_lm_step(c, bb, &c->talker, &bb->tk, &c->tk, B, bb->pos, 0, bb->hidden, last ? c->codec_head : NULL, c->talker.vocab)) return -1;
        qtc_launch_pdl(c->stream, k_b_pos_add, dim3((B + 63) / 64), dim3(64), 0, bb->pos, B, 1);
        BLAUNCH(c);
    }
    const int key = b_params_key(c, B);
    if (bb->graph_key != key || bb->graph_B != B || !bb->g_front || !bb->g_talker) {
        if (b_capture(c, &bb->g_front, 0, B, &bb->n_front) || b_capture(c, &bb->g_talker, 1, B, &bb->n_talker)) return -1;
        bb->graph_key = key; bb->graph_B = B;
    }
    QTC_CUDA(cudaEventRecord(c->ev0, c->stream));
    int launched = 0;
    const int poll = fixed > 0 ? limit : 8;
    while (launched < limit) {
        const int upto = launched + poll < limit ? launched + poll : limit;
        for (; launched < upto; launched++) {
            QTC_CUDA(cudaGraphLaunch(bb->g_front, c->stream));
            c->launches += bb->n_front;
            if (launched + 1 < limit) { QTC_CUDA(cudaGraphLaunch(bb->g_talker, c->stream)); c->launches += bb->n_talker; }
        }
        if (b_pull_state(c, bb, B)) return -1;
        bool all_done = true;
        for (int b = 0; b < B; b++) all_done = all_done && bb->state_host[b].done;
        if (all_done) break;
    }
    QTC_CUDA(cudaEventRecord(c->ev1, c->stream));
    if (b_pull_state(c, bb, B)) return -1;
    std::vector<int> host_codes((size_t)bb->codes_cap * G);
    int total_frames = 0;
    for (int b = 0; b < B; b++) {
        const int n = bb->state_host[b].step;
        if (out_n_frames) out_n_frames[b] = n;
        total_frames += n;
        if (n > 0) {
            QTC_CUDA(cudaMemcpy(host_codes.data(), bb->codes + (size_t)b * bb->codes_cap * G, (size_t)n * G * 4, cudaMemcpyDeviceToHost));
            memcpy(out_codes + (size_t)b * max_tokens * G, host_codes.data(), (size_t)n * G * 4);
        }
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->perf.talker_ms = ms; c->perf.codec_tokens = total_frames;
    if (pcm_out) {
        if (!c->codec) { qtc_set_error("generate_batch: context has no codec weights"); return -1; }
        std::vector<int> nfr(B);
        for (int b = 0; b < B; b++) nfr[b] = bb->state_host[b].step;
        if (b_decode_all(c, bb, B, bb->codes, (size_t)bb->codes_cap * G, nfr.data(), pcm_out, max_tokens)) return -1;
        float dms = 0;
        cudaEventElapsedTime(&dms, c->ev2, c->ev3);
        c->perf.device_ms = dms;
    } else {
        float dms = 0;
        cudaEventElapsedTime(&dms, c->ev2, c->ev1);
        c->perf.device_ms = dms;
    }
    return 0;
}

/* ---- continuous batching: a queue of utterances through `rows` lock-step rows -------------------------------------------
 * The reference is batch 1 (c/qwen_tts.h:8) and stops an utterance at EOS (c/qwen_tts.c:1324-1330); in a lock-step batch the
 * utterances stop at different frames and qwen_tts_cuda_generate_batch keeps the finished rows as no-ops until the slowest ends.
 * Here a row whose utterance has ended is RETIRED at the next poll (every 8 frames; with a fixed frame budget exactly when the
 * first row's budget ends): its codes are saved, its KV pages go back to the pool's free list, and the next utterance of the
 * queue is ADMITTED into the row - pages reserved, prompt built, prefilled on its own (qtc_prefill.cuh, KV rows straight into
 * the row's pages), frame-0 logits from the codec head, per-row state / RNG / repetition history reset - while the other rows
 * are mid-utterance.  Rows never see each other, so utterance u (seed sample_seed + u) produces what it produces alone.
 * frame_budgets (optional, [n_utt]): utterance u stops after min(budget, the call's limit) frames - requests of different
 * lengths are where a freed row pays: lock-step batches run every row to the longest request. */
__global__ void k_pf_fill_rows_imm(int *slot, int *pos, int R, int s, int base) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < R) { slot[r] = s; pos[r] = base + r; }
}
#define QTC_Q_POLL 8
#define QTC_Q_PARKED (-(1 << 30))         /* position of a row with no utterance: negative = every kernel skips it, and 2^30 steps stay negative */
extern "C" int qwen_tts_cuda_generate_queue(qtc_ctx_t *c, int n_utt, int rows, const int *token_ids, const int *offsets, const int *speaker_ids,
                                            const int *language_ids, const int *frame_budgets, int *out_codes, int *out_n_frames, float *pcm_out, int max_tokens) {
    if (!c || c->talker.layers <= 0) { qtc_set_error("generate_queue: context has no talker weights"); return -1; }
    if (n_utt < 1 || rows < 1 || rows > BROWS || !token_ids || !offsets || !out_codes || max_tokens < 1) { qtc_set_error("generate_queue: rows must be 1..64 and buffers non-null"); return -1; }
    if (c->talker.hidden % 64 || c->talker.inter % 64 || (c->talker.heads * c->talker.head_dim) % 64 || c->sub.hidden % 64 || c->sub.inter % 64 ||
        (c->sub.heads * c->sub.head_dim) % 64) { qtc_set_error("generate_queue: model dimensions must be multiples of 64"); return -1; }
    if (!(pf_ok(c) && prefill_gemm_on())) { qtc_set_error("generate_queue: needs fast arithmetic and the GEMM prefill (a row is prefilled on its own while the others are mid-utterance)"); return -1; }
    const int B = rows < n_utt ? rows : n_utt, H = c->H, G = c->G, V = c->talker.vocab;
    if (B > c->kv.slots) { qtc_set_error("generate_queue: %d rows, the context was uploaded with max_batch %d", B, c->kv.slots); return -1; }
    const int fixed = c->params.fixed_codec_tokens > 0 ? c->params.fixed_codec_tokens : 0;
    int limit = fixed > 0 ? fixed : c->params.max_new_tokens;
    if (limit > max_tokens) limit = max_tokens;
    int max_tr = 1;
    const int max_plen = 10;
    for (int u = 0; u < n_utt; u++) {
        const int n = offsets[u + 1] - offsets[u];
        if (n < 8) { qtc_set_error("generate_queue: utterance %d has %d token ids, need >= 8", u, n); return -1; }
        if (n - 8 > max_tr) max_tr = n - 8;
    }
    /* a row that ends between two polls keeps stepping until the poll: its pages cover those positions too */
    const int kv_need = max_plen + limit + QTC_Q_POLL + 4;
    if (batch_ensure(c, B, kv_need, limit, max_plen + 3, max_tr + 1)) return -1;
    BatchBuf *bb = c->bb;
    if (pcm_out && !c->codec) { qtc_set_error("generate_queue: context has no codec weights"); return -1; }
    if (pcm_out && bb->qcodes_cap < (size_t)n_utt * limit * G) {          /* finished utterances' codes wait on the device for the decode at the end */
        bb->qcodes_cap = (size_t)n_utt * limit * G;
        bb->qcodes = (int *)balloc(bb, bb->qcodes_cap * 4);
        if (!bb->qcodes) return -1;
        c->dev_mallocs++;
        QTC_CUDA(cudaDeviceSynchronize());               /* balloc clears on the NULL stream; the copies into the stash run on c->stream */
    }
    /* QTC_QUEUE_OVERLAP (default on): a retired request is decoded right away on the second stream while the rows go on generating -
     * the frame loop's kernels are short and latency-bound and leave most of the SMs to the vocoder; off: all decodes after the last frame */
    static int overlap_env = -1;
    if (overlap_env < 0) { const char *e = getenv("QTC_QUEUE_OVERLAP"); overlap_env = e ? atoi(e) : 1; }
    const bool overlap = pcm_out && overlap_env != 0 && c->stream2 != NULL;
    const size_t per = (size_t)max_tokens * 1920;
    struct EvGuard { cudaEvent_t e = NULL; ~EvGuard() { if (e) cudaEventDestroy(e); } } evg;      /* destroyed on every way out */
    cudaEvent_t evq = NULL;
    std::vector<int> qns(n_utt, 0);
    if (overlap) {
        if (bb->qpcm_cap < (size_t)n_utt * per) {
            bb->qpcm_cap = (size_t)n_utt * per;
            bb->qpcm = (float *)balloc(bb, bb->qpcm_cap * 4);
            if (!bb->qpcm) { bb->qpcm_cap = 0; return -1; }
            c->dev_mallocs++;
            QTC_CUDA(cudaDeviceSynchronize());
        }
        QTC_CUDA(cudaEventCreateWithFlags(&evg.e, cudaEventDisableTiming));
        evq = evg.e;
    }
    cudaEventRecord(c->ev2, c->stream);
    std::vector<int> row_utt(B, -1), row_end(B, 0), row_lim(B, 0), pos_host(B, QTC_Q_PARKED), nframes(n_utt, 0);
    std::vector<uint32_t> seed_host(B, 0);
    int frame_now = 0, next_u = 0, admitted = 0;
    /* the next utterances of the queue -> the rows in `free_rows` (as many as there are of both); their prompts are prefilled
     * together as ONE ragged pass.  Everything is enqueued on c->stream behind the frames already launched. */
    std::vector<int> hs, hp, plen(B, 0), ntr(B, 0);
    auto admit = [&](const std::vector<int> &free_rows) -> int {
        std::vector<int> rows_in;
        for (int b : free_rows) {
            if (next_u >= n_utt) break;
            const int u = next_u++;
            KvPool &kp = c->kv;
            for (int i = kp.used[b] - 1; i >= 0; i--) kp.free_pages.push_back(kp.tables_host[(size_t)b * kp.max_seq_pages + i]);
            kp.used[b] = 0;
            if (qtc_kv_reserve(c, b, kv_need)) return -1;
            const int spk = speaker_ids ? speaker_ids[u] : -1, lang = language_ids ? language_ids[u] : -1;
            if (qtc_build_prompt_dev(c, token_ids + offsets[u], offsets[u + 1] - offsets[u], spk, lang, bb->prefill + (size_t)b * bb->prefill_cap * H, &plen[b],
                                     bb->trailing + (size_t)b * bb->trailing_cap * H, &ntr[b], c->tts_pad)) return -1;
            row_utt[b] = u; row_lim[b] = frame_budgets && frame_budgets[u] < limit ? (frame_budgets[u] > 1 ? frame_budgets[u] : 1) : limit;
            row_end[b] = frame_now + row_lim[b]; admitted++;
            rows_in.push_back(b);
        }
        if (rows_in.empty()) return 0;
        hs.clear(); hp.clear();
        int maxlen = 0;
        for (int b : rows_in) { for (int t = 0; t < plen[b]; t++) { hs.push_back(b); hp.push_back(t); } if (plen[b] > maxlen) maxlen = plen[b]; }
        const int R = (int)hs.size();
        if (pf_reserve(c, R)) return -1;
        QTC_CUDA(cudaMemcpyAsync(c->pf->slot, hs.data(), (size_t)R * 4, cudaMemcpyHostToDevice, c->stream));
        QTC_CUDA(cudaMemcpyAsync(c->pf->pos, hp.data(), (size_t)R * 4, cudaMemcpyHostToDevice, c->stream));
        k_pf_gather_rows<<<dim3((H + 255) / 256, R), 256, 0, c->stream>>>(c->pf->x, bb->prefill, bb->prefill_cap, c->pf->slot, c->pf->pos, H);
        BLAUNCH(c);
        if (pf_run(c, R, maxlen - 1)) return -1;
        int r_last = -1;
        for (int b : rows_in) {
            r_last += plen[b];
            /* final norm -> the row's hidden (ctx->tk_x), codec head -> its logits of frame 0; the other rows' are left alone */
            if (qtc_rmsnorm(c, bb->hidden + (size_t)b * H, c->pf->x + (size_t)r_last * H, c->talker.final_norm, H, c->talker.eps)) return -1;
            GemvSeg hseg = {c->codec_head, bb->tk.logits + (size_t)b * V, V};
            if (qtc_gemv(c, &hseg, 1, bb->hidden + (size_t)b * H, H, EPI_STORE, NULL)) return -1;
            GenState s;
            memset(&s, 0, sizeof(s));
            const float f = (float)(c->params.sample_seed + row_utt[b]);
            memcpy(&s.rng, &f, 4); memcpy(&seed_host[b], &f, 4);
            s.limit = row_lim[b]; s.fixed = fixed > 0; s.n_trailing = ntr[b]; s.stop_step = row_lim[b];
            bb->state_host[b] = s;
            pos_host[b] = plen[b];
            QTC_CUDA(cudaMemcpyAsync(bb->state + b, bb->state_host + b, sizeof(GenState), cudaMemcpyHostToDevice, c->stream));
            QTC_CUDA(cudaMemcpyAsync(bb->pos + b, &pos_host[b], 4, cudaMemcpyHostToDevice, c->stream));
            QTC_CUDA(cudaMemcpyAsync(bb->seeds + b, &seed_host[b], 4, cudaMemcpyHostToDevice, c->stream));
            QTC_CUDA(cudaMemsetAsync(bb->hist + (size_t)b * V, 0, (size_t)V * 4, c->stream));
            QTC_CUDA(cudaMemsetAsync(bb->frame_codes + (size_t)b * G, 0, (size_t)G * 4, c->stream));
        }
        QTC_CUDA(cudaStreamSynchronize(c->stream));      /* the host-side staging (state_host, hs, hp, ...) is rewritten by the next poll */
        return 0;
    };
    {
        std::vector<int> all(B);
        for (int b = 0; b < B; b++) all[b] = b;
        if (admit(all)) return -1;
    }
    const int key = b_params_key(c, B);
    if (bb->graph_key != key || bb->graph_B != B || !bb->g_front || !bb->g_talker) {
        if (b_capture(c, &bb->g_front, 0, B, &bb->n_front) || b_capture(c, &bb->g_talker, 1, B, &bb->n_talker)) return -1;
        bb->graph_key = key; bb->graph_B = B;
    }
    QTC_CUDA(cudaEventRecord(c->ev0, c->stream));
    std::vector<int> host_codes((size_t)bb->codes_cap * G);
    int active = B, total_frames = 0;
    long long guard = ((long long)n_utt / B + 2) * ((long long)limit + 2 * QTC_Q_POLL);      /* frames this call can need at most */
    while (active > 0) {
        int poll = QTC_Q_POLL;
        if (fixed > 0) {                                  /* the host knows when the next row's budget ends */
            poll = limit;
            for (int b = 0; b < B; b++) if (row_utt[b] >= 0 && row_end[b] - frame_now < poll) poll = row_end[b] - frame_now;
            if (poll < 1) poll = 1;
        }
        for (int i = 0; i < poll; i++, frame_now++) {
            QTC_CUDA(cudaGraphLaunch(bb->g_front, c->stream));
            QTC_CUDA(cudaGraphLaunch(bb->g_talker, c->stream));
            c->launches += bb->n_front + bb->n_talker;
        }
        if ((guard -= poll) < 0) { qtc_set_error("generate_queue: the frame loop ran past its bound"); return -1; }
        if (b_pull_state(c, bb, B)) return -1;
        std::vector<int> freed;
        for (int b = 0; b < B; b++) {
            const int u = row_utt[b];
            if (u < 0 || !bb->state_host[b].done) continue;
            /* retire: codes to the caller (and to the device-side stash for the decode at the end) */
            const int n = bb->state_host[b].step;
            nframes[u] = n; total_frames += n;
            if (out_n_frames) out_n_frames[u] = n;
            if (n > 0) {
                QTC_CUDA(cudaMemcpy(host_codes.data(), bb->codes + (size_t)b * bb->codes_cap * G, (size_t)n * G * 4, cudaMemcpyDeviceToHost));
                memcpy(out_codes + (size_t)u * max_tokens * G, host_codes.data(), (size_t)n * G * 4);
                if (pcm_out) QTC_CUDA(cudaMemcpyAsync(bb->qcodes + (size_t)u * limit * G, bb->codes + (size_t)b * bb->codes_cap * G, (size_t)n * G * 4, cudaMemcpyDeviceToDevice, c->stream));
                if (overlap) {
                    /* the codes are in the stash once the frame loop's stream reaches this point; the decode waits for exactly that */
                    QTC_CUDA(cudaEventRecord(evq, c->stream));
                    QTC_CUDA(cudaStreamWaitEvent(c->stream2, evq, 0));
                    cudaStream_t keep = c->cstream;
                    c->cstream = c->stream2;
                    float *wav = NULL;
                    int L = -1;
                    if (qtc_codec_stream_begin(c, n) == 0) L = qtc_codec_stream_push(c, bb->qcodes + (size_t)u * limit * G, n, &wav);
                    if (L > 0 && (size_t)L <= per) {
                        if (cudaMemcpyAsync(bb->qpcm + (size_t)u * per, wav, (size_t)L * 4, cudaMemcpyDeviceToDevice, c->stream2) != cudaSuccess) L = -1;
                    } else L = -1;
                    c->cstream = keep;
                    if (L <= 0) { qtc_set_error("generate_queue: decode of request %d failed", u); return -1; }
                    qns[u] = L;
                }
            }
            row_utt[b] = -1;
            freed.push_back(b);
        }
        if (admit(freed)) return -1;
        bool parked = false;
        for (int b : freed) {
            if (row_utt[b] >= 0) continue;               /* took a new utterance */
            active--;
            pos_host[b] = QTC_Q_PARKED;
            QTC_CUDA(cudaMemcpyAsync(bb->pos + b, &pos_host[b], 4, cudaMemcpyHostToDevice, c->stream));
            parked = true;
        }
        if (parked) QTC_CUDA(cudaStreamSynchronize(c->stream));
    }
    QTC_CUDA(cudaEventRecord(c->ev1, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->perf.talker_ms = ms; c->perf.codec_tokens = total_frames;
    if (overlap) {
        /* every decode is already enqueued on the second stream: one more hop to the host */
        cudaEventRecord(c->ev3, c->stream2);
        bool full = true;
        for (int u = 0; u < n_utt; u++) full = full && (size_t)qns[u] == per;
        if (full) QTC_CUDA(cudaMemcpyAsync(pcm_out, bb->qpcm, (size_t)n_utt * per * 4, cudaMemcpyDeviceToHost, c->stream2));
        else for (int u = 0; u < n_utt; u++) if (qns[u] > 0) QTC_CUDA(cudaMemcpyAsync(pcm_out + (size_t)u * per, bb->qpcm + (size_t)u * per, (size_t)qns[u] * 4, cudaMemcpyDeviceToHost, c->stream2));
        QTC_CUDA(cudaStreamSynchronize(c->stream2));
        c->last_pcm_dev = NULL; c->last_pcm_n = 0;
        float dms = 0;
        cudaEventElapsedTime(&dms, c->ev2, c->ev3);
        c->perf.device_ms = dms;
    } else if (pcm_out) {
        if (b_decode_all(c, bb, n_utt, bb->qcodes, (size_t)limit * G, nframes.data(), pcm_out, max_tokens)) return -1;
        float dms = 0;
        cudaEventElapsedTime(&dms, c->ev2, c->ev3);
        c->perf.device_ms = dms;
    } else {
        float dms = 0;
        cudaEventElapsedTime(&dms, c->ev2, c->ev1);
        c->perf.device_ms = dms;
    }
    c->perf.queue_admitted = admitted;
    return 0;
}

/* ---- stage-level test surface of the batched path (host buffers) -------------------------------------------------- */
/* all B utterances at KV length 0 */
extern "C" int qwen_tts_cuda_batch_reset(qtc_ctx_t *c, int batch, int kv_max) {
    if (!c || c->talker.layers <= 0 || batch < 1 || batch > BROWS) { qtc_set_error("batch_reset: bad arguments"); return -1; }
    if (batch_ensure(c, batch, kv_max, 8, 16, 2)) return -1;
    QTC_CUDA(cudaMemsetAsync(c->bb->pos, 0, (size_t)batch * 4, c->stream));
    return 0;
}
/* one talker decode step for every utterance: embeds [B,H] -> logits [B,vocab] (optional), hidden [B,H] (optional) */
extern "C" int qwen_tts_cuda_batch_forward(qtc_ctx_t *c, int batch, const float *embeds, float *logits, float *hidden) {
    if (!c || !c->bb || batch < 1 || batch > c->bb->B || !embeds) { qtc_set_error("batch_forward: call batch_reset first"); return -1; }
    BatchBuf *bb = c->bb;
    const int H = c->H, V = c->talker.vocab;
    QTC_CUDA(cudaMemcpyAsync(bb->next_embed, embeds, (size_t)batch * H * 4, cudaMemcpyHostToDevice, c->stream));
    if (b_talker_forward(c, bb, batch, bb->next_embed, 1)) return -1;
    if (logits) QTC_CUDA(cudaMemcpyAsync(logits, bb->tk.logits, (size_t)batch * V * 4, cudaMemcpyDeviceToHost, c->stream));
    if (hidden) QTC_CUDA(cudaMemcpyAsync(hidden, bb->hidden, (size_t)batch * H * 4, cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
/* the 15-codebook loop for every utterance: hidden [B,H], first_codes [B] -> out_codes [B,G]; utterance b seeds with sample_seed + b */
extern "C" int qwen_tts_cuda_batch_subtalker(qtc_ctx_t *c, int batch, const float *hidden, const int *first_codes, int *out_codes) {
    if (!c || !c->bb || batch < 1 || batch > c->bb->B || !hidden || !first_codes || !out_codes) { qtc_set_error("batch_subtalker: call batch_reset first"); return -1; }
    BatchBuf *bb = c->bb;
    const int H = c->H, G = c->G;
    std::vector<int> fc((size_t)batch * G, 0);
    std::vector<uint32_t> seeds(batch);
    for (int b = 0; b < batch; b++) { fc[(size_t)b * G] = first_codes[b]; const float f = (float)(c->params.sample_seed + b); memcpy(&seeds[b], &f, 4); }
    QTC_CUDA(cudaMemcpyAsync(bb->hidden, hidden, (size_t)batch * H * 4, cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaMemcpyAsync(bb->frame_codes, fc.data(), fc.size() * 4, cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaMemcpyAsync(bb->seeds, seeds.data(), (size_t)batch * 4, cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    if (b_subtalker(c, bb, batch, bb->seeds)) return -1;
    QTC_CUDA(cudaMemcpyAsync(out_codes, bb->frame_codes, (size_t)batch * G * 4, cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
/* 3: activations split exactly into three bf16 pieces (default); 1: rounded to bf16 */
extern "C" int qwen_tts_cuda_set_batch_pieces(qtc_ctx_t *c, int pieces) {
    if (!c || (pieces != 1 && pieces != 3)) return -1;
    c->batch_pieces = pieces;
    return 0;
}
