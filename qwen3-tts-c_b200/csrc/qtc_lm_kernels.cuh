/*
 * qtc_lm_kernels.cuh - the non-GEMV kernels of one decode step (batch 1):
 * RMSNorm, per-head QK-norm + RoPE + KV append + GQA decode attention,
 * embedding gathers and the next-input embedding sum.
 *
 * Every kernel reads the KV position from device memory (GenState.pos or an
 * int counter) so the same captured CUDA graph can be replayed every frame.
 */
#pragma once
#include "qtc_device.cuh"
#include "qtc_internal.h"

/* c/qwen_tts_kernels.c:27-39.  One CTA per vector (blockIdx.x), vectors `n` apart;
 * w is shared by all vectors (per-head norms) */
template <int ARITH>
__global__ void __launch_bounds__(256) k_rmsnorm(float *out, const float *x, const float *w, int n, float eps) {
    __shared__ float red[32];
    __shared__ float s_inv;
    const float *xi = x + (size_t)blockIdx.x * n;
    float *yo = out + (size_t)blockIdx.x * n;
    float inv;
    if (ARITH == QTC_STRICT) {
        if (threadIdx.x == 0) {
            float ss = 0.0f;
            for (int i = 0; i < n; i++) ss = __fadd_rn(ss, __fmul_rn(xi[i], xi[i]));
            s_inv = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)n), eps)));
        }
        __syncthreads();
        inv = s_inv;
        for (int i = threadIdx.x; i < n; i += blockDim.x) yo[i] = __fmul_rn(__fmul_rn(xi[i], inv), w[i]);
    } else {
        float ss = 0.0f;
        for (int i = threadIdx.x; i < n; i += blockDim.x) ss = fmaf(xi[i], xi[i], ss);
        ss = block_sum(ss, red);
        inv = 1.0f / sqrtf(ss / (float)n + eps);
        for (int i = threadIdx.x; i < n; i += blockDim.x) yo[i] = xi[i] * inv * w[i];
    }
}

struct AttnArgs {
    const float *q, *k, *v;        /* raw projections of this token: [heads*hd], [kvh*hd], [kvh*hd] */
    const float *q_norm, *k_norm;  /* [hd] */
    const float *cos, *sin;        /* [kv_max, hd] */
    float *kc, *vc;                /* this layer's cache: pool slice (paged) or [kv_max, kvh*hd] */
    const int *table;              /* page table of the utterance, or NULL (contiguous) */
    float *att;                    /* out [heads*hd] */
    const int *pos_ptr; int pos_add;
    int heads, kv_heads, hd;
    float eps, scale;
};

/* talker_attention_single steps 3-6, c/qwen_tts_talker.c:150-225 (and ST_FORWARD :647-672).
 * One CTA per Q head, 128 threads.  dynamic smem: 2*hd + (pos+1) floats (+32) */
template <int ARITH>
__global__ void __launch_bounds__(128) k_attn_decode(const AttnArgs a) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[32];
    __shared__ float s_b[2];
    const int hd = a.hd, half = hd >> 1;
    float *qs = sm, *ks = sm + hd, *sc = sm + 2 * hd;
    const int h = blockIdx.x, rep = a.heads / a.kv_heads, g = h / rep, kvd = a.kv_heads * hd;
    const int pos = *a.pos_ptr + a.pos_add;
    const float *qr = a.q + h * hd, *kr = a.k + g * hd;
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;

    /* per-head RMSNorm of q and k, then rotate-half */
    float qinv, kinv;
    if (ARITH == QTC_STRICT) {
        if (threadIdx.x < 2) {
            const float *p = threadIdx.x == 0 ? qr : kr;
            float ss = 0.0f;
            for (int i = 0; i < hd; i++) ss = __fadd_rn(ss, __fmul_rn(p[i], p[i]));
            s_b[threadIdx.x] = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)hd), a.eps)));
        }
        __syncthreads();
        qinv = s_b[0]; kinv = s_b[1];
        for (int i = threadIdx.x; i < hd; i += blockDim.x) {
            qs[i] = __fmul_rn(__fmul_rn(qr[i], qinv), a.q_norm[i]);
            ks[i] = __fmul_rn(__fmul_rn(kr[i], kinv), a.k_norm[i]);
        }
    } else {
        float s1 = 0.0f, s2 = 0.0f;
        for (int i = threadIdx.x; i < hd; i += blockDim.x) { s1 = fmaf(qr[i], qr[i], s1); s2 = fmaf(kr[i], kr[i], s2); }
        s1 = block_sum(s1, red);
        s2 = block_sum(s2, red);
        qinv = 1.0f / sqrtf(s1 / (float)hd + a.eps);
        kinv = 1.0f / sqrtf(s2 / (float)hd + a.eps);
        for (int i = threadIdx.x; i < hd; i += blockDim.x) { qs[i] = qr[i] * qinv * a.q_norm[i]; ks[i] = kr[i] * kinv * a.k_norm[i]; }
    }
    __syncthreads();
    float rq0 = 0, rq1 = 0, rk0 = 0, rk1 = 0;
    const int i0 = threadIdx.x;
    if (i0 < half) {
        const float qa = qs[i0], qb = qs[i0 + half], ka = ks[i0], kb = ks[i0 + half];
        if (ARITH == QTC_STRICT) {
            rq0 = __fsub_rn(__fmul_rn(qa, cs[i0]), __fmul_rn(qb, sn[i0]));
            rq1 = __fadd_rn(__fmul_rn(qb, cs[i0 + half]), __fmul_rn(qa, sn[i0 + half]));
            rk0 = __fsub_rn(__fmul_rn(ka, cs[i0]), __fmul_rn(kb, sn[i0]));
            rk1 = __fadd_rn(__fmul_rn(kb, cs[i0 + half]), __fmul_rn(ka, sn[i0 + half]));
        } else {
            rq0 = qa * cs[i0] - qb * sn[i0]; rq1 = qb * cs[i0 + half] + qa * sn[i0 + half];
            rk0 = ka * cs[i0] - kb * sn[i0]; rk1 = kb * cs[i0 + half] + ka * sn[i0 + half];
        }
    }
    __syncthreads();
    if (i0 < half) { qs[i0] = rq0; qs[i0 + half] = rq1; ks[i0] = rk0; ks[i0 + half] = rk1; }
    __syncthreads();
    const float *vr = a.v + g * hd;
    if (h % rep == 0)                                         /* KV append, c/qwen_tts_talker.c:192-196 */
        for (int i = threadIdx.x; i < hd; i += blockDim.x) {
            a.kc[kv_off(a.table, pos, kvd) + g * hd + i] = ks[i];
            a.vc[kv_off(a.table, pos, kvd) + g * hd + i] = vr[i];
        }

    /* scores */
    const int n = pos + 1;
    if (ARITH == QTC_STRICT) {
        for (int t = threadIdx.x; t < n; t += blockDim.x) {
            float d = 0.0f;
            if (t < pos) {
                const float4 *kt = reinterpret_cast<const float4 *>(a.kc + kv_off(a.table, t, kvd) + g * hd);
                for (int i = 0; i < hd / 4; i++) {
                    const float4 kv = kt[i];
                    d = __fadd_rn(d, __fmul_rn(qs[4 * i], kv.x)); d = __fadd_rn(d, __fmul_rn(qs[4 * i + 1], kv.y));
                    d = __fadd_rn(d, __fmul_rn(qs[4 * i + 2], kv.z)); d = __fadd_rn(d, __fmul_rn(qs[4 * i + 3], kv.w));
                }
            } else {
                for (int i = 0; i < hd; i++) d = __fadd_rn(d, __fmul_rn(qs[i], ks[i]));
            }
            sc[t] = __fmul_rn(d, a.scale);
        }
    } else {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
        for (int t = warp; t < n; t += nw) {
            const float *kt = t < pos ? a.kc + kv_off(a.table, t, kvd) + g * hd : ks;
            float d = 0.0f;
            for (int i = lane; i < hd; i += 32) d = fmaf(qs[i], kt[i], d);
            d = warp_sum(d);
            if (lane == 0) sc[t] = d * a.scale;
        }
    }
    __syncthreads();
    /* softmax, c/qwen_tts_kernels.c:371-378 */
    float mx = -3.402823466e+38f;
    for (int t = threadIdx.x; t < n; t += blockDim.x) mx = fmaxf(mx, sc[t]);
    mx = block_max(mx, red);
    if (ARITH == QTC_STRICT) {
        for (int t = threadIdx.x; t < n; t += blockDim.x) sc[t] = expf_libm(__fsub_rn(sc[t], mx));
        __syncthreads();
        if (threadIdx.x == 0) {
            float tot = 0.0f;
            for (int t = 0; t < n; t++) tot = __fadd_rn(tot, sc[t]);
            s_b[0] = __fdiv_rn(1.0f, tot);
        }
        __syncthreads();
        const float inv = s_b[0];
        for (int t = threadIdx.x; t < n; t += blockDim.x) sc[t] = __fmul_rn(sc[t], inv);
    } else {
        float tot = 0.0f;
        for (int t = threadIdx.x; t < n; t += blockDim.x) { const float e = __expf(sc[t] - mx); sc[t] = e; tot += e; }
        tot = block_sum(tot, red);
        const float inv = 1.0f / tot;
        for (int t = threadIdx.x; t < n; t += blockDim.x) sc[t] *= inv;
    }
    __syncthreads();
    /* weighted sum of V, t ascending */
    for (int i = threadIdx.x; i < hd; i += blockDim.x) {
        float o = 0.0f;
        const float *vcol = a.vc + g * hd + i;
        if (ARITH == QTC_STRICT) {
            for (int t = 0; t < pos; t++) o = __fadd_rn(o, __fmul_rn(sc[t], vcol[kv_off(a.table, t, kvd)]));
            o = __fadd_rn(o, __fmul_rn(sc[pos], vr[i]));
        } else {
            float o1 = 0.0f;
            int t = 0;
            for (; t + 1 < pos; t += 2) { o = fmaf(sc[t], vcol[kv_off(a.table, t, kvd)], o); o1 = fmaf(sc[t + 1], vcol[kv_off(a.table, t + 1, kvd)], o1); }
            for (; t < pos; t++) o = fmaf(sc[t], vcol[kv_off(a.table, t, kvd)], o);
            o = fmaf(sc[pos], vr[i], o + o1);
        }
        a.att[h * hd + i] = o;
    }
}

/* out[i] = f32(table[idx, i]); idx from device memory when idx_ptr != NULL (kernel_bf16_to_f32 of one row) */
__global__ void k_embed_row(float *out, const uint16_t *table, const int *idx_ptr, int idx_imm, int n) {
    const int idx = idx_ptr ? *idx_ptr : idx_imm;
    const uint16_t *row = table + (size_t)idx * n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = bf16_to_f32(row[i]);
}

__global__ void k_copy_pad(float *out, const float *in, int n_in, int n_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_out) out[i] = i < n_in ? in[i] : 0.0f;
}

/* SiLU in place (text projection, c/qwen_tts.c:836-838) */
template <int ARITH>
__global__ void k_silu(float *x, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float f = x[i];
    x[i] = ARITH == QTC_STRICT ? __fdiv_rn(f, __fadd_rn(1.0f, expf_libm(-f))) : f / (1.0f + __expf(-f));
}

struct NextEmbedArgs {
    float *out;                       /* [H] */
    const uint16_t *codec_embedding;  /* talker table */
    const uint16_t *sub_emb[QTC_MAX_GROUPS - 1];
    const int *frame_codes;           /* device [G]: this frame's codes (talker token + sub-talker codes) */
    int *all_codes;                   /* device [cap, G]: row state->step receives a copy */
    const float *trailing, *tts_pad;
    GenState *st;
    int H, G;
    /* batched launch (blockIdx.y = utterance): strides between utterances, 0 for one utterance */
    int bs_fc, bs_codes, bs_tr;
};
/* c/qwen_tts.c:1345-1363: zero + talker row + sub rows (g ascending) + text row; one chain per element,
 * the same order in both arithmetic modes.  Also advances step (n_generated). */
__global__ void k_next_embed(const NextEmbedArgs a0) {
    pdl_enter();
    NextEmbedArgs a = a0;
    {
        const size_t b = blockIdx.y;
        a.out += b * a.H; a.frame_codes += b * a.bs_fc; a.all_codes += b * a.bs_codes; a.trailing += b * a.bs_tr; a.st += b;
    }
    if (a.st->done) return;
    const int step = a.st->step;                 /* index of the frame just produced */
    const int *codes = a.frame_codes;
    if (blockIdx.x == 0 && threadIdx.x < a.G) a.all_codes[(size_t)step * a.G + threadIdx.x] = codes[threadIdx.x];
    const float *txt = step < a.st->n_trailing ? a.trailing + (size_t)step * a.H : a.tts_pad;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.H; i += gridDim.x * blockDim.x) {
        float acc = __fadd_rn(0.0f, bf16_to_f32(a.codec_embedding[(size_t)codes[0] * a.H + i]));
        for (int g = 1; g < a.G; g++) acc = __fadd_rn(acc, bf16_to_f32(a.sub_emb[g - 1][(size_t)codes[g] * a.H + i]));
        a.out[i] = __fadd_rn(acc, txt[i]);
    }
}
/* separate 1-thread kernel so every CTA of k_next_embed saw the same step */
__global__ void k_advance_step(GenState *st0) {
    pdl_enter();
    GenState *st = st0 + blockIdx.x;
    if (st->done) return;
    st->step += 1;
    if (st->step >= st->limit) { st->done = 2; st->stop_step = st->limit; }
}
