/*
 * qtc_codec_kernels.cuh - kernels of the codec decoder / vocoder (12.5 Hz codes
 * -> 24 kHz PCM), replacing c/qwen_tts_codec.c + the conv / SnakeBeta / LayerNorm
 * / GELU kernels of c/qwen_tts_kernels.c.  Everything fp32 (the reference keeps
 * the codec in f32), FMA accumulation in the reference's (channel, tap) order.
 *
 *  k_conv_tile   causal (dilated) conv1d and transposed conv1d as one register-
 *                blocked direct convolution: [BM out-channels x 128 samples] per
 *                CTA, input slab staged once per channel chunk in shared memory
 *                with SnakeBeta applied on the way in, bias / residual fused in
 *                the epilogue, polyphase scatter for the transposed form.
 *  k_sgemm_nt    C = A . B^T for the transformer / ConvNeXt projections with a
 *                fused epilogue (bias, GELU, LayerScale, residual, transposed store).
 */
#pragma once
#include "qtc_device.cuh"
#include "../../include/qwen_tts_cuda.h"

/* ------------------------------------------------------------------ SGEMM NT */
struct GemmEpi {
    const float *bias;    /* [N] or NULL */
    const float *scale;   /* [N] or NULL: LayerScale / gamma */
    const float *res;     /* residual, indexed like the output, or NULL */
    int gelu;             /* tanh-GELU after bias (c/qwen_tts_kernels.c:244-249) */
    int transposed;       /* store at [n*ldc + m] instead of [m*ldc + n] */
    int ldc;
};

#define SG_BM 128
#define SG_BN 128
#define SG_BK 8

__device__ __forceinline__ float gelu_tanh(float v) {
    return 0.5f * v * (1.0f + tanhf(0.7978845608028654f * (v + 0.044715f * v * v * v)));
}

/* A [M,K] row-major, B [N,K] row-major, K % 4 == 0.  256 threads, 8x8 outputs each. */
__global__ void __launch_bounds__(256) k_sgemm_nt(float *C, const float *A, const float *B, int M, int N, int K, const GemmEpi e) {
    __shared__ __align__(16) float As[2][SG_BK][SG_BM + 4];
    __shared__ __align__(16) float Bs[2][SG_BK][SG_BN + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
    /* loader: thread -> (row 0..127, k-quad 0..1) */
    const int lr = tid >> 1, lk = (tid & 1) * 4;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = 0.0f;
    float4 ra, rb;
    auto gload = [&](int k0) {
        const int am = m0 + lr, bn = n0 + lr;
        ra = (am < M && k0 + lk < K) ? *reinterpret_cast<const float4 *>(A + (size_t)am * K + k0 + lk) : make_float4(0, 0, 0, 0);
        rb = (bn < N && k0 + lk < K) ? *reinterpret_cast<const float4 *>(B + (size_t)bn * K + k0 + lk) : make_float4(0, 0, 0, 0);
    };
    auto sstore = [&](int buf) {
        As[buf][lk + 0][lr] = ra.x; As[buf][lk + 1][lr] = ra.y; As[buf][lk + 2][lr] = ra.z; As[buf][lk + 3][lr] = ra.w;
        Bs[buf][lk + 0][lr] = rb.x; Bs[buf][lk + 1][lr] = rb.y; Bs[buf][lk + 2][lr] = rb.z; Bs[buf][lk + 3][lr] = rb.w;
    };
    gload(0);
    sstore(0);
    __syncthreads();
    int buf = 0;
    for (int k0 = 0; k0 < K; k0 += SG_BK) {
        if (k0 + SG_BK < K) gload(k0 + SG_BK);
#pragma unroll
        for (int kk = 0; kk < SG_BK; kk++) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][kk][ty * 8]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][kk][ty * 8 + 4]);
            const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][tx * 8]);
            const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][tx * 8 + 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
                for (int j = 0; j < 8; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (k0 + SG_BK < K) { sstore(buf ^ 1); __syncthreads(); buf ^= 1; }
    }
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int m = m0 + ty * 8 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int n = n0 + tx * 8 + j;
            if (n >= N) continue;
            float v = acc[i][j];
            if (e.bias) v += e.bias[n];
            if (e.gelu) v = gelu_tanh(v);
            if (e.scale) v *= e.scale[n];
            const size_t idx = e.transposed ? (size_t)n * e.ldc + m : (size_t)m * e.ldc + n;
            if (e.res) v += e.res[idx];
            C[idx] = v;
        }
    }
}

/* ------------------------------------------------------------------ conv tile */
struct ConvArgs {
    const float *in;      /* [Cin, Lin] */
    const float *w;       /* [rows, Cin*K] (transposed convs are repacked to this form at upload) */
    const float *bias;    /* [Cout] or NULL (indexed by out channel, not by virtual row) */
    const float *res;     /* [Cout, Lout] or NULL */
    const float *alpha, *inv_beta;   /* SnakeBeta on the input, per input channel, or NULL */
    float *out;           /* [Cout, Lout] */
    int Cin, rows, K, dil, Lin;
    int up;               /* 1: plain conv (rows = Cout, Lout = Lin); s>1: transposed conv phase count
                             (rows = Cout*s, row = oc*s + p, Lout = Lin*s, element (row, m) -> out[oc, m*s+p]) */
};

#define CV_BN 128
#define CV_MAXCK 56      /* max (channels per chunk) * K: keeps static tiles under 48 KB */

__device__ __forceinline__ float snake_beta(float x, float a, float ib) {
    const float s = sinf(x * a);
    return fmaf(ib * s, s, x);
}

/* threads = (BM/8)*16; thread (ty,tx) owns rows ty*8..+7, samples tx*8..+7.
 * dynamic smem: Ws[CK][BM+4] + slab[CI][CV_BN + halo (+pad)] */
template <int BM>
__global__ void __launch_bounds__((BM / 8) * 16) k_conv_tile(const ConvArgs a) {
    extern __shared__ __align__(16) float smem[];
    const int K = a.K, dil = a.dil, halo = (K - 1) * dil;
    int CI = CV_MAXCK / K; if (CI > 32) CI = 32; if (CI < 1) CI = 1;
    const int CK = CI * K, SW = CV_BN + halo + 1;
    float *Ws = smem;                       /* [CK][BM+4] */
    float *slab = smem + CK * (BM + 4);     /* [CI][SW]   */
    const int nthr = (BM / 8) * 16, tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int t0 = blockIdx.x * CV_BN, r0 = blockIdx.y * BM;
    const int CinK = a.Cin * K;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = 0.0f;

    for (int ic0 = 0; ic0 < a.Cin; ic0 += CI) {
        const int nci = a.Cin - ic0 < CI ? a.Cin - ic0 : CI;
        __syncthreads();
        /* weights: Ws[kk][r] = w[(r0+r), ic0*K + kk] */
        for (int i = tid; i < BM * CK; i += nthr) {
            const int r = i / CK, kk = i - r * CK;
            float v = 0.0f;
            if (r0 + r < a.rows && kk < nci * K) v = a.w[(size_t)(r0 + r) * CinK + (size_t)ic0 * K + kk];
            Ws[kk * (BM + 4) + r] = v;
        }
        /* input slab with the activation applied once: slab[ci][j] = act(in[ic0+ci, t0 - halo + j]) */
        for (int i = tid; i < nci * (CV_BN + halo); i += nthr) {
            const int ci = i / (CV_BN + halo), j = i - ci * (CV_BN + halo);
            const int t = t0 - halo + j;
            float v = 0.0f;
            if (t >= 0 && t < a.Lin) {
                v = a.in[(size_t)(ic0 + ci) * a.Lin + t];
                if (a.alpha) v = snake_beta(v, a.alpha[ic0 + ci], a.inv_beta[ic0 + ci]);
            }
            slab[ci * SW + j] = v;
        }
        __syncthreads();
        for (int ci = 0; ci < nci; ci++) {
            const float *srow = slab + ci * SW + tx * 8;
            for (int k = 0; k < K; k++) {
                /* out[t] += w[k] * in[t - (K-1-k)*dil]  ->  slab offset (t - t0) + k*dil */
                const float *wp = Ws + (ci * K + k) * (BM + 4) + ty * 8;
                const float4 a0 = *reinterpret_cast<const float4 *>(wp), a1 = *reinterpret_cast<const float4 *>(wp + 4);
                const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                float bv[8];
                const float *bp = srow + k * dil;
#pragma unroll
                for (int j = 0; j < 8; j++) bv[j] = bp[j];
#pragma unroll
                for (int i = 0; i < 8; i++)
#pragma unroll
                    for (int j = 0; j < 8; j++) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
        }
    }
    const int s = a.up;
    const size_t Lout = (size_t)a.Lin * s;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int r = r0 + ty * 8 + i;
        if (r >= a.rows) continue;
        const int oc = s > 1 ? r / s : r, p = s > 1 ? r - oc * s : 0;
        const float b = a.bias ? a.bias[oc] : 0.0f;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int t = t0 + tx * 8 + j;
            if (t >= a.Lin) continue;
            const size_t idx = (size_t)oc * Lout + (size_t)t * s + p;
            float v = acc[i][j] + b;
            if (a.res) v += a.res[idx];
            a.out[idx] = v;
        }
    }
}

/* Cout == 1 (final conv, c/qwen_tts_codec.c:727-735): memory-bound, one thread per sample; clamp fused */
__global__ void k_conv_to_mono(float *out, const float *in, const float *w, const float *bias, const float *alpha,
                               const float *inv_beta, int Cin, int K, int L, int clamp) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= L) return;
    float acc = bias ? bias[0] : 0.0f;
    for (int ic = 0; ic < Cin; ic++) {
        const float al = alpha ? alpha[ic] : 0.0f, ib = alpha ? inv_beta[ic] : 0.0f;
        for (int k = 0; k < K; k++) {
            const int tt = t - (K - 1 - k);
            if (tt < 0) continue;
            float v = in[(size_t)ic * L + tt];
            if (alpha) v = snake_beta(v, al, ib);
            acc = fmaf(w[ic * K + k], v, acc);
        }
    }
    if (clamp) acc = fminf(1.0f, fmaxf(-1.0f, acc));
    out[t] = acc;
}

/* depthwise causal conv (groups == channels), ConvNeXt dwconv k=7 (c/qwen_tts_codec.c:477) */
__global__ void k_dwconv(float *out, const float *in, const float *w, const float *bias, int C, int K, int L, int dil) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
    if (t >= L) return;
    float acc = bias ? bias[c] : 0.0f;
    for (int k = 0; k < K; k++) {
        const int tt = t - (K - 1 - k) * dil;
        if (tt >= 0) acc = fmaf(w[c * K + k], in[(size_t)c * L + tt], acc);
    }
    out[(size_t)c * L + t] = acc;
}

/* general grouped conv fallback for the kernel-level test surface (any groups) */
__global__ void k_conv_grouped(float *out, const float *in, const float *w, const float *bias, int Cin, int Cout, int K, int L,
                               int dil, int groups) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x, oc = blockIdx.y;
    if (t >= L) return;
    const int cpg = Cin / groups, opg = Cout / groups, ic0 = (oc / opg) * cpg;
    float acc = bias ? bias[oc] : 0.0f;
    for (int j = 0; j < cpg; j++)
        for (int k = 0; k < K; k++) {
            const int tt = t - (K - 1 - k) * dil;
            if (tt >= 0) acc = fmaf(w[((size_t)oc * cpg + j) * K + k], in[(size_t)(ic0 + j) * L + tt], acc);
        }
    out[(size_t)oc * L + t] = acc;
}

/* SnakeBeta stand-alone (kernel-level test surface only; the decoder fuses it into k_conv_tile) */
__global__ void k_snake(float *out, const float *x, const float *alpha, const float *inv_beta, int C, size_t L) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)C * L) return;
    const int c = (int)(i / L);
    out[i] = snake_beta(x[i], alpha[c], inv_beta[c]);
}

/* dst[c, r] = src[r, c]; 32x32 tiles through shared memory */
__global__ void k_transpose(float *dst, const float *src, int rows, int cols) {
    __shared__ float tile[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < rows && c < cols) ? src[(size_t)r * cols + c] : 0.0f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < cols && r < rows) dst[(size_t)c * rows + r] = tile[threadIdx.x][i];
    }
}

__device__ __forceinline__ void put_planes(uint16_t *ph, uint16_t *pl, size_t i, float v) {
    const uint32_t u = __float_as_uint(v), h = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
    const uint32_t r = __float_as_uint(v - __uint_as_float(h << 16));
    ph[i] = (uint16_t)h; pl[i] = (uint16_t)((r + 0x7fffu + ((r >> 16) & 1u)) >> 16);
}
/* row-wise norms over [rows, n]; one CTA per row.  mode 0: RMSNorm(w), 1: LayerNorm(w,b) (c/qwen_tts_kernels.c:45-58).
 * ph / pl != NULL: the result goes out as bf16 planes (input of a tensor-core GEMM) instead of fp32 */
__global__ void __launch_bounds__(256) k_row_norm(float *out, const float *x, const float *w, const float *b, int n, float eps, int mode,
                                                  uint16_t *ph = NULL, uint16_t *pl = NULL) {
    __shared__ float red[32];
    const float *xi = x + (size_t)blockIdx.x * n;
    float *yo = out + (size_t)blockIdx.x * n;
    const size_t pb = (size_t)blockIdx.x * n;
    if (mode == 0) {
        float ss = 0.0f;
        for (int i = threadIdx.x; i < n; i += blockDim.x) ss = fmaf(xi[i], xi[i], ss);
        ss = block_sum(ss, red);
        const float inv = 1.0f / sqrtf(ss / (float)n + eps);
        for (int i = threadIdx.x; i < n; i += blockDim.x) { const float y = xi[i] * inv * w[i]; if (ph) put_planes(ph, pl, pb + i, y); else yo[i] = y; }
    } else {
        float s = 0.0f;
        for (int i = threadIdx.x; i < n; i += blockDim.x) s += xi[i];
        const float mean = block_sum(s, red) / (float)n;
        float v = 0.0f;
        for (int i = threadIdx.x; i < n; i += blockDim.x) { const float d = xi[i] - mean; v = fmaf(d, d, v); }
        const float var = block_sum(v, red) / (float)n;
        const float inv = 1.0f / sqrtf(var + eps);
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            float y = (xi[i] - mean) * inv;
            if (w) y *= w[i];
            if (b) y += b[i];
            if (ph) put_planes(ph, pl, pb + i, y); else yo[i] = y;
        }
    }
}

/* RVQ de-quantise (c/qwen_tts_codec.c:127-261): sem[t,:] = E0[c0], aco[t,:] = sum_{q>=1} Eq[cq]; out-of-range code -> 0 */
struct RvqArgs { const float *cb[QTC_MAX_QUANTIZERS]; const int *codes; float *sem, *aco; int T, nq, vq, cb_size; };
__global__ void k_rvq_gather(const RvqArgs a) {
    const int t = blockIdx.x;
    for (int d = threadIdx.x; d < a.vq; d += blockDim.x) {
        float s = 0.0f, acc = 0.0f;
        for (int q = 0; q < a.nq; q++) {
            int code = a.codes[(size_t)t * a.nq + q];
            if (code < 0 || code >= a.cb_size) code = 0;
            const float v = a.cb[q][(size_t)code * a.vq + d];
            if (q == 0) s = v; else acc += v;
        }
        a.sem[(size_t)t * a.vq + d] = s;
        a.aco[(size_t)t * a.vq + d] = acc;
    }
}

/* rotate-half RoPE on q and k rows of the codec transformer (c/qwen_tts_codec.c:347-352); cos/sin [T, hd] */
__global__ void k_rope_rows(float *q, float *k, const float *cs, const float *sn, int T, int nh, int nkv, int hd) {
    const int t = blockIdx.x, half = hd / 2;
    for (int i = threadIdx.x; i < (nh + nkv) * half; i += blockDim.x) {
        const int h = i / half, e = i - h * half;
        float *p = h < nh ? q + ((size_t)t * nh + h) * hd : k + ((size_t)t * nkv + (h - nh)) * hd;
        const float a = p[e], b = p[e + half], c = cs[(size_t)t * hd + e], s = sn[(size_t)t * hd + e];
        p[e] = a * c - b * s;
        p[e + half] = b * c + a * s;
    }
}

/* sliding-window causal attention (c/qwen_tts_codec.c:355-383): one warp per (query, head), window <= 128.
 * q / out hold T queries whose absolute positions are q0 .. q0+T-1; k / v are indexed by absolute position (the streaming
 * decoder keeps them across chunks), so a chunk attends to keys of earlier chunks exactly as a one-shot decode does. */
__global__ void __launch_bounds__(128) k_window_attn(float *out, const float *q, const float *k, const float *v, int T, int nh, int nkv,
                                                      int hd, int W, float scale, int q0, uint16_t *ph = NULL, uint16_t *pl = NULL) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x * 4 + warp;
    if (item >= T * nh) return;
    const int ql = item / nh, h = item - ql * nh, g = h / (nh / nkv), qi = q0 + ql;
    int start = qi - W + 1; if (start < 0) start = 0;
    const int n = qi - start + 1;
    const float *qh = q + ((size_t)ql * nh + h) * hd;
    float sc[4];
    float mx = -3.402823466e+38f;
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int i = lane + 32 * r;
        float d = -3.402823466e+38f;
        if (i < n) {
            const float *kh = k + ((size_t)(start + i) * nkv + g) * hd;
            d = 0.0f;
            for (int e = 0; e < hd; e += 4) {
                const float4 a = *reinterpret_cast<const float4 *>(qh + e), b = *reinterpret_cast<const float4 *>(kh + e);
                d = fmaf(a.x, b.x, d); d = fmaf(a.y, b.y, d); d = fmaf(a.z, b.z, d); d = fmaf(a.w, b.w, d);
            }
            d *= scale;
        }
        sc[r] = d; mx = fmaxf(mx, d);
    }
    mx = warp_max(mx);
    float tot = 0.0f;
#pragma unroll
    for (int r = 0; r < 4; r++) { const int i = lane + 32 * r; sc[r] = i < n ? expf(sc[r] - mx) : 0.0f; tot += sc[r]; }
    tot = warp_sum(tot);
    const float inv = 1.0f / tot;
    /* lanes own output dims e = lane, lane+32, ... */
    float o[8];
#pragma unroll
    for (int u = 0; u < 8; u++) o[u] = 0.0f;
    for (int i = 0; i < n; i++) {
        const float p = __shfl_sync(0xffffffffu, sc[i >> 5], i & 31) * inv;
        const float *vh = v + ((size_t)(start + i) * nkv + g) * hd;
#pragma unroll
        for (int u = 0; u < 8; u++) { const int e = lane + 32 * u; if (e < hd) o[u] = fmaf(p, vh[e], o[u]); }
    }
    float *oh = out + ((size_t)ql * nh + h) * hd;
#pragma unroll
    for (int u = 0; u < 8; u++) {
        const int e = lane + 32 * u;
        if (e < hd) { if (ph) put_planes(ph, pl, ((size_t)ql * nh + h) * hd + e, o[u]); else oh[e] = o[u]; }
    }
}

/* dst[i] = src[i]: PCM of a streamed chunk into the utterance's device buffer */
__global__ void k_copy_pcm(float *dst, const float *src, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

/* g[i] = silu(g[i]) * u[i] (c/qwen_tts_codec.c:411-414) */
__global__ void k_swiglu_mul(float *g, const float *u, size_t n, uint16_t *ph = NULL, uint16_t *pl = NULL) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const float x = g[i], y = (x / (1.0f + expf(-x))) * u[i]; if (ph) put_planes(ph, pl, i, y); else g[i] = y; }
}
