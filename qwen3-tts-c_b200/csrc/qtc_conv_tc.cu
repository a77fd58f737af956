/*
 * qtc_conv_tc.cu - the vocoder's convolutions on the 5th-generation tensor cores ("dense-contraction
 * rule applies: tensor cores"): kernel_causal_conv1d and kernel_transposed_conv1d (c/qwen_tts_kernels.c:659-972) as implicit GEMMs
 *
 *      D[t, co] = sum_taps  A[t - shift_tap, 0:Cin] . W_tap[co, 0:Cin]^T          (one tcgen05 accumulation chain per output tile)
 *
 * Layout.  Inside the vocoder activations are TIME-MAJOR [L, C]: a tile of 128 output samples is the UMMA M side, the input
 * channels are K, and a tap of a (dilated) causal convolution is the same K-major operand fetched at a smaller time coordinate -
 * the TMA box simply starts (K-1-k)*dil rows earlier, and rows before t = 0 come back as zeros (out-of-bounds fill), which IS the
 * causal left padding.  A transposed convolution with stride s is s independent phase convolutions with J = ceil(k/s) taps whose
 * output rows interleave (row m*s + p): same kernel, blockIdx.z = phase.
 *
 * Precision.  The reference computes in fp32.  Every activation and weight is carried as TWO bf16 pieces (hi = rn(x),
 * lo = rn(x - hi): 16 mantissa bits) and each product as three MMAs (hi.hi + hi.lo + lo.hi) accumulated in fp32 in TMEM - the
 * hi.hi products in one accumulator, the two cross terms in a second one that the epilogue adds (the accumulate step of the
 * tensor core drops low bits of the running sum, an error that grows with the number of MMAs into the same accumulator:
 * 93.7 dB at K = 7168 with one accumulator, 102.3 dB with two).  Measured 102-110 dB per layer and 82-85 dB through the whole
 * decoder against the reference's fp32 path (tests/test_gpu_conv_tc.py, test_gpu_fullshape_parity.py; profiles/r02_snr_*).
 * That is 3 of 2250 TFLOP/s-class MMAs per product instead of 1 of 74 TFLOP/s-class FFMAs.
 *
 * Fusion.  SnakeBeta is applied where a tensor is PRODUCED, not where it is consumed: the epilogue adds bias and residual,
 * stores the raw fp32 row (only where a later residual needs it), applies the NEXT layer's SnakeBeta and stores the result as
 * the two bf16 planes the next convolution's TMA reads.  So a ResUnit (Snake, conv7, Snake, conv1, +res) is two launches and no
 * element-wise pass; nothing is transposed between layers.
 *
 * Kernel.  Persistent, one CTA per SM; a tile = 128 samples x N output channels (N <= 128, a divisor of Cout).  Warp 0: TMA
 * producer (A hi/lo boxes of [128 x 64], B hi/lo boxes of [N x 64] per 64-channel block and tap, 2-4 stage mbarrier pipeline
 * that runs on across tiles); warp 1: MMA issuer (12 tcgen05.mma.kind::f16 per stage) + TMEM owner; warps 2-5: epilogue
 * (tcgen05.ld 32x32b.x16, one output row per thread).  The accumulator is double-buffered in TMEM, so the epilogue of a tile
 * (sinf per element) overlaps the MMAs of the next (one tile per CTA and one CTA per SM measured 20 % tensor-pipe activity on
 * the 96-channel layers: profiles/r02_ncu_full_k_conv_tc_v1.csv).
 */
#include "qtc_internal.h"
#include "qtc_device.cuh"
#include "qtc_tc.cuh"

#define CT_BK 64                        /* K per stage: 64 bf16 = one 128-byte swizzle row */
#define CT_A_BYTES (TC_BM * CT_BK * 2)  /* one plane of the A tile: 16 KB */
#define CT_EPI_WARPS 8                   /* two warps per TMEM lane quarter, each takes half of the tile's columns */
#define CT_THREADS (64 + 32 * CT_EPI_WARPS)

struct ConvTcArgs {
    int L;                  /* input rows (samples) */
    int Cin, Cout, N;       /* channels; N = output channels per CTA */
    int taps, dil;          /* taps per phase; tap j reads input row t - (taps-1-j)*dil */
    int os;                 /* output row = t*os + phase (1: plain conv; s: transposed conv, gridDim.z = s phases) */
    int stages, kb_per_tap; /* pipeline depth; ceil(Cin / 64) */
    const float *bias;      /* [Cout] or NULL */
    const float *res;       /* [Lout, Cout] fp32 or NULL (may alias out_raw: each thread reads its element before it writes it) */
    float *out_raw;         /* [Lout, Cout] fp32 or NULL */
    uint16_t *out_hi, *out_lo;          /* [Lout, Cout] bf16 planes of act(out), or NULL */
    const float *alpha, *inv_beta;      /* SnakeBeta of the NEXT layer, per output channel; NULL: planes hold out itself */
    int aplanes;                        /* 2: activations as hi + lo planes; 3: hi + mid + lo, the EXACT split of an fp32 value (prefill: products exact) */
    int wplanes;                        /* 2: weights as hi + lo planes (fp32 originals); 1: the weights ARE bf16 (talker / sub-talker): no lo plane, 2 MMAs */
    const float *scale;                 /* [Cout] or NULL: LayerScale / gamma, applied to (acc + bias) before the residual */
    int gelu;                           /* tanh-GELU after the bias (c/qwen_tts_kernels.c:244-249) */
};

__device__ __forceinline__ float ct_snake(float x, float a, float ib) {
    const float s = sinf(x * a);
    return fmaf(ib * s, s, x);
}
__device__ __forceinline__ uint32_t ct_bf16_rn(float v) {
    const uint32_t u = __float_as_uint(v);
    return (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
}

/* Persistent: grid = min(tiles, SMs); CTA walks tiles (n tile fastest, then time, then phase) with the TMA / MMA warps running
 * ahead into the next tile while the epilogue warps drain the previous accumulator (two TMEM buffers).
 * mapA*: activation planes [L, Cin]; mapW*: weight planes [phases*taps*Cout, Cin] */
__global__ void __launch_bounds__(CT_THREADS, 1) k_conv_tc(const __grid_constant__ CUtensorMap mapAh, const __grid_constant__ CUtensorMap mapAl,
                                                            const __grid_constant__ CUtensorMap mapWh, const __grid_constant__ CUtensorMap mapWl,
                                                            const __grid_constant__ CUtensorMap mapAx, const ConvTcArgs a) {
    extern __shared__ __align__(1024) unsigned char ct_smem_raw[];
    __shared__ __align__(8) uint64_t full[4], empty[4], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_s;
    unsigned char *tiles = reinterpret_cast<unsigned char *>(((uintptr_t)ct_smem_raw + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b_bytes = a.N * CT_BK * 2, a_bytes = a.aplanes * CT_A_BYTES, stage_bytes = a_bytes + a.wplanes * b_bytes;
    const int iters = a.taps * a.kb_per_tap;
    /* TMEM: two buffers (tile j, tile j + 1), each with TWO accumulators: hi.hi products go to the first, the two cross terms
     * (2^-8 of the magnitude) to the second; the epilogue adds them in fp32.  The tensor core's accumulate step loses low bits
     * of the running sum (toward zero: the error grows ~linearly with the number of MMAs, 93.7 dB at K = 7168 with one
     * accumulator), so keeping two of every three MMAs out of the large sum is worth ~9 dB on the long contractions. */
    const uint32_t acc_cols = a.N <= 32 ? 32u : a.N <= 64 ? 64u : 128u, tmem_cols = 4 * acc_cols;
    const int tiles_n = a.Cout / a.N, tiles_t = (a.L + TC_BM - 1) / TC_BM, ntiles = tiles_n * tiles_t * a.os;

    if (threadIdx.x == 0) {
        for (int i = 0; i < a.stages; i++) { tc_mbar_init(&full[i], 1); tc_mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; i++) { tc_mbar_init(&acc_full[i], 1); tc_mbar_init(&acc_empty[i], CT_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapAh) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapAl) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapWh) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapWl) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapAx) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(&tmem_base_s)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;

    if (warp == 0) {
        if (lane == 0) {
            int g = 0;                                                              /* pipeline slot counter across tiles */
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const int ny = tile % tiles_n, rest = tile / tiles_n, tx = rest % tiles_t, phase = rest / tiles_t;
                const int t0 = tx * TC_BM, n0 = ny * a.N;
                for (int it = 0; it < iters; it++, g++) {
                    const int st = g % a.stages, ph = (g / a.stages) & 1;
                    const int tap = it / a.kb_per_tap, kb = it - tap * a.kb_per_tap;
                    if (g >= a.stages) tc_mbar_wait(&empty[st], ph ^ 1);
                    unsigned char *sa = tiles + (size_t)st * stage_bytes, *sb = sa + a_bytes;
                    tc_mbar_expect_tx(&full[st], stage_bytes);
                    const int arow = t0 - (a.taps - 1 - tap) * a.dil;              /* may be negative: rows before t = 0 are zeros */
                    const int wrow = (phase * a.taps + tap) * a.Cout + n0;
                    tc_tma_2d(sa, &mapAh, kb * CT_BK, arow, &full[st]);
                    tc_tma_2d(sa + CT_A_BYTES, &mapAl, kb * CT_BK, arow, &full[st]);
                    if (a.aplanes == 3) tc_tma_2d(sa + 2 * CT_A_BYTES, &mapAx, kb * CT_BK, arow, &full[st]);
                    tc_tma_2d(sb, &mapWh, kb * CT_BK, wrow, &full[st]);
                    if (a.wplanes == 2) tc_tma_2d(sb + b_bytes, &mapWl, kb * CT_BK, wrow, &full[st]);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = tc_idesc(a.N);
            int g = 0, j = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, j++) {
                const int buf = j & 1;
                if (j >= 2) { tc_mbar_wait(&acc_empty[buf], ((j >> 1) & 1) ^ 1); asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
                const uint32_t acc = tmem_d + (uint32_t)buf * 2 * acc_cols, acc2 = acc + acc_cols;
                for (int it = 0; it < iters; it++, g++) {
                    const int st = g % a.stages, ph = (g / a.stages) & 1;
                    tc_mbar_wait(&full[st], ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = tc_smem_u32(tiles + (size_t)st * stage_bytes), sb = sa + a_bytes;
                    const uint64_t ah = tc_smem_desc(sa), al = tc_smem_desc(sa + CT_A_BYTES), ax = tc_smem_desc(sa + 2 * CT_A_BYTES);
                    const uint64_t bh = tc_smem_desc(sb), bl = tc_smem_desc(sb + b_bytes);
#pragma unroll
                    for (int k = 0; k < CT_BK / 16; k++) {       /* UMMA K = 16 bf16 = 32 bytes inside the swizzle row */
                        const uint64_t o = (uint64_t)(k * 2);
                        tc_mma(acc, ah + o, bh + o, idesc, (it | k) != 0);
                        if (a.wplanes == 2) { tc_mma(acc2, ah + o, bl + o, idesc, (it | k) != 0); tc_mma(acc2, al + o, bh + o, idesc, 1); }
                        else {
                            tc_mma(acc2, al + o, bh + o, idesc, (it | k) != 0);
                            if (a.aplanes == 3) tc_mma(acc2, ax + o, bh + o, idesc, 1);
                        }
                    }
                    tc_commit(&empty[st]);
                }
                tc_commit(&acc_full[buf]);
            }
        }
    } else {
        /* epilogue: warp w may touch TMEM lanes [32 (w % 4), +32); thread = one output row, half of its columns (the residual
         * rows of the next 16 columns are requested before the current 16 are processed: the chain per chunk was one L2/HBM round
         * trip long and the k = 1 convolutions were bound by it) */
        const int q = warp & 3, half = (warp - 2) >> 2, ncol = a.N / 2, cbeg = half * ncol, cend = cbeg + ncol;
        int j = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, j++) {
            const int ny = tile % tiles_n, rest = tile / tiles_n, tx = rest % tiles_t, phase = rest / tiles_t;
            const int n0 = ny * a.N, t = tx * TC_BM + q * 32 + lane, buf = j & 1;
            const bool live = t < a.L;
            const size_t orow = ((size_t)t * a.os + phase) * a.Cout + n0;
            tc_mbar_wait(&acc_full[buf], (j >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            float4 rn[4];
            if (a.res && live) {
                const float4 *r4 = reinterpret_cast<const float4 *>(a.res + orow + cbeg);
#pragma unroll
                for (int jj = 0; jj < 4; jj++) rn[jj] = r4[jj];
            }
            for (int c0 = cbeg; c0 < cend; c0 += 16) {
                uint32_t v[16], v2[16];
                float4 rc[4];
#pragma unroll
                for (int jj = 0; jj < 4; jj++) rc[jj] = rn[jj];
                if (a.res && live && c0 + 16 < cend) {
                    const float4 *r4 = reinterpret_cast<const float4 *>(a.res + orow + c0 + 16);
#pragma unroll
                    for (int jj = 0; jj < 4; jj++) rn[jj] = r4[jj];
                }
                const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * 2 * acc_cols + (uint32_t)c0;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                               "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                             : "r"(taddr));
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                             : "=r"(v2[0]), "=r"(v2[1]), "=r"(v2[2]), "=r"(v2[3]), "=r"(v2[4]), "=r"(v2[5]), "=r"(v2[6]), "=r"(v2[7]),
                               "=r"(v2[8]), "=r"(v2[9]), "=r"(v2[10]), "=r"(v2[11]), "=r"(v2[12]), "=r"(v2[13]), "=r"(v2[14]), "=r"(v2[15])
                             : "r"(taddr + acc_cols));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (c0 + 16 >= cend) {
                    /* the accumulator is in registers: hand the buffer back before the (long) SnakeBeta part of the last chunk */
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc_smem_u32(&acc_empty[buf])) : "memory");
                }
                if (!live) continue;
                float y[16];
#pragma unroll
                for (int jj = 0; jj < 16; jj++) y[jj] = (__uint_as_float(v[jj]) + __uint_as_float(v2[jj])) + (a.bias ? a.bias[n0 + c0 + jj] : 0.0f);
                if (a.gelu) {
#pragma unroll
                    for (int jj = 0; jj < 16; jj++) y[jj] = 0.5f * y[jj] * (1.0f + tanhf(0.7978845608028654f * (y[jj] + 0.044715f * y[jj] * y[jj] * y[jj])));
                }
                if (a.scale) {
#pragma unroll
                    for (int jj = 0; jj < 16; jj++) y[jj] *= a.scale[n0 + c0 + jj];
                }
                if (a.res) {
#pragma unroll
                    for (int jj = 0; jj < 4; jj++) { const float4 r = rc[jj]; y[4 * jj] += r.x; y[4 * jj + 1] += r.y; y[4 * jj + 2] += r.z; y[4 * jj + 3] += r.w; }
                }
                if (a.out_raw) {
                    float4 *o4 = reinterpret_cast<float4 *>(a.out_raw + orow + c0);
#pragma unroll
                    for (int jj = 0; jj < 4; jj++) o4[jj] = make_float4(y[4 * jj], y[4 * jj + 1], y[4 * jj + 2], y[4 * jj + 3]);
                }
                if (a.out_hi) {
                    uint32_t hi[8], lo[8];
#pragma unroll
                    for (int jj = 0; jj < 16; jj += 2) {
                        float z0 = y[jj], z1 = y[jj + 1];
                        if (a.alpha) { z0 = ct_snake(z0, a.alpha[n0 + c0 + jj], a.inv_beta[n0 + c0 + jj]); z1 = ct_snake(z1, a.alpha[n0 + c0 + jj + 1], a.inv_beta[n0 + c0 + jj + 1]); }
                        const uint32_t h0 = ct_bf16_rn(z0), h1 = ct_bf16_rn(z1);
                        const uint32_t l0 = ct_bf16_rn(z0 - __uint_as_float(h0 << 16)), l1 = ct_bf16_rn(z1 - __uint_as_float(h1 << 16));
                        hi[jj >> 1] = h0 | (h1 << 16); lo[jj >> 1] = l0 | (l1 << 16);
                    }
                    uint4 *ph = reinterpret_cast<uint4 *>(a.out_hi + orow + c0), *pl = reinterpret_cast<uint4 *>(a.out_lo + orow + c0);
                    ph[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]); ph[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                    pl[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]); pl[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(tmem_cols) : "memory");
    }
}

/* ---- element-wise helpers around the tensor-core path ------------------------------------------------------------------ */
/* channels-first fp32 [C, L] -> time-major bf16 planes [L, C] (hi, lo) of act(x); 32 x 32 tiles through shared memory */
__global__ void k_cf_to_planes(uint16_t *hi, uint16_t *lo, const float *src, int C, int L, const float *alpha, const float *inv_beta) {
    __shared__ float tile[32][33];
    const int l0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, l = l0 + threadIdx.x;
        float v = (c < C && l < L) ? src[(size_t)c * L + l] : 0.0f;
        if (alpha && c < C) v = ct_snake(v, alpha[c], inv_beta[c]);
        tile[i][threadIdx.x] = v;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int l = l0 + i, c = c0 + threadIdx.x;
        if (l < L && c < C) {
            const float v = tile[threadIdx.x][i];
            const uint32_t h = ct_bf16_rn(v);
            hi[(size_t)l * C + c] = (uint16_t)h;
            lo[(size_t)l * C + c] = (uint16_t)ct_bf16_rn(v - __uint_as_float(h << 16));
        }
    }
}
/* fp32 rows -> bf16 planes (hi, lo), element-wise; n % 4 == 0 */
__global__ void k_rows_to_planes(uint16_t *hi, uint16_t *lo, const float *src, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 v = reinterpret_cast<const float4 *>(src)[i];
    const uint32_t h0 = ct_bf16_rn(v.x), h1 = ct_bf16_rn(v.y), h2 = ct_bf16_rn(v.z), h3 = ct_bf16_rn(v.w);
    const uint32_t l0 = ct_bf16_rn(v.x - __uint_as_float(h0 << 16)), l1 = ct_bf16_rn(v.y - __uint_as_float(h1 << 16));
    const uint32_t l2 = ct_bf16_rn(v.z - __uint_as_float(h2 << 16)), l3 = ct_bf16_rn(v.w - __uint_as_float(h3 << 16));
    reinterpret_cast<uint2 *>(hi)[i] = make_uint2(h0 | (h1 << 16), h2 | (h3 << 16));
    reinterpret_cast<uint2 *>(lo)[i] = make_uint2(l0 | (l1 << 16), l2 | (l3 << 16));
}
/* depthwise causal conv, time-major [L, C] (ConvNeXt dwconv k = 7, c/qwen_tts_codec.c:477): out[t, c] = b[c] + sum_k w[c, k] in[t - (K-1-k), c] */
__global__ void k_dwconv_tm(float *out, const float *in, const float *w, const float *bias, int C, int K, int L) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)L * C) return;
    const int t = (int)(i / C), c = (int)(i - (size_t)t * C);
    float acc = bias ? bias[c] : 0.0f;
    for (int k = 0; k < K; k++) {
        const int tt = t - (K - 1 - k);
        if (tt >= 0) acc = fmaf(w[c * K + k], in[(size_t)tt * C + c], acc);
    }
    out[i] = acc;
}

/* final conv (Cout = 1, k = 7, c/qwen_tts_codec.c:727-735) on time-major planes that already hold SnakeBeta(x): a CTA stages the
 * 256 + K - 1 rows its samples need in shared memory as fp32 (coalesced 16-byte loads of both planes), then a thread walks the
 * K x C window of one sample (row stride C + 1 floats: conflict-free); clamp fused */
#define MONO_T 256
__global__ void __launch_bounds__(MONO_T) k_mono_from_planes(float *out, const uint16_t *hi, const uint16_t *lo, const float *w /*[C, K]*/, const float *bias, int C, int K, int L) {
    extern __shared__ float msm[];                       /* [K][C] weights, then [MONO_T + K - 1][C + 1] rows */
    float *wsm = msm, *rows = msm + K * C;
    const int t0 = blockIdx.x * MONO_T, nrows = MONO_T + K - 1, C8 = C >> 3;
    for (int i = threadIdx.x; i < C * K; i += MONO_T) { const int c = i / K, k = i - c * K; wsm[k * C + c] = w[i]; }
    for (int i = threadIdx.x; i < nrows * C8; i += MONO_T) {
        const int r = i / C8, c8 = i - r * C8, t = t0 - (K - 1) + r;
        float *d = rows + r * (C + 1) + c8 * 8;
        if (t >= 0 && t < L) {
            const uint4 h = *reinterpret_cast<const uint4 *>(hi + (size_t)t * C + c8 * 8), l = *reinterpret_cast<const uint4 *>(lo + (size_t)t * C + c8 * 8);
            d[0] = bf16_lo(h.x) + bf16_lo(l.x); d[1] = bf16_hi(h.x) + bf16_hi(l.x); d[2] = bf16_lo(h.y) + bf16_lo(l.y); d[3] = bf16_hi(h.y) + bf16_hi(l.y);
            d[4] = bf16_lo(h.z) + bf16_lo(l.z); d[5] = bf16_hi(h.z) + bf16_hi(l.z); d[6] = bf16_lo(h.w) + bf16_lo(l.w); d[7] = bf16_hi(h.w) + bf16_hi(l.w);
        } else {
#pragma unroll
            for (int e = 0; e < 8; e++) d[e] = 0.0f;
        }
    }
    __syncthreads();
    const int t = t0 + threadIdx.x;
    if (t >= L) return;
    float acc = bias ? bias[0] : 0.0f;
    for (int c = 0; c < C; c++)
        for (int k = 0; k < K; k++) acc = fmaf(wsm[k * C + c], rows[(threadIdx.x + k) * (C + 1) + c], acc);
    out[t] = fminf(1.0f, fmaxf(-1.0f, acc));
}

/* ---- host side ---------------------------------------------------------------------------------------------------------- */
static void split_bf16(const float *src, size_t n, uint16_t *hi, uint16_t *lo) {
    for (size_t i = 0; i < n; i++) {
        uint32_t u; memcpy(&u, &src[i], 4);
        const uint32_t h = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
        const uint32_t hb = h << 16; float hf; memcpy(&hf, &hb, 4);
        const float r = src[i] - hf;
        uint32_t ur; memcpy(&ur, &r, 4);
        hi[i] = (uint16_t)h; lo[i] = (uint16_t)((ur + 0x7fffu + ((ur >> 16) & 1u)) >> 16);
    }
}
/* causal conv weight [Cout, Cin, K] (c/qwen_tts_kernels.c:659) -> tap-major planes [K*Cout, Cin] */
int qtc_conv_tc_prep(qtc_ctx *c, ConvTcW *o, const float *w, int cout, int cin, int K) {
    std::vector<float> r((size_t)K * cout * cin);
    for (int oc = 0; oc < cout; oc++)
        for (int ic = 0; ic < cin; ic++)
            for (int k = 0; k < K; k++) r[((size_t)k * cout + oc) * cin + ic] = w[((size_t)oc * cin + ic) * K + k];
    std::vector<uint16_t> hi(r.size()), lo(r.size());
    split_bf16(r.data(), r.size(), hi.data(), lo.data());
    o->hi = qtc_upload(c, hi.data(), hi.size()); o->lo = qtc_upload(c, lo.data(), lo.size());
    o->cin = cin; o->cout = cout; o->taps = K; o->phases = 1;
    return (o->hi && o->lo) ? 0 : -1;
}
/* transposed conv weight [Cin, Cout, k], stride s (c/qwen_tts_kernels.c:873-972): phase p of the output (rows m*s + p) is a
 * J = ceil(k/s)-tap causal conv of the input with taps w[:, :, p + (J-1-j)*s] -> planes [(p*J + j)*Cout + oc, ic] */
int qtc_tconv_tc_prep(qtc_ctx *c, ConvTcW *o, const float *w, int cin, int cout, int k, int s) {
    const int J = (k + s - 1) / s;
    std::vector<float> r((size_t)s * J * cout * cin, 0.0f);
    for (int p = 0; p < s; p++)
        for (int j = 0; j < J; j++) {
            const int tap = p + (J - 1 - j) * s;
            if (tap >= k) continue;
            for (int oc = 0; oc < cout; oc++)
                for (int ic = 0; ic < cin; ic++) r[(((size_t)p * J + j) * cout + oc) * cin + ic] = w[((size_t)ic * cout + oc) * k + tap];
        }
    std::vector<uint16_t> hi(r.size()), lo(r.size());
    split_bf16(r.data(), r.size(), hi.data(), lo.data());
    o->hi = qtc_upload(c, hi.data(), hi.size()); o->lo = qtc_upload(c, lo.data(), lo.size());
    o->cin = cin; o->cout = cout; o->taps = J; o->phases = s;
    return (o->hi && o->lo) ? 0 : -1;
}

/* N: the largest multiple of 32 (two epilogue halves of whole 16-column chunks) <= 128 (four accumulators of <= 128 TMEM
 * columns: two tiles in flight x {hi.hi, cross terms}) that divides Cout */
static int pick_n(int cout) {
    for (int n = 128; n >= 32; n -= 32) if (cout % n == 0) return n;
    return 0;
}
bool qtc_conv_tc_ok(int cin, int cout) { return cin % 8 == 0 && cout % 32 == 0 && pick_n(cout) > 0; }

static int g_ct_attr = 0;
/* out rows = L * phases.  in_hi/in_lo: [L, Cin] planes.  Any of out_raw / out_hi may be NULL; alpha NULL: planes hold the raw output */
int qtc_conv_tc(qtc_ctx *c, const ConvTcW *w, const uint16_t *in_hi, const uint16_t *in_lo, int L, int dil, const float *bias, const float *res,
                float *out_raw, uint16_t *out_hi, uint16_t *out_lo, const float *alpha, const float *inv_beta, const float *scale, int gelu,
                const uint16_t *in_x) {
    if (L <= 0) return 0;
    ConvTcArgs a;
    memset(&a, 0, sizeof(a));
    a.scale = scale; a.gelu = gelu; a.wplanes = w->lo ? 2 : 1; a.aplanes = in_x ? 3 : 2;
    if (in_x && w->lo) { qtc_set_error("conv_tc: a third activation plane goes with single-plane (bf16) weights only"); return -1; }
    a.L = L; a.Cin = w->cin; a.Cout = w->cout; a.N = pick_n(w->cout); a.taps = w->taps; a.dil = dil; a.os = w->phases;
    a.kb_per_tap = (w->cin + CT_BK - 1) / CT_BK;
    a.bias = bias; a.res = res; a.out_raw = out_raw; a.out_hi = out_hi; a.out_lo = out_lo; a.alpha = alpha; a.inv_beta = inv_beta;
    if (a.N <= 0 || w->cin % 8) { qtc_set_error("conv_tc: Cout %d / Cin %d not tileable", w->cout, w->cin); return -1; }
    const int stage_bytes = a.aplanes * CT_A_BYTES + a.wplanes * a.N * CT_BK * 2;
    /* one persistent CTA per SM: as many stages as fit (2-4) */
    int stages = (220 * 1024) / stage_bytes;
    if (stages > 4) stages = 4;
    if (stages < 2) stages = 2;
    a.stages = stages;
    const size_t smem = (size_t)stages * stage_bytes + 1024;
    if (!g_ct_attr) { QTC_CUDA(cudaFuncSetAttribute(k_conv_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024)); g_ct_attr = 1; }
    const CUtensorMap *mah = qtc_tmap_bf16(in_hi, L, w->cin, TC_BM), *mal = qtc_tmap_bf16(in_lo, L, w->cin, TC_BM);
    const int wrows = w->phases * w->taps * w->cout;
    const CUtensorMap *mwh = qtc_tmap_bf16(w->hi, wrows, w->cin, a.N), *mwl = w->lo ? qtc_tmap_bf16(w->lo, wrows, w->cin, a.N) : mwh;
    const CUtensorMap *max_ = in_x ? qtc_tmap_bf16(in_x, L, w->cin, TC_BM) : mah;
    if (!mah || !mal || !mwh || !mwl || !max_) return -1;
    const long long ntiles = (long long)((L + TC_BM - 1) / TC_BM) * (w->cout / a.N) * w->phases;
    const int sms = c->num_sms > 0 ? c->num_sms : 148;
    const dim3 grid((unsigned)(ntiles < sms ? ntiles : sms));
    k_conv_tc<<<grid, CT_THREADS, smem, c->cstream>>>(*mah, *mal, *mwh, *mwl, *max_, a);
    c->launches++;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { qtc_set_error("conv_tc launch -> %s", cudaGetErrorString(e)); return -1; }
    return 0;
}
int qtc_cf_to_planes(qtc_ctx *c, uint16_t *hi, uint16_t *lo, const float *src, int C, int L, const float *alpha, const float *inv_beta) {
    k_cf_to_planes<<<dim3((L + 31) / 32, (C + 31) / 32), dim3(32, 8), 0, c->cstream>>>(hi, lo, src, C, L, alpha, inv_beta);
    c->launches++;
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
int qtc_rows_to_planes(qtc_ctx *c, uint16_t *hi, uint16_t *lo, const float *src, size_t n) {
    if (n % 4) { qtc_set_error("rows_to_planes: n %% 4"); return -1; }
    k_rows_to_planes<<<(unsigned)((n / 4 + 255) / 256), 256, 0, c->cstream>>>(hi, lo, src, n / 4);
    c->launches++;
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
int qtc_dwconv_tm(qtc_ctx *c, float *out, const float *in, const float *w, const float *bias, int C, int K, int L) {
    const size_t n = (size_t)L * C;
    k_dwconv_tm<<<(unsigned)((n + 255) / 256), 256, 0, c->cstream>>>(out, in, w, bias, C, K, L);
    c->launches++;
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
int qtc_mono_from_planes(qtc_ctx *c, float *out, const uint16_t *hi, const uint16_t *lo, const float *w, const float *bias, int C, int K, int L) {
    if (C % 8) { qtc_set_error("mono_from_planes: C %% 8"); return -1; }
    const size_t smem = ((size_t)C * K + (size_t)(MONO_T + K - 1) * (C + 1)) * 4;
    static int attr = 0;
    if (!attr) { if (cudaFuncSetAttribute(k_mono_from_planes, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) return -1; attr = 1; }
    if (smem > 200 * 1024) { qtc_set_error("mono_from_planes: %d channels do not fit the row stage", C); return -1; }
    k_mono_from_planes<<<(L + MONO_T - 1) / MONO_T, MONO_T, smem, c->cstream>>>(out, hi, lo, w, bias, C, K, L);
    c->launches++;
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
