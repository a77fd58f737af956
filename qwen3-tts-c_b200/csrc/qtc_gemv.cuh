/*
 * qtc_gemv.cuh - bf16 weight-streaming matrix-vector kernels (batch 1).
 *
 * Replaces kernel_matvec_bf16 (c/qwen_tts_kernels.c:95-149) and
 * kernel_swiglu_matvec_bf16 (:213-233) on the decode path.
 *
 *  FAST   : one warp per row pair, 128-bit coalesced weight loads that bypass
 *           L1, x staged once per CTA in shared memory, fp32 FMA, warp-shuffle
 *           reduction.  Pure HBM streaming: 2 bytes per weight, read once.
 *  STRICT : one THREAD per row walking k ascending with round-to-nearest
 *           multiply then add (no FMA) - the reference's exact evaluation
 *           order - over [32 rows x 64 cols] tiles staged per warp with
 *           cp.async (padded rows, conflict-free 128-bit shared loads).
 */
#pragma once
#include "qtc_device.cuh"

struct GemvArgs {
    const uint16_t *w[3];
    float *out[3];
    int rows[3];
    int nseg, total_rows, cols, epi;
    const float *x;
    const float *aux;     /* bias[r] (EPI_BIAS) or residual[r] (EPI_RESID), indexed like out of segment 0 */
};

__device__ __forceinline__ void gemv_locate(const GemvArgs &a, int r, int &seg, int &lr) {
    seg = 0; lr = r;
    if (a.nseg > 1 && lr >= a.rows[0]) { lr -= a.rows[0]; seg = 1; if (a.nseg > 2 && lr >= a.rows[1]) { lr -= a.rows[1]; seg = 2; } }
}

__device__ __forceinline__ float dot8(const uint4 w, const float4 x0, const float4 x1, float acc) {
    acc = fmaf(bf16_lo(w.x), x0.x, acc); acc = fmaf(bf16_hi(w.x), x0.y, acc);
    acc = fmaf(bf16_lo(w.y), x0.z, acc); acc = fmaf(bf16_hi(w.y), x0.w, acc);
    acc = fmaf(bf16_lo(w.z), x1.x, acc); acc = fmaf(bf16_hi(w.z), x1.y, acc);
    acc = fmaf(bf16_lo(w.w), x1.z, acc); acc = fmaf(bf16_hi(w.w), x1.w, acc);
    return acc;
}

#define GEMV_FAST_THREADS 256
#define GEMV_FAST_ROWS 16      /* rows per CTA: 8 warps x 2 rows */

/* dynamic smem: cols floats */
__global__ void __launch_bounds__(GEMV_FAST_THREADS) k_gemv_fast(const GemvArgs a) {
    extern __shared__ __align__(16) float xs[];
    const int cols = a.cols;
    for (int i = threadIdx.x * 4; i < cols; i += GEMV_FAST_THREADS * 4)
        *reinterpret_cast<float4 *>(xs + i) = *reinterpret_cast<const float4 *>(a.x + i);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r0 = blockIdx.x * GEMV_FAST_ROWS + warp * 2;
    if (r0 >= a.total_rows) return;
    const bool two = r0 + 1 < a.total_rows;
    int s0, l0, s1, l1;
    gemv_locate(a, r0, s0, l0);
    gemv_locate(a, two ? r0 + 1 : r0, s1, l1);
    const uint4 *w0 = reinterpret_cast<const uint4 *>(a.w[s0] + (size_t)l0 * cols);
    const uint4 *w1 = reinterpret_cast<const uint4 *>(a.w[s1] + (size_t)l1 * cols);
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    const int n8 = cols >> 3;
    float acc0 = 0.0f, acc1 = 0.0f;
    int i = lane;
    for (; i + 96 < n8; i += 128) {            /* 8 independent 16-byte loads in flight per lane */
        uint4 a0 = ldg_stream(w0 + i), a1 = ldg_stream(w0 + i + 32), a2 = ldg_stream(w0 + i + 64), a3 = ldg_stream(w0 + i + 96);
        uint4 b0 = ldg_stream(w1 + i), b1 = ldg_stream(w1 + i + 32), b2 = ldg_stream(w1 + i + 64), b3 = ldg_stream(w1 + i + 96);
        acc0 = dot8(a0, x4[2 * i], x4[2 * i + 1], acc0);
        acc1 = dot8(b0, x4[2 * i], x4[2 * i + 1], acc1);
        acc0 = dot8(a1, x4[2 * (i + 32)], x4[2 * (i + 32) + 1], acc0);
        acc1 = dot8(b1, x4[2 * (i + 32)], x4[2 * (i + 32) + 1], acc1);
        acc0 = dot8(a2, x4[2 * (i + 64)], x4[2 * (i + 64) + 1], acc0);
        acc1 = dot8(b2, x4[2 * (i + 64)], x4[2 * (i + 64) + 1], acc1);
        acc0 = dot8(a3, x4[2 * (i + 96)], x4[2 * (i + 96) + 1], acc0);
        acc1 = dot8(b3, x4[2 * (i + 96)], x4[2 * (i + 96) + 1], acc1);
    }
    for (; i < n8; i += 32) {
        uint4 a0 = ldg_stream(w0 + i), b0 = ldg_stream(w1 + i);
        acc0 = dot8(a0, x4[2 * i], x4[2 * i + 1], acc0);
        acc1 = dot8(b0, x4[2 * i], x4[2 * i + 1], acc1);
    }
    acc0 = warp_sum(acc0);
    acc1 = warp_sum(acc1);
    if (lane == 0) {
        float v = acc0;
        if (a.epi != 0) v += a.aux[r0];
        a.out[s0][l0] = v;
        if (two) {
            v = acc1;
            if (a.epi != 0) v += a.aux[r0 + 1];
            a.out[s1][l1] = v;
        }
    }
}

/* gate/up rows share x; out[r] = silu(gate_r.x) * (up_r.x) */
__global__ void __launch_bounds__(GEMV_FAST_THREADS) k_gemv_swiglu_fast(float *out, const uint16_t *gate, const uint16_t *up,
                                                                         const float *x, int inter, int cols) {
    extern __shared__ __align__(16) float xs[];
    for (int i = threadIdx.x * 4; i < cols; i += GEMV_FAST_THREADS * 4)
        *reinterpret_cast<float4 *>(xs + i) = *reinterpret_cast<const float4 *>(x + i);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    const int n8 = cols >> 3;
    for (int r = blockIdx.x * 8 + warp; r < inter; r += gridDim.x * 8) {
        const uint4 *w0 = reinterpret_cast<const uint4 *>(gate + (size_t)r * cols);
        const uint4 *w1 = reinterpret_cast<const uint4 *>(up + (size_t)r * cols);
        float acc0 = 0.0f, acc1 = 0.0f;
        int i = lane;
        for (; i + 96 < n8; i += 128) {
            uint4 a0 = ldg_stream(w0 + i), a1 = ldg_stream(w0 + i + 32), a2 = ldg_stream(w0 + i + 64), a3 = ldg_stream(w0 + i + 96);
            uint4 b0 = ldg_stream(w1 + i), b1 = ldg_stream(w1 + i + 32), b2 = ldg_stream(w1 + i + 64), b3 = ldg_stream(w1 + i + 96);
            acc0 = dot8(a0, x4[2 * i], x4[2 * i + 1], acc0);
            acc1 = dot8(b0, x4[2 * i], x4[2 * i + 1], acc1);
            acc0 = dot8(a1, x4[2 * (i + 32)], x4[2 * (i + 32) + 1], acc0);
            acc1 = dot8(b1, x4[2 * (i + 32)], x4[2 * (i + 32) + 1], acc1);
            acc0 = dot8(a2, x4[2 * (i + 64)], x4[2 * (i + 64) + 1], acc0);
            acc1 = dot8(b2, x4[2 * (i + 64)], x4[2 * (i + 64) + 1], acc1);
            acc0 = dot8(a3, x4[2 * (i + 96)], x4[2 * (i + 96) + 1], acc0);
            acc1 = dot8(b3, x4[2 * (i + 96)], x4[2 * (i + 96) + 1], acc1);
        }
        for (; i < n8; i += 32) {
            uint4 a0 = ldg_stream(w0 + i), b0 = ldg_stream(w1 + i);
            acc0 = dot8(a0, x4[2 * i], x4[2 * i + 1], acc0);
            acc1 = dot8(b0, x4[2 * i], x4[2 * i + 1], acc1);
        }
        acc0 = warp_sum(acc0);
        acc1 = warp_sum(acc1);
        if (lane == 0) out[r] = swiglu_act<QTC_FAST>(acc0, acc1);
    }
}

/* ---- STRICT ------------------------------------------------------------------ */
#define STRICT_TILE_K 64
#define STRICT_ROW_BYTES 144                       /* 64 bf16 + 16 B pad: conflict-free LDS.128 per row */
#define STRICT_STAGES 3
#define STRICT_WARP_SMEM (STRICT_STAGES * 32 * STRICT_ROW_BYTES)
#define STRICT_WARPS 4

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

/* Rows row0..row0+31 of W (clamped to rows-1), all `cols` columns (cols % 64 == 0);
 * returns, in lane l, the reference-order sum for row row0+l.  xs: x in shared memory. */
__device__ __forceinline__ float strict_rows32(const uint16_t *W, int row0, int rows, int cols, const float *xs,
                                               unsigned char *tile) {
    const int lane = threadIdx.x & 31;
    const int ntile = cols / STRICT_TILE_K;
    /* piece j = lane + 32*i covers row j/8, 16-byte segment j%8 of a tile */
    auto issue = [&](int t, int stage) {
        unsigned char *dst = tile + stage * 32 * STRICT_ROW_BYTES;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int j = lane + 32 * i, rr = j >> 3, sg = j & 7;
            int r = row0 + rr; if (r >= rows) r = rows - 1;
            cp_async16(dst + rr * STRICT_ROW_BYTES + sg * 16, W + (size_t)r * cols + t * STRICT_TILE_K + sg * 8);
        }
    };
#pragma unroll
    for (int s = 0; s < STRICT_STAGES - 1; s++) { if (s < ntile) issue(s, s); cp_async_commit(); }
    float acc = 0.0f;
    for (int t = 0; t < ntile; t++) {
        if (t + STRICT_STAGES - 1 < ntile) issue(t + STRICT_STAGES - 1, (t + STRICT_STAGES - 1) % STRICT_STAGES);
        cp_async_commit();
        cp_async_wait<STRICT_STAGES - 1>();
        __syncwarp();
        const uint4 *row = reinterpret_cast<const uint4 *>(tile + (t % STRICT_STAGES) * 32 * STRICT_ROW_BYTES + lane * STRICT_ROW_BYTES);
        const float4 *x4 = reinterpret_cast<const float4 *>(xs + t * STRICT_TILE_K);
#pragma unroll
        for (int p = 0; p < 8; p++) {
            const uint4 w = row[p];
            const float4 x0 = x4[2 * p], x1 = x4[2 * p + 1];
            acc = __fadd_rn(acc, __fmul_rn(bf16_lo(w.x), x0.x)); acc = __fadd_rn(acc, __fmul_rn(bf16_hi(w.x), x0.y));
            acc = __fadd_rn(acc, __fmul_rn(bf16_lo(w.y), x0.z)); acc = __fadd_rn(acc, __fmul_rn(bf16_hi(w.y), x0.w));
            acc = __fadd_rn(acc, __fmul_rn(bf16_lo(w.z), x1.x)); acc = __fadd_rn(acc, __fmul_rn(bf16_hi(w.z), x1.y));
            acc = __fadd_rn(acc, __fmul_rn(bf16_lo(w.w), x1.z)); acc = __fadd_rn(acc, __fmul_rn(bf16_hi(w.w), x1.w));
        }
        __syncwarp();
    }
    return acc;
}

/* dynamic smem: cols floats, then STRICT_WARPS * STRICT_WARP_SMEM bytes of tiles */
__global__ void __launch_bounds__(STRICT_WARPS * 32) k_gemv_strict(const GemvArgs a) {
    extern __shared__ __align__(16) float xs[];
    const int cols = a.cols;
    for (int i = threadIdx.x * 4; i < cols; i += blockDim.x * 4)
        *reinterpret_cast<float4 *>(xs + i) = *reinterpret_cast<const float4 *>(a.x + i);
    __syncthreads();
    unsigned char *tiles = reinterpret_cast<unsigned char *>(xs + cols);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    /* 32-row groups never straddle a segment: group index -> (segment, first row) */
    int grp = blockIdx.x * STRICT_WARPS + warp, seg = 0;
    while (seg < a.nseg && grp >= (a.rows[seg] + 31) / 32) { grp -= (a.rows[seg] + 31) / 32; seg++; }
    if (seg >= a.nseg) return;
    const int row0 = grp * 32;
    const float acc = strict_rows32(a.w[seg], row0, a.rows[seg], cols, xs, tiles + warp * STRICT_WARP_SMEM);
    const int r = row0 + lane;
    if (r < a.rows[seg]) {
        float v = acc;
        if (a.epi != 0) v = __fadd_rn(a.aux[r], v);     /* bias: f[i] += b[i]; residual: x[i] += proj[i] */
        a.out[seg][r] = v;
    }
}

/* warps 0,1: gate rows; warps 2,3: the same 64 up rows; combined through shared memory */
__global__ void __launch_bounds__(STRICT_WARPS * 32) k_gemv_swiglu_strict(float *out, const uint16_t *gate, const uint16_t *up,
                                                                           const float *x, int inter, int cols) {
    extern __shared__ __align__(16) float xs[];
    __shared__ float gbuf[64];
    for (int i = threadIdx.x * 4; i < cols; i += blockDim.x * 4)
        *reinterpret_cast<float4 *>(xs + i) = *reinterpret_cast<const float4 *>(x + i);
    __syncthreads();
    unsigned char *tiles = reinterpret_cast<unsigned char *>(xs + cols);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * 64 + (warp & 1) * 32;
    float acc = 0.0f;
    if (row0 < inter) acc = strict_rows32(warp < 2 ? gate : up, row0, inter, cols, xs, tiles + warp * STRICT_WARP_SMEM);
    if (warp < 2) gbuf[(warp & 1) * 32 + lane] = acc;
    __syncthreads();
    if (warp >= 2 && row0 + lane < inter) out[row0 + lane] = swiglu_act<QTC_STRICT>(gbuf[(warp & 1) * 32 + lane], acc);
}

/* any cols: thread per row straight from global memory (tiny shapes only) */
__global__ void k_gemv_strict_generic(const GemvArgs a) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.total_rows) return;
    int seg, lr;
    gemv_locate(a, r, seg, lr);
    const uint16_t *w = a.w[seg] + (size_t)lr * a.cols;
    float acc = 0.0f;
    for (int c = 0; c < a.cols; c++) acc = __fadd_rn(acc, __fmul_rn(bf16_to_f32(w[c]), a.x[c]));
    if (a.epi != 0) acc = __fadd_rn(a.aux[r], acc);
    a.out[seg][lr] = acc;
}
