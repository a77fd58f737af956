/*
 * qtc_device.cuh - device-side building blocks shared by every kernel of the
 * B200 backend: bf16 unpacking, warp/block reductions, the two arithmetic
 * policies, and a bit-faithful expf.
 *
 * Arithmetic policies
 *   FAST   : fp32 with FMA contraction and tree reductions (weight-streaming speed).
 *   STRICT : the reference's evaluation order, operation by operation:
 *            round-to-nearest multiply, then add, k ascending, no FMA
 *            (gcc -std=c11 => -ffp-contract=off).  Used to
 *            prove the kernels differ from the C path only by summation order.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define QTC_WARP 32

enum { QTC_FAST = 0, QTC_STRICT = 1 };

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ float bf16_to_f32(uint16_t h) { return __uint_as_float(((uint32_t)h) << 16); }

/* streaming 128-bit load that does not pollute L1 (weights are read once per step) */
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

/* ---- paged KV cache -----------------------------------------------------------
 * The talker's K and V rows live in one pool per layer, allocated once at upload: pages of 64 positions x (kv_heads*head_dim)
 * fp32, handed out to utterances by a host-side free list; table[i] = pool page that holds positions [64 i, 64 i + 64) of
 * the utterance.  Row t of a layer's cache is at layer_base + kv_off(table, t, kvd) floats.  table == NULL: a contiguous
 * cache (the sub-talker's 18 positions, reset every frame).  The reference grows one contiguous fp32 cache by realloc
 * (c/qwen_tts_talker.c:265-271, 489-496). */
#define QTC_KV_PAGE 64
__device__ __forceinline__ size_t kv_off(const int *table, int t, int kvd) {
    return table ? ((size_t)table[t >> 6] * QTC_KV_PAGE + (size_t)(t & (QTC_KV_PAGE - 1))) * (size_t)kvd : (size_t)t * (size_t)kvd;
}

/* programmatic dependent launch (batched path): every kernel of a captured frame starts with pdl_enter() - the next kernel may
 * be scheduled as soon as all CTAs of this one are resident (its launch latency and prologue overlap this kernel's run), and
 * nothing of this kernel's own work happens before the previous kernel has completed and flushed.  Without the launch
 * attribute both instructions are no-ops. */
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() { pdl_trigger(); pdl_wait(); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

/* block-wide sum; `red` is >= 32 floats of shared memory; result broadcast to all threads */
__device__ __forceinline__ float block_sum(float v, float *red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = lane < nw ? red[lane] : 0.0f;
    t = warp_sum(t);
    return t;
}
__device__ __forceinline__ float block_max(float v, float *red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = lane < nw ? red[lane] : -3.402823466e+38f;
    t = warp_max(t);
    return t;
}

/* ---- expf with glibc's algorithm -------------------------------------------
 * glibc >= 2.28 computes expf in double precision from a 32-entry table
 * (sysdeps/ieee754/flt-32/e_expf.c, the ARM optimized-routines algorithm):
 * exp(x) = 2^(k/32) * 2^(r/32), polynomial of degree 3 in r, one final
 * rounding to float.  Re-stating it in fp64 on the device reproduces the
 * oracle's expf bit for bit (checked against the host libm in tests); CUDA's
 * own expf is only within 2 ulp.  The table is 2^(i/32) with i<<47 subtracted.
 */
__device__ __constant__ uint64_t QTC_EXP2F_TAB[32] = {
    0x3ff0000000000000ULL, 0x3fefd9b0d3158574ULL, 0x3fefb5586cf9890fULL, 0x3fef9301d0125b51ULL,
    0x3fef72b83c7d517bULL, 0x3fef54873168b9aaULL, 0x3fef387a6e756238ULL, 0x3fef1e9df51fdee1ULL,
    0x3fef06fe0a31b715ULL, 0x3feef1a7373aa9cbULL, 0x3feedea64c123422ULL, 0x3feece086061892dULL,
    0x3feebfdad5362a27ULL, 0x3feeb42b569d4f82ULL, 0x3feeab07dd485429ULL, 0x3feea47eb03a5585ULL,
    0x3feea09e667f3bcdULL, 0x3fee9f75e8ec5f74ULL, 0x3feea11473eb0187ULL, 0x3feea589994cce13ULL,
    0x3feeace5422aa0dbULL, 0x3feeb737b0cdc5e5ULL, 0x3feec49182a3f090ULL, 0x3feed503b23e255dULL,
    0x3feee89f995ad3adULL, 0x3feeff76f2fb5e47ULL, 0x3fef199bdd85529cULL, 0x3fef3720dcef9069ULL,
    0x3fef5818dcfba487ULL, 0x3fef7c97337b9b5fULL, 0x3fefa4afa2a490daULL, 0x3fefd0765b6e4540ULL};

__device__ __forceinline__ float expf_libm(float x) {
    const uint32_t ax = __float_as_uint(x) & 0x7fffffffu;
    if (ax >= 0x42b00000u) {                       /* |x| >= 88 or non-finite */
        if (__float_as_uint(x) == 0xff800000u) return 0.0f;
        if (ax >= 0x7f800000u) return x + x;
        if (x > 88.72283172607421875f) return __uint_as_float(0x7f800000u);      /* 0x1.62e42ep6f */
        if (x < -103.972076416015625f) return 0.0f;                              /* -0x1.9fe368p6f */
    }
    const double xd = (double)x;
    const double z0 = __dmul_rn(46.16624130844683, xd);                          /* 32/ln2 */
    double kd = __dadd_rn(z0, 6755399441055744.0);                               /* 0x1.8p52 */
    const uint64_t ki = (uint64_t)__double_as_longlong(kd);
    kd = __dadd_rn(kd, -6755399441055744.0);
    const double r = __dadd_rn(z0, -kd);
    const uint64_t t = QTC_EXP2F_TAB[ki & 31] + (ki << 47);
    const double s = __longlong_as_double((long long)t);
    /* glibc's x86-64 build of this file selects an FMA variant on every CPU of
     * the last decade; the fused forms below follow that variant. */
    const double zz = __fma_rn(0x1.c6af84b912394p-20, r, 0x1.ebfce50fac4f3p-13);
    const double r2 = __dmul_rn(r, r);
    double y = __fma_rn(0x1.62e42ff0c52d6p-6, r, 1.0);
    y = __fma_rn(zz, r2, y);
    y = __dmul_rn(y, s);
    return (float)y;
}

/* xorshift32 on the storage of a float: rand_uniform, c/qwen_tts_kernels.c:384-393 */
__device__ __forceinline__ float rng_next(uint32_t *state) {
    uint32_t s = *state;
    s ^= s << 13; s ^= s >> 17; s ^= s << 5;
    *state = s;
    return __fdiv_rn((float)(s & 0x7fffffffu), (float)0x7fffffff);
}

/* SiLU(g)*u exactly as c/qwen_tts_kernels.c:229-232: (g / (1 + expf(-g))) * u */
template <int ARITH>
__device__ __forceinline__ float swiglu_act(float g, float u) {
    if (ARITH == QTC_STRICT) return __fmul_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf_libm(-g))), u);
    return (g / (1.0f + __expf(-g))) * u;
}
