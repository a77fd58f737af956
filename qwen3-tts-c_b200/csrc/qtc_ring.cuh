/*
 * qtc_ring.cuh - weight streaming + tensor-core GEMV for the decode-step kernel k_lm_ring (mode 2)
 * (talker_attention_single / ST_FORWARD, c/qwen_tts_talker.c:119-248, 638-691; kernel_matvec_bf16,
 * c/qwen_tts_kernels.c:95-149).
 *
 * Streaming.  The bf16 weights a CTA reads during a whole launch are known before the launch starts: its
 * contiguous row range of every matrix, in phase order.  A producer warp walks that list and keeps up to
 * nst x 16 KB of TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx) in flight or landed per SM; it
 * never waits for activations, only for a free stage.  The sixteen consumer warps wait for activations
 * (grid barriers) but not for HBM: when a phase may start, its first stages are already in shared memory,
 * so HBM streams through grid barriers, RMSNorms, the attention phase and the sampler.
 *
 * Stage = 8 weight rows x one 1024-column chunk (one 2 KB bulk copy per row, rows skewed by 16 bytes);
 * gate/up stages hold 4 gate rows + the same 4 up rows.  A row block's chunks are consecutive stages.
 *
 * Arithmetic.  Batch-1 GEMV on CUDA cores costs > 2 issue slots per weight (unpack + FMA) and a shuffle
 * tree per row; with 16 warps sharing four schedulers that, not HBM, bounded the step.  Here the 8 rows of
 * a stage are the N side of mma.sync.m16n8k16 (bf16 x bf16 -> fp32) and the activation vector is the M
 * side, split EXACTLY into three bf16 pieces (x = hi + mid + lo, 8 + 8 + 8 mantissa bits) that occupy rows
 * 0..2 of the A tile: every bf16 x bf16 product is exact in fp32, so the result differs from the fp32 FMA
 * path only by summation order - the same class of difference "fast" arithmetic already has - while one
 * ldmatrix + two MMAs replace ~40 issue slots per 256 weights.  Warp w owns columns [64w, 64w + 64) of
 * every chunk; per row block it leaves 8 partial sums in shared memory and one thread per row folds the 16
 * warps in a fixed order (deterministic), applies the epilogue (residual / bias / SwiGLU) and stores.
 */
#pragma once
#include "qtc_fused_kernel.cuh"

#define RING_KC 1024                              /* weight columns of one row per stage */
#define RING_ROW_STRIDE (RING_KC * 2 + 16)        /* bytes between stage rows: the 16-byte skew keeps ldmatrix conflict-free */
#define RING_STAGE_BYTES (8 * RING_ROW_STRIDE)    /* 8 rows x 1024 bf16 (+ skew) = 16512 bytes */
#define RING_MAX_STAGES 13
#define RING_THREADS (FUSED_THREADS + 32)     /* 16 consumer warps + the producer warp */

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct Ring {
    unsigned char *buf;        /* nst * RING_STAGE_BYTES, 128-byte aligned */
    uint64_t *full, *empty;    /* [nst] each */
    float *part;               /* [local rows][16] partial sums of the phase in flight */
    int nst;
};
/* position in the ring; `phase` flips every time the stage index wraps (mbarrier parity) */
struct RingPos { int stage; unsigned phase; };
__device__ __forceinline__ void ring_advance(RingPos &rp, int nst) {
    if (++rp.stage == nst) { rp.stage = 0; rp.phase ^= 1u; }
}


/* can a matrix with rows of `cols` bf16 go through the ring? */
__device__ __host__ __forceinline__ bool ring_shape_ok(int cols) { return cols > 0 && cols % 64 == 0; }

/* ---- geometry of one weight phase for one CTA; identical in producer and consumers -----------------------
 * local rows (units for gate/up) 0..n-1 across segment boundaries; row block b = local rows [8b, 8b+8)
 * (units [4b, 4b+4)); stages in order: for each row block, for each 1024-column chunk. */
struct PhaseGeom {
    int cols, swiglu;
    int r0, n;        /* first phase-global row (unit) of this CTA, count */
    int nrb, C;       /* row blocks, column chunks */
};
/* 32-bit row split (rows * ncta < 2^31 is checked on the host): same result as row_range */
__device__ __forceinline__ void row_range32(int rows, int cta, int ncta, int &r0, int &r1) {
    r0 = (int)(((unsigned)rows * (unsigned)cta) / (unsigned)ncta);
    r1 = (int)(((unsigned)rows * (unsigned)(cta + 1)) / (unsigned)ncta);
}
__device__ __forceinline__ PhaseGeom phase_geom(const Seg3 &s, int cols, int swiglu, int cta, int ncta) {
    PhaseGeom g;
    g.cols = cols; g.swiglu = swiglu;
    int r1;
    const int total = swiglu ? s.rows[0] : s.rows[0] + (s.n > 1 ? s.rows[1] : 0) + (s.n > 2 ? s.rows[2] : 0);
    row_range32(total, cta, ncta, g.r0, r1);
    g.n = r1 - g.r0;
    g.nrb = swiglu ? (g.n + 3) >> 2 : (g.n + 7) >> 3;
    g.C = (cols + RING_KC - 1) / RING_KC;
    return g;
}

/* producer warp: every stage of one phase, in consumption order (chunk-major: all row blocks of column chunk 0,
 * then chunk 1, ...); lane r copies stage row r */
__device__ __forceinline__ void ring_produce(const Ring &rg, RingPos &rp, const Seg3 &s, int cols, int swiglu, int cta, int ncta, long long *pf = NULL) {
    const PhaseGeom g = phase_geom(s, cols, swiglu, cta, ncta);
    const int lane = threadIdx.x & 31;
    for (int c = 0; c < g.C; c++) {
        const int kc = cols - c * RING_KC < RING_KC ? cols - c * RING_KC : RING_KC;
        for (int b = 0; b < g.nrb; b++) {
            /* source of my row of this block (lanes 0..7) */
            const uint16_t *src = NULL;
            int cnt;
            if (swiglu) {
                const int u0 = b * 4, nu = g.n - u0 < 4 ? g.n - u0 : 4, u = lane & 3;
                cnt = 2 * nu;
                if (lane < 8 && u < nu) src = (lane < 4 ? s.w[0] : s.w[1]) + (size_t)(g.r0 + u0 + u) * cols;
            } else {
                const int l0 = b * 8;
                cnt = g.n - l0 < 8 ? g.n - l0 : 8;
                if (lane < cnt) {
                    int q = g.r0 + l0 + lane;
                    src = s.w[0];
                    if (s.n > 1 && q >= s.rows[0]) {
                        q -= s.rows[0]; src = s.w[1];
                        if (s.n > 2 && q >= s.rows[1]) { q -= s.rows[1]; src = s.w[2]; }
                    }
                    src += (size_t)q * cols;
                }
            }
            if (pf && lane == 0 && b == 0 && c == 0) pf[27] = clock64();
            mbar_wait(&rg.empty[rp.stage], rp.phase ^ 1u);
            if (pf && lane == 0 && b == 0 && c == 0) pf[28] = clock64();
            uint64_t *full = &rg.full[rp.stage];
            if (lane == 0) mbar_expect_tx(full, (unsigned)(cnt * kc * 2));
            __syncwarp();
            if (src) bulk_g2s(rg.buf + (size_t)rp.stage * RING_STAGE_BYTES + (size_t)lane * RING_ROW_STRIDE, src + (size_t)c * RING_KC, (unsigned)(kc * 2), full);
            ring_advance(rp, rg.nst);
        }
    }
}

/* ---- activation vector as three bf16 pieces: xa[p][cols] (piece stride cols + 8 elements) ---------------- */
__device__ __forceinline__ uint32_t bf16_rn_bits(float v) {      /* round-to-nearest-even, finite inputs */
    const uint32_t u = __float_as_uint(v);
    return (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
}
__device__ __forceinline__ void split3(float v, uint32_t &h, uint32_t &m, uint32_t &l) {
    h = bf16_rn_bits(v);
    const float r1 = v - __uint_as_float(h << 16);               /* exact */
    m = bf16_rn_bits(r1);
    const float r2 = r1 - __uint_as_float(m << 16);              /* exact */
    l = bf16_rn_bits(r2);
}
/* src: shared memory (XG = false) or a vector produced by other CTAs (XG = true, read through L2); cols even */
template <bool XG>
__device__ __noinline__ void xa_build(uint16_t *xa, const float *src, int cols) {
    const int ps = cols + 8;
    for (int i = threadIdx.x * 2; i < cols; i += FUSED_THREADS * 2) {
        const float2 v = XG ? __ldcg(reinterpret_cast<const float2 *>(src + i)) : *reinterpret_cast<const float2 *>(src + i);
        uint32_t h0, m0, l0, h1, m1, l1;
        split3(v.x, h0, m0, l0); split3(v.y, h1, m1, l1);
        *reinterpret_cast<uint32_t *>(xa + i) = h0 | (h1 << 16);
        *reinterpret_cast<uint32_t *>(xa + ps + i) = m0 | (m1 << 16);
        *reinterpret_cast<uint32_t *>(xa + 2 * ps + i) = l0 | (l1 << 16);
    }
    compute_sync();
}

__device__ __forceinline__ void ldsm_x4(uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3, uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float &c0, float &c1, float &c2, float &c3, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c0), "+f"(c1), "+f"(c2), "+f"(c3) : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ring_advance_n(RingPos &rp, int k, int nst) {
    int st = rp.stage + k;
    while (st >= nst) { st -= nst; rp.phase ^= 1u; }
    rp.stage = st;
}

#define RING_GROUPS 4                 /* warp groups; group j takes stages j, j+4, ... of a phase */
#define RING_GW (FUSED_WARPS / RING_GROUPS)   /* warps per group: each owns 256 columns of the chunk */

/* consumers (the 512 compute threads): this CTA's share of one weight phase -> partials in rg.part
 * ([stage of the phase][warp of the group][8 rows]).  xa: the split input vector in shared memory.
 * Every stage is independent (MMAs from zero, fold of the three pieces, 2 floats per lane < 4).  The four warp
 * groups work on four consecutive stages at the same time, so the wait / ldmatrix / MMA / shuffle latency chain
 * of one stage overlaps with three others without unrolling. */
__device__ __noinline__ RingPos ring_consume(const Ring rg, RingPos rp, const PhaseGeom g, const uint16_t *xa) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, grp = warp / RING_GW, wq = warp % RING_GW;
    const int gi = lane >> 2, t = lane & 3, ps = g.cols + 8;
    const int k0 = wq * 256;                                    /* my columns inside every chunk */
    /* A fragments: piece gi (rows 0..2 of the tile), columns k + 2t, 2t+1 and + 8 */
    const uint16_t *xrow = xa + (gi < 3 ? gi : 0) * ps + k0 + 2 * t;
    /* B fragments: ldmatrix.x4 = four 8x8 tiles (8 rows x 8 columns each); lane -> (tile lane/8, row lane%8) */
    const uint32_t boff = (uint32_t)(lane & 7) * RING_ROW_STRIDE + (uint32_t)(k0 + (lane >> 3) * 8) * 2u;
    const uint32_t buf0 = smem_u32(rg.buf);
    const int total = g.nrb * g.C;
    RingPos q = rp;
    ring_advance_n(q, grp, rg.nst);
    int c = grp / g.nrb, b = grp - c * g.nrb;                   /* (chunk, block) of my first stage; stages are chunk-major */
    float *pdst = rg.part + (grp * RING_GW + wq) * 8 + 2 * t;
    for (int i = grp; i < total; i += RING_GROUPS) {
        const int kc = g.cols - c * RING_KC < RING_KC ? g.cols - c * RING_KC : RING_KC;
        mbar_wait(&rg.full[q.stage], q.phase);
        float r0 = 0.0f, r1 = 0.0f;
        if (k0 < kc) {
            float acc[4][4];
#pragma unroll
            for (int j = 0; j < 4; j++) { acc[j][0] = 0.0f; acc[j][1] = 0.0f; acc[j][2] = 0.0f; acc[j][3] = 0.0f; }
            const uint32_t sb = buf0 + (uint32_t)q.stage * RING_STAGE_BYTES + boff;
            const uint16_t *xp = xrow + c * RING_KC;
#pragma unroll
            for (int s4 = 0; s4 < 4; s4++) {
                if (k0 + s4 * 64 < kc) {
                    uint32_t w[8], a[8];
                    ldsm_x4(w[0], w[1], w[2], w[3], sb + s4 * 128);
                    ldsm_x4(w[4], w[5], w[6], w[7], sb + s4 * 128 + 64);
#pragma unroll
                    for (int j = 0; j < 8; j++) a[j] = gi < 3 ? *reinterpret_cast<const uint32_t *>(xp + s4 * 64 + j * 8) : 0u;
#pragma unroll
                    for (int j = 0; j < 4; j++) mma_bf16(acc[j][0], acc[j][1], acc[j][2], acc[j][3], a[2 * j], a[2 * j + 1], w[2 * j], w[2 * j + 1]);
                }
            }
            r0 = (acc[0][0] + acc[1][0]) + (acc[2][0] + acc[3][0]);
            r1 = (acc[0][1] + acc[1][1]) + (acc[2][1] + acc[3][1]);
            /* fold the three pieces (lanes gi = 0,1,2 -> gi = 0); lanes 0..3 hold rows 2t, 2t+1 of the block */
            r0 += __shfl_down_sync(0xffffffffu, r0, 4) + __shfl_down_sync(0xffffffffu, r0, 8);
            r1 += __shfl_down_sync(0xffffffffu, r1, 4) + __shfl_down_sync(0xffffffffu, r1, 8);
        }
        if (lane < 4) { pdst[0] = r0; pdst[1] = r1; }
        __syncwarp();
        if (lane == 0) mbar_arrive(&rg.empty[q.stage]);
        ring_advance_n(q, RING_GROUPS, rg.nst);
        pdst += RING_GROUPS * RING_GW * 8;
        b += RING_GROUPS;
        while (b >= g.nrb) { b -= g.nrb; c++; }
    }
    ring_advance_n(rp, total, rg.nst);
    return rp;
}

/* fold the partials of the phase and store this CTA's outputs.  aux: residual or bias indexed by the phase-global
 * row (read through L2), or NULL.  swiglu phases write g[unit] = silu(gate.x) * (up.x).
 * stage (c, b) of the phase left part[((c*nrb + b)*4 + warp of the group)*8 + row]. */
__device__ __noinline__ void ring_finalize(const Ring rg, const PhaseGeom g, const Seg3 s, const float *aux) {
    compute_sync();
    const int cstride = g.nrb * RING_GW * 8;
    for (int lr = threadIdx.x; lr < g.n; lr += FUSED_THREADS) {
        if (g.swiglu) {
            const float *p = rg.part + (lr >> 2) * (RING_GW * 8) + (lr & 3);
            float gv = 0.0f, uv = 0.0f;
            for (int c = 0; c < g.C; c++, p += cstride)
#pragma unroll
                for (int k = 0; k < RING_GW; k++) { gv += p[k * 8]; uv += p[k * 8 + 4]; }
            s.out[0][g.r0 + lr] = swiglu_act<QTC_FAST>(gv, uv);
        } else {
            const float *p = rg.part + (lr >> 3) * (RING_GW * 8) + (lr & 7);
            float v = 0.0f;
            for (int c = 0; c < g.C; c++, p += cstride)
#pragma unroll
                for (int k = 0; k < RING_GW; k++) v += p[k * 8];
            const int gr = g.r0 + lr;
            int q = gr;
            float *dst = s.out[0];
            if (s.n > 1 && q >= s.rows[0]) {
                q -= s.rows[0]; dst = s.out[1];
                if (s.n > 2 && q >= s.rows[1]) { q -= s.rows[1]; dst = s.out[2]; }
            }
            dst[q] = aux ? __ldcg(aux + gr) + v : v;
        }
    }
}
