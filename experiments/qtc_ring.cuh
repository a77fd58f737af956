/*
 * qtc_ring.cuh - weight streaming for the fused decode-step kernel: a shared-memory ring filled by
 * TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx) that runs AHEAD of the data dependencies.
 *
 * The weights a CTA will read during a whole decoder-stack step are known before the step starts
 * (its contiguous row range of every matrix, in phase order).  One producer warp walks that list and
 * keeps RING_STAGES x 32 KB of bulk copies in flight per SM; it never waits for activations, only for a
 * free stage.  The sixteen consumer warps wait for activations (grid barriers) but never for HBM latency:
 * by the time a phase may start, its first stages have landed.  HBM therefore streams through grid
 * barriers, RMSNorms and the attention phase, and a phase is bounded by shared-memory bandwidth
 * (128 B/clk/SM) instead of by one DRAM round trip per 8 loads per lane.
 *
 * Chunk = up to R whole rows (R a power of two <= 16, R*row_bytes <= 32 KB) of ONE matrix segment, or for
 * gate/up R/2 gate rows + the same R/2 up rows.  A warp owns (row w/F, K-slice w%F), F = 16/R, keeps its
 * slice of x in registers for the whole phase, and reads only weights from the ring; slices are folded
 * in a fixed order (deterministic) by one thread per row.
 */
#pragma once
#include "qtc_fused.cuh"

#define RING_STAGES 4
#define RING_STAGE_BYTES 32768
#define RING_MAX_PPL 4              /* 16-byte weight pieces per lane per row slice */
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
    unsigned char *buf;        /* RING_STAGES * RING_STAGE_BYTES, 128-byte aligned */
    uint64_t *full, *empty;    /* [RING_STAGES] each */
    float *part;               /* [2][16][16] partial sums */
};

/* rows per chunk for a row of `cols` bf16: largest power of two <= 16 that fits a stage */
__device__ __host__ __forceinline__ int ring_rows(int cols) {
    int r = 16;
    while (r > 1 && (size_t)r * cols * 2 > RING_STAGE_BYTES) r >>= 1;
    return r;
}
/* can this shape go through the ring?  (host-side check mirrors the device layout rules) */
__device__ __host__ __forceinline__ bool ring_shape_ok(int cols) {
    if (cols % 8 || (size_t)cols * 2 > RING_STAGE_BYTES) return false;
    const int R = ring_rows(cols), F = 16 / R, np = cols / 8;
    return np % F == 0 && (np / F + 31) / 32 <= RING_MAX_PPL;
}

/* iterator over the chunks of one weight phase for one CTA; identical in producer and consumers */
struct PhaseIter {
    Seg3 s; int cols, swiglu, R;
    int k, cur, hi;            /* segment, next row (segment-local), end row */
    int r0, r1, base;          /* CTA range in phase-global rows; first global row of segment k */
};
struct Chunk { const uint16_t *src0, *src1; int rows; float *out; int seg_row; int grow; };

__device__ __forceinline__ void iter_begin(PhaseIter &it, const Seg3 &s, int cols, int swiglu, int cta, int ncta) {
    it.s = s; it.cols = cols; it.swiglu = swiglu; it.R = ring_rows(cols);
    if (swiglu) {             /* units u in [r0,r1): gate row u + up row u; segment 0 = gate, 1 = up */
        row_range(s.rows[0], cta, ncta, it.r0, it.r1);
        it.k = 0; it.base = 0; it.cur = it.r0; it.hi = it.r1;
        if (it.R > 1) it.R >>= 1;                  /* R/2 units per chunk -> R stage rows */
        return;
    }
    const int total = s.rows[0] + (s.n > 1 ? s.rows[1] : 0) + (s.n > 2 ? s.rows[2] : 0);
    row_range(total, cta, ncta, it.r0, it.r1);
    it.k = -1; it.base = 0; it.cur = 0; it.hi = 0;
}
__device__ __forceinline__ bool iter_next(PhaseIter &it, Chunk &c) {
    if (it.swiglu) {
        if (it.cur >= it.hi) return false;
        c.rows = it.hi - it.cur < it.R ? it.hi - it.cur : it.R;
        c.src0 = it.s.w[0] + (size_t)it.cur * it.cols; c.src1 = it.s.w[1] + (size_t)it.cur * it.cols;
        c.out = it.s.out[0]; c.seg_row = it.cur; c.grow = it.cur;
        it.cur += c.rows;
        return true;
    }
    while (it.cur >= it.hi) {                       /* advance to the next segment that intersects [r0,r1) */
        if (it.k >= 0) it.base += it.s.rows[it.k];
        it.k++;
        if (it.k >= it.s.n) return false;
        const int lo = it.r0 > it.base ? it.r0 - it.base : 0;
        const int hi = (it.r1 < it.base + it.s.rows[it.k] ? it.r1 : it.base + it.s.rows[it.k]) - it.base;
        it.cur = lo; it.hi = hi > lo ? hi : lo;
    }
    c.rows = it.hi - it.cur < it.R ? it.hi - it.cur : it.R;
    c.src0 = it.s.w[it.k] + (size_t)it.cur * it.cols; c.src1 = NULL;
    c.out = it.s.out[it.k]; c.seg_row = it.cur; c.grow = it.base + it.cur;
    it.cur += c.rows;
    return true;
}

/* producer: one elected lane issues every chunk of every phase, in consumption order */
__device__ __forceinline__ void ring_produce_phase(const Ring &rg, PhaseIter &it, unsigned &n) {
    Chunk c;
    while (iter_next(it, c)) {
        const unsigned stage = n % RING_STAGES, round = n / RING_STAGES;
        if (round > 0) mbar_wait(&rg.empty[stage], (round - 1) & 1);
        unsigned char *dst = rg.buf + (size_t)stage * RING_STAGE_BYTES;
        const unsigned bytes = (unsigned)c.rows * it.cols * 2u;
        if (it.swiglu) {
            mbar_expect_tx(&rg.full[stage], 2u * bytes);
            bulk_g2s(dst, c.src0, bytes, &rg.full[stage]);
            bulk_g2s(dst + (size_t)it.R * it.cols * 2, c.src1, bytes, &rg.full[stage]);
        } else {
            mbar_expect_tx(&rg.full[stage], bytes);
            bulk_g2s(dst, c.src0, bytes, &rg.full[stage]);
        }
        n++;
    }
}

/* consumers (all 512 compute threads): this CTA's share of one weight phase.
 * xs: the phase's input vector in shared memory; resid: optional residual read through L2 (plain phases) */
__device__ __forceinline__ void ring_consume_phase(const Ring &rg, PhaseIter &it, unsigned &n, const float *xs, const float *resid) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cols = it.cols, np = cols >> 3;
    const int Rs = it.swiglu ? it.R * 2 : it.R;           /* stage rows */
    const int F = 16 / Rs > 0 ? 16 / Rs : 1;
    const int row = warp / F, slice = warp - row * F, nps = np / F;
    /* my K-slice of x stays in registers for the whole phase */
    float4 xr[RING_MAX_PPL][2];
#pragma unroll
    for (int j = 0; j < RING_MAX_PPL; j++) {
        const int p = lane + 32 * j;
        if (p < nps) {
            const float4 *xp = reinterpret_cast<const float4 *>(xs + (size_t)(slice * nps + p) * 8);
            xr[j][0] = xp[0]; xr[j][1] = xp[1];
        } else { xr[j][0] = make_float4(0, 0, 0, 0); xr[j][1] = xr[j][0]; }
    }
    Chunk c;
    while (iter_next(it, c)) {
        const unsigned stage = n % RING_STAGES, round = n / RING_STAGES;
        mbar_wait(&rg.full[stage], round & 1);
        const unsigned char *st = rg.buf + (size_t)stage * RING_STAGE_BYTES;
        /* stage row -> chunk row: plain: row; swiglu: rows [0,R) gate, [R,2R) up */
        const int crow = it.swiglu ? (row < it.R ? row : row - it.R) : row;
        float acc = 0.0f;
        if (row < Rs && crow < c.rows) {
            const uint4 *wp = reinterpret_cast<const uint4 *>(st + (size_t)row * cols * 2) + slice * nps;
#pragma unroll
            for (int j = 0; j < RING_MAX_PPL; j++) {
                const int p = lane + 32 * j;
                if (p < nps) acc = dot8f(wp[p], xr[j][0], xr[j][1], acc);
            }
        }
        acc = warp_sum(acc);
        float *part = rg.part + (n & 1) * 256;
        if (lane == 0) part[row * 16 + slice] = acc;
        compute_sync();
        if (tid == 0) mbar_arrive(&rg.empty[stage]);        /* every warp has finished reading the stage */
        if (tid < c.rows) {
            if (it.swiglu) {
                float g = 0.0f, u = 0.0f;
                for (int f = 0; f < F; f++) { g += part[tid * 16 + f]; u += part[(tid + it.R) * 16 + f]; }
                c.out[c.seg_row + tid] = swiglu_act<QTC_FAST>(g, u);
            } else {
                float v = 0.0f;
                for (int f = 0; f < F; f++) v += part[tid * 16 + f];
                if (resid) v += __ldcg(resid + c.grow + tid);
                c.out[c.seg_row + tid] = v;
            }
        }
        n++;
    }
}
