/*
 * qtc_fused_main.cuh - k_lm_fused: one decoder-stack step as one persistent cooperative kernel,
 * 148 CTAs x 16 warps.  See qtc_fused.cuh (barriers, staging, row split) and qtc_fused_kernel.cuh (attention).
 *
 * Weights do not depend on activations, so HBM must not wait for the exchanges between phases.  Two mechanisms keep
 * it streaming through every grid barrier, RMSNorm and the weight-less attention phase:
 *
 *   landing zone   when a CTA finishes a weight phase it issues TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx)
 *                  of ITS rows of the NEXT phase into ~110-130 KB of shared memory: as many whole rows as fit (all of
 *                  them for QKV and O of the 1.7B talker and every phase but gate/up of the sub-talker), one bulk copy per
 *                  contiguous run of rows.  They land while the barrier spins and the next input vector is staged; the phase
 *                  then reads them with conflict-free 128-bit shared-memory loads.
 *   first trip     each warp's first row pair of the rows that do NOT fit the zone - four 128-bit loads per lane per row,
 *                  64 registers - is issued right after the CTA ARRIVES at the barrier and consumed after it.
 *
 * (A ring of TMA stages feeding tensor-core GEMVs and an L2 bulk prefetch two phases ahead were both measured slower
 * on B200 at batch 1: profiles/README.md.  The landing zone differs from that ring in keeping the fp32 FMA datapath
 * and the row-per-warp ownership, so results are bit-identical to the plain path: same products, same order.)
 *
 * dynamic shared memory: [ xs | attention scratch | KV stage 2 x 64 x hd | landing zone ]
 */
#pragma once
#include "qtc_fused_kernel.cuh"

#define FUSED_BLOCK FUSED_THREADS

#define PFJ 4                      /* 16-byte pieces per lane per row held across the barrier */
struct WTrip { uint4 a[PFJ], b[PFJ]; };

/* ---- landing zone ------------------------------------------------------------------------------------------- */
struct Lz {
    unsigned char *buf;            /* shared memory, 128-byte aligned */
    uint32_t bar;                  /* one mbarrier (shared-window address): each prefetch generation is one phase of it */
    int bytes;
    unsigned parity;               /* parity of the generation the next wait belongs to (same in every thread) */
};
__device__ __forceinline__ uint32_t lz_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void lz_bar_init(uint64_t *bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(lz_u32(bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void lz_expect(uint32_t bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void lz_wait(Lz &lz) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "LZW_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra LZD_%=;\n"
        "bra LZW_%=;\n"
        "LZD_%=:\n"
        "}\n" ::"r"(lz.bar), "r"(lz.parity) : "memory");
    lz.parity ^= 1u;
}
__device__ __forceinline__ void lz_copy(void *dst, const void *src, unsigned bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(lz_u32(dst)), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
/* whole rows of `cols` bf16 that fit the zone (row-wise: a CTA's rows of one matrix are contiguous in HBM, so the zone is
 * filled by one bulk copy per matrix segment - per-row copies cost ~50-100 ns EACH of TMA issue time on the arrival path) */
__device__ __forceinline__ int lz_rows(const Lz &lz, int nrows, int cols) {
    if (nrows <= 0 || lz.bytes <= 0) return 0;
    const int fit = lz.bytes / (cols * 2);
    return fit < nrows ? fit : nrows;
}

/* rows (ra, rb) of `cols` bf16: lane's pieces lane+32j, j < PFJ */
__device__ __forceinline__ void trip_issue(WTrip &t, const uint16_t *ra, const uint16_t *rb, int cols) {
    const int lane = threadIdx.x & 31, n8 = cols >> 3;
    const uint4 *w0 = reinterpret_cast<const uint4 *>(ra), *w1 = reinterpret_cast<const uint4 *>(rb);
#pragma unroll
    for (int j = 0; j < PFJ; j++) {
        const int p = lane + 32 * j;
        if (p < n8) { t.a[j] = ldg_stream(w0 + p); t.b[j] = ldg_stream(w1 + p); }
    }
}
/* finish the two dot products: the prefetched pieces, then (cols > 1024) the rest of the rows */
__device__ __forceinline__ void trip_finish(const WTrip &t, const uint16_t *ra, const uint16_t *rb, const float *xs, int cols, float &oa, float &ob) {
    const int lane = threadIdx.x & 31, n8 = cols >> 3;
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    float a = 0.0f, b = 0.0f;
#pragma unroll
    for (int j = 0; j < PFJ; j++) {
        const int p = lane + 32 * j;
        if (p < n8) { a = dot8f(t.a[j], x4[2 * p], x4[2 * p + 1], a); b = dot8f(t.b[j], x4[2 * p], x4[2 * p + 1], b); }
    }
    if (n8 > 32 * PFJ) {
        const uint4 *w0 = reinterpret_cast<const uint4 *>(ra), *w1 = reinterpret_cast<const uint4 *>(rb);
        int i = lane + 32 * PFJ;
        for (; i + 96 < n8; i += 128) {
            const uint4 a0 = ldg_stream(w0 + i), a1 = ldg_stream(w0 + i + 32), a2 = ldg_stream(w0 + i + 64), a3 = ldg_stream(w0 + i + 96);
            const uint4 b0 = ldg_stream(w1 + i), b1 = ldg_stream(w1 + i + 32), b2 = ldg_stream(w1 + i + 64), b3 = ldg_stream(w1 + i + 96);
            a = dot8f(a0, x4[2 * i], x4[2 * i + 1], a);           b = dot8f(b0, x4[2 * i], x4[2 * i + 1], b);
            a = dot8f(a1, x4[2 * i + 64], x4[2 * i + 65], a);     b = dot8f(b1, x4[2 * i + 64], x4[2 * i + 65], b);
            a = dot8f(a2, x4[2 * i + 128], x4[2 * i + 129], a);   b = dot8f(b2, x4[2 * i + 128], x4[2 * i + 129], b);
            a = dot8f(a3, x4[2 * i + 192], x4[2 * i + 193], a);   b = dot8f(b3, x4[2 * i + 192], x4[2 * i + 193], b);
        }
        for (; i < n8; i += 32) {
            const uint4 a0 = ldg_stream(w0 + i), b0 = ldg_stream(w1 + i);
            a = dot8f(a0, x4[2 * i], x4[2 * i + 1], a);
            b = dot8f(b0, x4[2 * i], x4[2 * i + 1], b);
        }
    }
    oa = warp_sum(a);
    ob = warp_sum(b);
}
/* the same two dot products for rows that sit in the landing zone (la, lb: shared memory) */
__device__ __forceinline__ void lz_dot2(const unsigned char *la, const unsigned char *lb, const float *xs, int cols, float &oa, float &ob) {
    const int lane = threadIdx.x & 31, n8 = cols >> 3;
    const uint4 *w0 = reinterpret_cast<const uint4 *>(la), *w1 = reinterpret_cast<const uint4 *>(lb);
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    float a = 0.0f, b = 0.0f;
    int i = lane;
    for (; i + 96 < n8; i += 128) {
        const uint4 a0 = w0[i], a1 = w0[i + 32], a2 = w0[i + 64], a3 = w0[i + 96];
        const uint4 b0 = w1[i], b1 = w1[i + 32], b2 = w1[i + 64], b3 = w1[i + 96];
        a = dot8f(a0, x4[2 * i], x4[2 * i + 1], a);           b = dot8f(b0, x4[2 * i], x4[2 * i + 1], b);
        a = dot8f(a1, x4[2 * i + 64], x4[2 * i + 65], a);     b = dot8f(b1, x4[2 * i + 64], x4[2 * i + 65], b);
        a = dot8f(a2, x4[2 * i + 128], x4[2 * i + 129], a);   b = dot8f(b2, x4[2 * i + 128], x4[2 * i + 129], b);
        a = dot8f(a3, x4[2 * i + 192], x4[2 * i + 193], a);   b = dot8f(b3, x4[2 * i + 192], x4[2 * i + 193], b);
    }
    for (; i < n8; i += 32) {
        const uint4 a0 = w0[i], b0 = w1[i];
        a = dot8f(a0, x4[2 * i], x4[2 * i + 1], a);
        b = dot8f(b0, x4[2 * i], x4[2 * i + 1], b);
    }
    oa = warp_sum(a);
    ob = warp_sum(b);
}

/* ---- plain phases (QKV, O, down, head): this CTA owns rows [r0, r1) of the phase; the first nz of them go through the zone ---- */
struct RowSpan { int r0, r1, nz; };
__device__ __forceinline__ RowSpan row_span(const Lz &lz, const Seg3 &s, int cols, int cta, int ncta) {
    const int total = s.rows[0] + (s.n > 1 ? s.rows[1] : 0) + (s.n > 2 ? s.rows[2] : 0);
    RowSpan sp;
    row_range(total, cta, ncta, sp.r0, sp.r1);
    sp.nz = lz_rows(lz, sp.r1 - sp.r0, cols);
    return sp;
}
/* after the barrier arrival of the previous phase (every warp is done with the zone): ONE thread issues one bulk copy per
 * matrix segment the CTA's zone rows touch (<= 3); local row i lands at buf + i * cols * 2 */
__device__ __forceinline__ void gemv_lz_issue(const Lz &lz, const Seg3 &s, int cols, int cta, int ncta) {
    const RowSpan sp = row_span(lz, s, cols, cta, ncta);
    if (sp.nz == 0 || threadIdx.x != FUSED_THREADS - 32) return;
    lz_expect(lz.bar, (unsigned)(sp.nz * cols * 2));
    int lo = sp.r0, left = sp.nz, base = 0;
    unsigned char *dst = lz.buf;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        if (k >= s.n || left <= 0) break;
        const int end = base + s.rows[k];
        if (lo < end) {
            int n = end - lo; if (n > left) n = left;
            const size_t bytes = (size_t)n * cols * 2;
            const unsigned char *src = reinterpret_cast<const unsigned char *>(s.w[k]) + (size_t)(lo - base) * cols * 2;
            /* bulk copies in <= 64 KB pieces */
            for (size_t off = 0; off < bytes; off += 65536) lz_copy(dst + off, src + off, (unsigned)(bytes - off < 65536 ? bytes - off : 65536), lz.bar);
            dst += bytes; lo += n; left -= n;
        }
        base = end;
    }
}
/* the streamed row pair a warp starts a plain phase with (both rows valid or rb == ra) */
__device__ __forceinline__ bool first_pair(const RowSpan &sp, int &r, int &rb) {
    r = sp.r0 + sp.nz + (threadIdx.x >> 5);
    rb = r + FUSED_WARPS < sp.r1 ? r + FUSED_WARPS : r;
    return r < sp.r1;
}
__device__ __forceinline__ void gemv_prefetch(WTrip &t, const Lz &lz, const Seg3 &s, int cols, int cta, int ncta) {
    const RowSpan sp = row_span(lz, s, cols, cta, ncta);
    int r, rb;
    if (!first_pair(sp, r, rb)) return;
    float *d0, *d1;
    trip_issue(t, seg_row(s, r, cols, d0), seg_row(s, rb, cols, d1), cols);
}
/* all threads: out[r] = (resid[r] +) W[r,:] . xs for the CTA's rows - streamed rows first (their first trip is in registers and
 * the rest of their loads go out at once), then, after the zone's copies have landed, the zone rows from shared memory */
__device__ __forceinline__ void gemv_run(const WTrip &t, Lz &lz, const Seg3 &s, int cols, const float *xs, const float *resid, int cta, int ncta) {
    const RowSpan sp = row_span(lz, s, cols, cta, ncta);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int r, rb;
    if (first_pair(sp, r, rb)) {
        float *da, *db;
        const uint16_t *pa = seg_row(s, r, cols, da), *pb = seg_row(s, rb, cols, db);
        float va, vb;
        /* the residual values are requested BEFORE the dot products: their L2 round trip hides behind the weights */
        float ea = resid ? __ldcg(resid + r) : 0.0f, eb = resid ? __ldcg(resid + rb) : 0.0f;
        trip_finish(t, pa, pb, xs, cols, va, vb);
        if (lane == 0) {
            *da = ea + va;
            if (rb != r) *db = eb + vb;
        }
        for (r += 2 * FUSED_WARPS; r < sp.r1; r += 2 * FUSED_WARPS) {
            rb = r + FUSED_WARPS < sp.r1 ? r + FUSED_WARPS : r;
            pa = seg_row(s, r, cols, da); pb = seg_row(s, rb, cols, db);
            ea = resid ? __ldcg(resid + r) : 0.0f; eb = resid ? __ldcg(resid + rb) : 0.0f;
            warp_dot2(pa, pb, xs, cols, va, vb);
            if (lane == 0) {
                *da = ea + va;
                if (rb != r) *db = eb + vb;
            }
        }
    }
    if (sp.nz == 0) return;
    lz_wait(lz);
    const size_t zrow = (size_t)cols * 2;
    for (int i = warp; i < sp.nz; i += 2 * FUSED_WARPS) {
        const int ib = i + FUSED_WARPS < sp.nz ? i + FUSED_WARPS : i;
        float *da, *db;
        seg_row(s, sp.r0 + i, cols, da); seg_row(s, sp.r0 + ib, cols, db);
        const float ea = resid ? __ldcg(resid + sp.r0 + i) : 0.0f, eb = resid ? __ldcg(resid + sp.r0 + ib) : 0.0f;
        float va, vb;
        lz_dot2(lz.buf + (size_t)i * zrow, lz.buf + (size_t)ib * zrow, xs, cols, va, vb);
        if (lane == 0) {
            *da = ea + va;
            if (ib != i) *db = eb + vb;
        }
    }
}

/* ---- gate/up: unit u = (gate row u, up row u) -> g[u] = silu(gate.x) * (up.x); the CTA owns units [u0, u1) ------- */
__device__ __forceinline__ RowSpan unit_span(const Lz &lz, int inter, int cols, int cta, int ncta) {
    RowSpan sp;
    row_range(inter, cta, ncta, sp.r0, sp.r1);
    sp.nz = lz_rows(lz, 2 * (sp.r1 - sp.r0), cols) >> 1;
    return sp;
}
/* zone layout: nz gate rows, then nz up rows (two bulk copies) */
__device__ __forceinline__ void swiglu_lz_issue(const Lz &lz, const uint16_t *gate, const uint16_t *up, int inter, int cols, int cta, int ncta) {
    const RowSpan sp = unit_span(lz, inter, cols, cta, ncta);
    if (sp.nz == 0 || threadIdx.x != FUSED_THREADS - 32) return;
    const size_t bytes = (size_t)sp.nz * cols * 2;
    lz_expect(lz.bar, (unsigned)(2 * bytes));
    const unsigned char *sg = reinterpret_cast<const unsigned char *>(gate + (size_t)sp.r0 * cols), *su = reinterpret_cast<const unsigned char *>(up + (size_t)sp.r0 * cols);
    for (size_t off = 0; off < bytes; off += 65536) {
        const unsigned n = (unsigned)(bytes - off < 65536 ? bytes - off : 65536);
        lz_copy(lz.buf + off, sg + off, n, lz.bar);
        lz_copy(lz.buf + bytes + off, su + off, n, lz.bar);
    }
}
__device__ __forceinline__ void swiglu_prefetch(WTrip &t, const Lz &lz, const uint16_t *gate, const uint16_t *up, int inter, int cols, int cta, int ncta) {
    const RowSpan sp = unit_span(lz, inter, cols, cta, ncta);
    const int u = sp.r0 + sp.nz + (threadIdx.x >> 5);
    if (u < sp.r1) trip_issue(t, gate + (size_t)u * cols, up + (size_t)u * cols, cols);
}
__device__ __forceinline__ void swiglu_run(const WTrip &t, Lz &lz, const uint16_t *gate, const uint16_t *up, float *g, int inter, int cols,
                                           const float *xs, int cta, int ncta) {
    const RowSpan sp = unit_span(lz, inter, cols, cta, ncta);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int u = sp.r0 + sp.nz + warp;
    if (u < sp.r1) {
        float va, vb;
        trip_finish(t, gate + (size_t)u * cols, up + (size_t)u * cols, xs, cols, va, vb);
        if (lane == 0) g[u] = swiglu_act<QTC_FAST>(va, vb);
        for (u += FUSED_WARPS; u < sp.r1; u += FUSED_WARPS) {
            warp_dot2(gate + (size_t)u * cols, up + (size_t)u * cols, xs, cols, va, vb);
            if (lane == 0) g[u] = swiglu_act<QTC_FAST>(va, vb);
        }
    }
    if (sp.nz == 0) return;
    lz_wait(lz);
    const size_t zrow = (size_t)cols * 2;
    const unsigned char *zu = lz.buf + (size_t)sp.nz * zrow;
    for (int i = warp; i < sp.nz; i += FUSED_WARPS) {
        float va, vb;
        lz_dot2(lz.buf + (size_t)i * zrow, zu + (size_t)i * zrow, xs, cols, va, vb);
        if (lane == 0) g[sp.r0 + i] = swiglu_act<QTC_FAST>(va, vb);
    }
}

#include "qtc_sample.cuh"

/* k_lm_fused(a0, sp): sp.G == 0 -> one stack step (talker: + codec head + KV length);
 *                     sp.G  > 0 -> the sub-talker's G steps with gathers, heads and sampling inside. */
__global__ void __launch_bounds__(FUSED_BLOCK, 1) k_lm_fused(const FusedArgs a0, const SubProg sp) {
    extern __shared__ __align__(128) float fused_smem[];
    float *sm = fused_smem;
    __shared__ float red[32];
    __shared__ SampleSmem S_samp;
    __shared__ int s_codes[QTC_MAX_GROUPS];
    __shared__ uint32_t s_rng;
    __shared__ __align__(8) uint64_t lz_bar;
    const FusedArgs &a = a0;                         /* stays in the constant bank: no registers held for its fields */
    const uint16_t *head_w = a0.head_w;             /* the sub-talker program switches heads per step */
    int head_rows = a0.head_rows;
    const LmDev &lm = *a.lm;
    const int cta = blockIdx.x, ncta = gridDim.x, tid = threadIdx.x;
    const int H = lm.hidden, I = lm.inter, L = lm.layers;
    float *att_sm = sm + a.xs_floats, *kvs = att_sm + a.att_floats;
    Lz lz;
    lz.buf = reinterpret_cast<unsigned char *>(fused_smem) + a.lz_off; lz.bar = lz_u32(&lz_bar); lz.bytes = a.lz_bytes; lz.parity = 0;
    if (tid == 0) lz_bar_init(&lz_bar);
    const int pos_base = *a.pos_ptr + a.pos_add;
    unsigned long long target = 0;
    if (tid == 0) { const unsigned long long v = ld_acquire_u64(a.bar); target = v - v % (unsigned long long)ncta; }
    const int nsteps = sp.G > 0 ? sp.G : 1;
    if (sp.G > 0 && tid == 0) s_codes[0] = sp.codes[0];
    const bool pf_l2 = (a.flags & 4) != 0;          /* experiment: L2 prefetch of the next phases at the END of a phase */
    Seg3 s; int cols;
    WTrip tp;
    __syncthreads();                                /* the mbarrier is initialised */
    if (sp.G == 0) { phase_seg(a, lm, 0, 0, s, cols, head_w, head_rows); gemv_lz_issue(lz, s, cols, cta, ncta); gemv_prefetch(tp, lz, s, cols, cta, ncta); }

    for (int g = 0; g < nsteps; g++) {
        const int pos = pos_base + g;
        /* flags & 8: the older 32-position splits with block-level reductions */
        const bool warp_att_ok = !(a.flags & 8) && lm.head_dim == 128 && a.att_floats >= 2 * FUSED_WARPS + FUSED_WARPS * 128;
        const AttPlan plan = att_plan(lm, pos, cta, ncta, ATT_STAGE_MAX, 0, warp_att_ok ? 64 : 32);
        const bool warp_att = warp_att_ok && plan.staged && plan.t1 - plan.t0 <= 64;
        const bool small_att = !(a.flags & 2) && plan.S == 1 && plan.staged && lm.head_dim == 128 && pos < 32;   /* every sub-talker step */
        if (sp.G > 0) {
            /* ST_PROJECT_INPUT: gather (every CTA for itself) -> input projection rows -> x.  The zone is idle here (the
             * previous step's head has been consumed, a grid barrier and the sampler lie in between): layer 0's QKV rows
             * start flying before the gather */
            compute_sync();
            head_w = g >= 1 ? sp.heads[g - 1] : NULL;
            head_rows = sp.vocab;
            phase_seg(a, lm, 0, 0, s, cols, head_w, head_rows);
            gemv_lz_issue(lz, s, cols, cta, ncta);
            if (g == 0) stage_vec(sm, sp.hidden, sp.Hin);
            else {
                const uint16_t *row = (g == 1 ? sp.emb0 : sp.emb[g - 2]) + (size_t)s_codes[g - 1] * sp.Hin;
                for (int i = tid; i < sp.Hin; i += FUSED_THREADS) sm[i] = bf16_to_f32(row[i]);
                compute_sync();
            }
            if (sp.in_w) {
                Seg3 si;
                si.n = 1; si.w[0] = sp.in_w; si.out[0] = a.x; si.rows[0] = H;
                cta_gemv(si, sp.Hin, sm, sp.in_b, cta, ncta);
            } else if (cta == 0) {
                for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = i < sp.Hin ? sm[i] : 0.0f;
            }
            barrier_arrive(a.bar, target, ncta);
            gemv_prefetch(tp, lz, s, cols, cta, ncta);
            barrier_wait(a.bar, target);
        }
#define PROF(i) do { if (a.prof && tid == 0 && l == a.prof_layer) a.prof[cta * 32 + (i)] = clock64(); } while (0)
        for (int l = 0; l < L; l++) {
            const LmLayerDev &w = lm.layer[l];
            /* P1: RMSNorm -> Q,K,V (rows in the landing zone, first trip of the rest in registers) */
            PROF(0);
            stage_rmsnorm(sm, a.x, w.in_norm, H, lm.eps, red);
            PROF(1);
            phase_seg(a, lm, l, 0, s, cols, head_w, head_rows);
            gemv_run(tp, lz, s, cols, sm, NULL, cta, ncta);
            att_stage_issue(a, lm, plan, l, pos, kvs);
            if (pf_l2) {       /* HBM idles through the attention phase and its two barriers: pull O and gate/up into L2 meanwhile */
                Seg3 s2; int c2;
                phase_seg(a, lm, l, 1, s2, c2, head_w, head_rows);
                prefetch_seg(s2, c2, cta, ncta);
                prefetch_swiglu(w.gate, w.up, I, H, cta, ncta);
            }
            PROF(2);
            barrier_arrive(a.bar, target, ncta);
            phase_seg(a, lm, l, 1, s, cols, head_w, head_rows);
            gemv_lz_issue(lz, s, cols, cta, ncta);                  /* O rows land during the attention phase */
            barrier_wait(a.bar, target);
            PROF(3);
            /* P2: attention partials */
            if (small_att) cta_attention_small(a, lm, plan, l, pos, kvs);
            else if (warp_att) cta_attention_warps(a, lm, plan, l, pos, att_sm, kvs);
            else cta_attention(a, lm, plan, l, pos, att_sm, kvs, red);
            PROF(4);
            barrier_arrive(a.bar, target, ncta);
            gemv_prefetch(tp, lz, s, cols, cta, ncta);
            barrier_wait(a.bar, target);
            PROF(5);
            /* P3: combine partials -> O projection + residual */
            if (plan.S == 1 && (small_att || !warp_att)) stage_vec(sm, a.att, lm.heads * lm.head_dim);
            else stage_attention_out(sm, att_sm, kvs, a, lm, plan.S, (a.prof && tid == 0 && l == a.prof_layer) ? a.prof + cta * 32 : NULL);
            PROF(15);
            gemv_run(tp, lz, s, cols, sm, a.x, cta, ncta);
            if (pf_l2) { Seg3 s2; int c2; phase_seg(a, lm, l, 3, s2, c2, head_w, head_rows); prefetch_seg(s2, c2, cta, ncta); }      /* down */
            PROF(6);
            barrier_arrive(a.bar, target, ncta);
            swiglu_lz_issue(lz, w.gate, w.up, I, H, cta, ncta);
            swiglu_prefetch(tp, lz, w.gate, w.up, I, H, cta, ncta);
            barrier_wait(a.bar, target);
            PROF(7);
            /* P4: RMSNorm -> gate/up -> SwiGLU */
            stage_rmsnorm(sm, a.x, w.post_norm, H, lm.eps, red);
            PROF(8);
            swiglu_run(tp, lz, w.gate, w.up, a.g, I, H, sm, cta, ncta);
            if (pf_l2) { Seg3 s2; int c2; if (phase_seg(a, lm, l + 1, 0, s2, c2, head_w, head_rows)) prefetch_seg(s2, c2, cta, ncta); }  /* next QKV / head */
            PROF(9);
            barrier_arrive(a.bar, target, ncta);
            phase_seg(a, lm, l, 3, s, cols, head_w, head_rows);
            gemv_lz_issue(lz, s, cols, cta, ncta);
            gemv_prefetch(tp, lz, s, cols, cta, ncta);
            barrier_wait(a.bar, target);
            PROF(10);
            /* P5: down projection + residual */
            stage_vec(sm, a.g, I);
            PROF(11);
            gemv_run(tp, lz, s, cols, sm, a.x, cta, ncta);
            PROF(12);
            barrier_arrive(a.bar, target, ncta);
            if (phase_seg(a, lm, l + 1, 0, s, cols, head_w, head_rows)) {                /* next layer's QKV, or the head */
                gemv_lz_issue(lz, s, cols, cta, ncta);
                gemv_prefetch(tp, lz, s, cols, cta, ncta);
            }
            barrier_wait(a.bar, target);
            PROF(13);
        }
#undef PROF
        /* final norm (every CTA, into shared memory).  Talker: x is then overwritten in place with the normed hidden (ctx->tk_x),
         * which needs every CTA to have staged x first - one grid barrier.  Inside the sub-talker loop nobody reads the hidden
         * state back, so there is no overwrite and no barrier; position 0 has no head either, so it skips the norm as well. */
        if (sp.G == 0 || head_w) stage_rmsnorm(sm, a.x, lm.final_norm, H, lm.eps, red);
        if (sp.G == 0) {
            grid_barrier(a.bar, target, ncta);
            if (cta == 0) {
                for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = sm[i];
                if (tid == 0 && a.st && !a.st->done) a.st->pos += 1;
            }
        }
        if (phase_seg(a, lm, L, 0, s, cols, head_w, head_rows)) gemv_run(tp, lz, s, cols, sm, NULL, cta, ncta);
        if (sp.G > 0 && g >= 1) {
            /* every CTA samples for itself from the same logits: same code everywhere, no broadcast needed */
            grid_barrier(a.bar, target, ncta);
            const int V = sp.vocab;
            float *lg = sm, *vv = sm + V;
            for (int i = tid; i < V; i += FUSED_THREADS) lg[i] = __ldcg(a.logits + i);
            if (g == 1 && tid == 0) s_rng = sp.seed_bits;
            __syncthreads();
            /* flags & 64 (diagnostic): a fixed code instead of the sample, to time the sampler's share of the frame */
            const int tok = (a.flags & 64) ? 1 : block_sample(lg, vv, V, sp.cfg.top_k, sp.cfg.top_p, sp.cfg.temperature, &s_rng, S_samp, sp.scratch_f, sp.scratch_i);
            if (tid == 0) { s_codes[g] = tok; if (cta == 0) sp.codes[g] = tok; }
        }
    }
}
