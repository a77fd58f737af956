/*
 * qtc_fused_main.cuh - k_lm_fused: one decoder-stack step as one persistent cooperative kernel,
 * 148 CTAs x 16 warps.  See qtc_fused.cuh (barriers, staging, row split) and qtc_fused_kernel.cuh (attention).
 *
 * Weights do not depend on activations, so the first "trip" of every GEMV phase - each warp's first two
 * rows, eight 128-bit loads per lane per row, 64 registers - is issued right after the CTA ARRIVES at the
 * grid barrier that precedes the phase and is consumed after the barrier and the input staging: one DRAM
 * round trip per phase moves off the critical path.  (A TMA-bulk shared-memory ring and an L2 bulk prefetch
 * two phases ahead were both measured slower on B200 at batch 1: profiles/README.md.)
 *
 * dynamic shared memory: [ xs | attention scratch | KV stage 2 x 64 x hd ]
 */
#pragma once
#include "qtc_fused_kernel.cuh"

#define FUSED_BLOCK FUSED_THREADS

#define PFJ 4                      /* 16-byte pieces per lane per row held across the barrier */
struct WTrip { uint4 a[PFJ], b[PFJ]; };

/* rows (ra, rb) of `cols` bf16: lane's pieces lane+32j, j<8 */
__device__ __forceinline__ void trip_issue(WTrip &t, const uint16_t *ra, const uint16_t *rb, int cols) {
    const int lane = threadIdx.x & 31, n8 = cols >> 3;
    const uint4 *w0 = reinterpret_cast<const uint4 *>(ra), *w1 = reinterpret_cast<const uint4 *>(rb);
#pragma unroll
    for (int j = 0; j < PFJ; j++) {
        const int p = lane + 32 * j;
        if (p < n8) { t.a[j] = ldg_stream(w0 + p); t.b[j] = ldg_stream(w1 + p); }
    }
}
/* finish the two dot products: the prefetched pieces, then (cols > 2048) the rest of the rows */
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

/* the row pair a warp starts a plain phase with (both rows valid or rb == ra) */
__device__ __forceinline__ bool first_pair(const Seg3 &s, int cols, int cta, int ncta, int &r, int &rb, int &r1) {
    const int total = s.rows[0] + (s.n > 1 ? s.rows[1] : 0) + (s.n > 2 ? s.rows[2] : 0);
    int r0;
    row_range(total, cta, ncta, r0, r1);
    r = r0 + (threadIdx.x >> 5);
    rb = r + FUSED_WARPS < r1 ? r + FUSED_WARPS : r;
    return r < r1;
}
__device__ __forceinline__ void gemv_prefetch(WTrip &t, const Seg3 &s, int cols, int cta, int ncta) {
    int r, rb, r1;
    if (!first_pair(s, cols, cta, ncta, r, rb, r1)) return;
    float *d0, *d1;
    trip_issue(t, seg_row(s, r, cols, d0), seg_row(s, rb, cols, d1), cols);
}
__device__ __forceinline__ void gemv_run(const WTrip &t, const Seg3 &s, int cols, const float *xs, const float *resid, int cta, int ncta) {
    int r, rb, r1;
    if (!first_pair(s, cols, cta, ncta, r, rb, r1)) return;
    const int lane = threadIdx.x & 31;
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
    for (r += 2 * FUSED_WARPS; r < r1; r += 2 * FUSED_WARPS) {
        rb = r + FUSED_WARPS < r1 ? r + FUSED_WARPS : r;
        pa = seg_row(s, r, cols, da); pb = seg_row(s, rb, cols, db);
        ea = resid ? __ldcg(resid + r) : 0.0f; eb = resid ? __ldcg(resid + rb) : 0.0f;
        warp_dot2(pa, pb, xs, cols, va, vb);
        if (lane == 0) {
            *da = ea + va;
            if (rb != r) *db = eb + vb;
        }
    }
}
/* gate/up: the pair is (gate row u, up row u) */
__device__ __forceinline__ void swiglu_prefetch(WTrip &t, const uint16_t *gate, const uint16_t *up, int inter, int cols, int cta, int ncta) {
    int u0, u1;
    row_range(inter, cta, ncta, u0, u1);
    const int u = u0 + (threadIdx.x >> 5);
    if (u < u1) trip_issue(t, gate + (size_t)u * cols, up + (size_t)u * cols, cols);
}
__device__ __forceinline__ void swiglu_run(const WTrip &t, const uint16_t *gate, const uint16_t *up, float *g, int inter, int cols,
                                           const float *xs, int cta, int ncta) {
    int u0, u1;
    row_range(inter, cta, ncta, u0, u1);
    const int lane = threadIdx.x & 31;
    int u = u0 + (threadIdx.x >> 5);
    if (u >= u1) return;
    float va, vb;
    trip_finish(t, gate + (size_t)u * cols, up + (size_t)u * cols, xs, cols, va, vb);
    if (lane == 0) g[u] = swiglu_act<QTC_FAST>(va, vb);
    for (u += FUSED_WARPS; u < u1; u += FUSED_WARPS) {
        warp_dot2(gate + (size_t)u * cols, up + (size_t)u * cols, xs, cols, va, vb);
        if (lane == 0) g[u] = swiglu_act<QTC_FAST>(va, vb);
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
    FusedArgs a = a0;
    const LmDev &lm = *a.lm;
    const int cta = blockIdx.x, ncta = gridDim.x, tid = threadIdx.x;
    const int H = lm.hidden, I = lm.inter, L = lm.layers;
    float *att_sm = sm + a.xs_floats, *kvs = att_sm + a.att_floats;
    const int pos_base = *a.pos_ptr + a.pos_add;
    unsigned long long target = 0;
    if (tid == 0) { const unsigned long long v = ld_acquire_u64(a.bar); target = v - v % (unsigned long long)ncta; }
    const int nsteps = sp.G > 0 ? sp.G : 1;
    if (sp.G > 0 && tid == 0) s_codes[0] = sp.codes[0];
    const bool pf_l2 = (a.flags & 4) != 0;          /* experiment: L2 prefetch of the next phases at the END of a phase */
    Seg3 s; int cols;
    WTrip tp;
    if (sp.G == 0) { phase_seg(a, lm, 0, 0, s, cols); gemv_prefetch(tp, s, cols, cta, ncta); }

    for (int g = 0; g < nsteps; g++) {
        const int pos = pos_base + g;
        /* flags & 8: the older 32-position splits with block-level reductions */
        const bool warp_att_ok = !(a.flags & 8) && lm.head_dim == 128 && a.att_floats >= 2 * FUSED_WARPS + FUSED_WARPS * 128;
        const AttPlan plan = att_plan(lm, pos, cta, ncta, ATT_STAGE_MAX, 0, warp_att_ok ? 64 : 32);
        const bool warp_att = warp_att_ok && plan.staged && plan.t1 - plan.t0 <= 64;
        const bool small_att = !(a.flags & 2) && plan.S == 1 && plan.staged && lm.head_dim == 128 && pos < 32;   /* every sub-talker step */
        if (sp.G > 0) {
            /* ST_PROJECT_INPUT: gather (every CTA for itself) -> input projection rows -> x */
            compute_sync();
            if (g == 0) stage_vec(sm, sp.hidden, sp.Hin);
            else {
                const uint16_t *row = (g == 1 ? sp.emb0 : sp.emb[g - 2]) + (size_t)s_codes[g - 1] * sp.Hin;
                for (int i = tid; i < sp.Hin; i += FUSED_THREADS) sm[i] = bf16_to_f32(row[i]);
                compute_sync();
            }
            if (sp.in_w) {
                s.n = 1; s.w[0] = sp.in_w; s.out[0] = a.x; s.rows[0] = H; cols = sp.Hin;
                cta_gemv(s, cols, sm, sp.in_b, cta, ncta);
            } else if (cta == 0) {
                for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = i < sp.Hin ? sm[i] : 0.0f;
            }
            a.head_w = g >= 1 ? sp.heads[g - 1] : NULL;
            a.head_rows = sp.vocab;
            barrier_arrive(a.bar, target, ncta);
            phase_seg(a, lm, 0, 0, s, cols);
            gemv_prefetch(tp, s, cols, cta, ncta);
            barrier_wait(a.bar, target);
        }
#define PROF(i) do { if (a.prof && tid == 0 && l == a.prof_layer) a.prof[cta * 32 + (i)] = clock64(); } while (0)
        for (int l = 0; l < L; l++) {
            const LmLayerDev &w = lm.layer[l];
            /* P1: RMSNorm -> Q,K,V (first trip already in registers) */
            PROF(0);
            stage_rmsnorm(sm, a.x, w.in_norm, H, lm.eps, red);
            PROF(1);
            phase_seg(a, lm, l, 0, s, cols);
            gemv_run(tp, s, cols, sm, NULL, cta, ncta);
            att_stage_issue(a, lm, plan, l, pos, kvs);
            if (pf_l2) {       /* HBM idles through the attention phase and its two barriers: pull O and gate/up into L2 meanwhile */
                Seg3 s2; int c2;
                phase_seg(a, lm, l, 1, s2, c2);
                prefetch_seg(s2, c2, cta, ncta);
                prefetch_swiglu(w.gate, w.up, I, H, cta, ncta);
            }
            PROF(2);
            grid_barrier(a.bar, target, ncta);
            PROF(3);
            /* P2: attention partials */
            if (small_att) cta_attention_small(a, lm, plan, l, pos, kvs);
            else if (warp_att) cta_attention_warps(a, lm, plan, l, pos, att_sm, kvs);
            else cta_attention(a, lm, plan, l, pos, att_sm, kvs, red);
            PROF(4);
            barrier_arrive(a.bar, target, ncta);
            phase_seg(a, lm, l, 1, s, cols);
            if (!(a.flags & 32)) gemv_prefetch(tp, s, cols, cta, ncta);
            barrier_wait(a.bar, target);
            PROF(5);
            /* P3: combine partials -> O projection + residual */
            if (plan.S == 1 && (small_att || !warp_att)) stage_vec(sm, a.att, lm.heads * lm.head_dim);
            else stage_attention_out(sm, att_sm, kvs, a, lm, plan.S, (a.prof && tid == 0 && l == a.prof_layer) ? a.prof + cta * 32 : NULL);
            PROF(15);
            if (a.flags & 32) gemv_prefetch(tp, s, cols, cta, ncta);      /* diagnostic: O-projection trip issued after the staging */
            gemv_run(tp, s, cols, sm, a.x, cta, ncta);
            if (pf_l2) { Seg3 s2; int c2; phase_seg(a, lm, l, 3, s2, c2); prefetch_seg(s2, c2, cta, ncta); }      /* down */
            PROF(6);
            barrier_arrive(a.bar, target, ncta);
            swiglu_prefetch(tp, w.gate, w.up, I, H, cta, ncta);
            barrier_wait(a.bar, target);
            PROF(7);
            /* P4: RMSNorm -> gate/up -> SwiGLU */
            stage_rmsnorm(sm, a.x, w.post_norm, H, lm.eps, red);
            PROF(8);
            swiglu_run(tp, w.gate, w.up, a.g, I, H, sm, cta, ncta);
            if (pf_l2) { Seg3 s2; int c2; if (phase_seg(a, lm, l + 1, 0, s2, c2)) prefetch_seg(s2, c2, cta, ncta); }  /* next QKV / head */
            PROF(9);
            barrier_arrive(a.bar, target, ncta);
            phase_seg(a, lm, l, 3, s, cols);
            gemv_prefetch(tp, s, cols, cta, ncta);
            barrier_wait(a.bar, target);
            PROF(10);
            /* P5: down projection + residual */
            stage_vec(sm, a.g, I);
            PROF(11);
            gemv_run(tp, s, cols, sm, a.x, cta, ncta);
            PROF(12);
            barrier_arrive(a.bar, target, ncta);
            if (phase_seg(a, lm, l + 1, 0, s, cols)) gemv_prefetch(tp, s, cols, cta, ncta);   /* next layer's QKV, or the head */
            barrier_wait(a.bar, target);
            PROF(13);
        }
#undef PROF
        /* final norm (every CTA, into shared memory).  Talker: x is then overwritten in place with the normed hidden (ctx->tk_x),
         * which needs every CTA to have staged x first - one grid barrier.  Inside the sub-talker loop nobody reads the hidden
         * state back, so there is no overwrite and no barrier; position 0 has no head either, so it skips the norm as well. */
        if (sp.G == 0 || a.head_w) stage_rmsnorm(sm, a.x, lm.final_norm, H, lm.eps, red);
        if (sp.G == 0) {
            grid_barrier(a.bar, target, ncta);
            if (cta == 0) {
                for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = sm[i];
                if (tid == 0 && a.st && !a.st->done) a.st->pos += 1;
            }
        }
        if (phase_seg(a, lm, L, 0, s, cols)) gemv_run(tp, s, cols, sm, NULL, cta, ncta);
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
