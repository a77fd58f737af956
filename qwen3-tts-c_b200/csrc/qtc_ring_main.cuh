/*
 * qtc_ring_main.cuh - k_lm_ring: one decoder-stack step (talker: + codec head + KV length), or the
 * sub-talker's whole code-group loop (qwen_tts_subtalker_generate, c/qwen_tts_talker.c:539-736), as ONE
 * persistent cooperative kernel: 148 CTAs x (16 consumer warps + 1 producer warp).
 *
 * Same phases and grid barriers as k_lm_fused (qtc_fused_main.cuh); what changes is where the weights come
 * from: the producer warp streams this CTA's rows of every matrix of the launch, in phase order, through the
 * shared-memory ring of qtc_ring.cuh, independent of the barriers the consumers wait on.
 *
 * dynamic shared memory: [ xs | attention scratch | KV stage / sampler | partials | ring ]
 */
#pragma once
#include "qtc_ring.cuh"
#include "qtc_sample.cuh"

/* the weight phases of one launch, in the order BOTH sides walk them:
 *   for each step g: [input projection] , layers x {QKV, O, gate/up, down} , [head] */
__device__ __forceinline__ const uint16_t *ring_head_of(const FusedArgs &a, const SubProg &sp, int g) {
    if (sp.G > 0) return g >= 1 ? sp.heads[g - 1] : NULL;
    return a.head_w;
}

__device__ __forceinline__ void ring_producer_main(const Ring &rg, const FusedArgs &a, const SubProg &sp, const LmDev &lm, int cta, int ncta) {
    RingPos rp = {0, 0u};
    const int nsteps = sp.G > 0 ? sp.G : 1;
    Seg3 s; int cols;
    for (int g = 0; g < nsteps; g++) {
        if (sp.G > 0 && sp.in_w) {
            s.n = 1; s.w[0] = sp.in_w; s.out[0] = a.x; s.rows[0] = lm.hidden;
            ring_produce(rg, rp, s, sp.Hin, 0, cta, ncta);
        }
        for (int l = 0; l < lm.layers; l++)
            for (int k = 0; k < 4; k++) {
                phase_seg(a, lm, l, k, s, cols);
                ring_produce(rg, rp, s, cols, k == 2, cta, ncta, (a.prof && k == 0 && l == a.prof_layer) ? a.prof + cta * 32 : NULL);
            }
        const uint16_t *hw = ring_head_of(a, sp, g);
        if (hw) {
            s.n = 1; s.w[0] = hw; s.out[0] = a.logits; s.rows[0] = a.head_rows;
            ring_produce(rg, rp, s, lm.hidden, 0, cta, ncta);
        }
    }
}

/* Short contexts (n = pos + 1 <= 32 positions; always true for the sub-talker, c/qwen_tts_talker.c:656-672): EVERY CTA
 * computes the attention of all heads for itself, a warp per head, and leaves the O projection's input vector in
 * shared memory.  That costs each CTA n x 2 x kvd x 4 bytes of L2 reads per layer (<= 139 KB at n = 17) and saves
 * the attention phase's two grid barriers plus the hand-over of the head outputs through HBM/L2.
 * Lane i owns dims 4i..4i+3 of the head; K/V rows of earlier positions come straight from L2, eight rows in flight.
 * scr: per-warp scratch [16][2][hd].  The KV rows of the new position are appended by CTA (kv head mod grid). */
__device__ __noinline__ void cta_attention_small(const FusedArgs &a, const LmDev &lm, int l, int pos, float *xs, float *scr, int cta, int ncta) {
    const int hd = lm.head_dim, half = hd >> 1, rep = lm.heads / lm.kv_heads, kvd = lm.kv_heads * hd;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n = pos + 1;
    const LmLayerDev &w = lm.layer[l];
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;
    float *kc = a.kc + (size_t)l * a.kv_max * kvd, *vc = a.vc + (size_t)l * a.kv_max * kvd;
    float *qs = scr + (size_t)warp * 2 * hd, *ks = qs + hd;
    const float scale = 1.0f / sqrtf((float)hd);
    const bool act = lane * 4 < hd;
    for (int h = warp; h < lm.heads; h += FUSED_WARPS) {
        const int g = h / rep;
        const float *qr = a.q + h * hd, *kr = a.kk + g * hd;
        /* per-head RMSNorm + rotate-half RoPE of q and of the new k (c/qwen_tts_talker.c:150-189) */
        float sq = 0.0f, sk = 0.0f;
        for (int i = lane; i < hd; i += 32) {
            const float vq = __ldcg(qr + i), vk = __ldcg(kr + i);
            qs[i] = vq; ks[i] = vk;
            sq = fmaf(vq, vq, sq); sk = fmaf(vk, vk, sk);
        }
        const float4 v_new = act ? __ldcg(reinterpret_cast<const float4 *>(a.vv + g * hd) + lane) : make_float4(0, 0, 0, 0);
        sq = warp_sum(sq); sk = warp_sum(sk);
        const float iq = 1.0f / sqrtf(sq / (float)hd + lm.eps), ik = 1.0f / sqrtf(sk / (float)hd + lm.eps);
        __syncwarp();
        for (int i = lane; i < half; i += 32) {
            const float c0 = cs[i], c1 = cs[i + half], s0 = sn[i], s1 = sn[i + half];
            float xa = qs[i] * iq * w.q_norm[i], xb = qs[i + half] * iq * w.q_norm[i + half];
            qs[i] = xa * c0 - xb * s0; qs[i + half] = xb * c1 + xa * s1;
            xa = ks[i] * ik * w.k_norm[i]; xb = ks[i + half] * ik * w.k_norm[i + half];
            ks[i] = xa * c0 - xb * s0; ks[i + half] = xb * c1 + xa * s1;
        }
        __syncwarp();
        const float4 q4 = act ? *reinterpret_cast<const float4 *>(qs + lane * 4) : make_float4(0, 0, 0, 0);
        const float4 k_new = act ? *reinterpret_cast<const float4 *>(ks + lane * 4) : make_float4(0, 0, 0, 0);
        if (act && h % rep == 0 && cta == g % ncta) {
            *reinterpret_cast<float4 *>(kc + (size_t)pos * kvd + g * hd + lane * 4) = k_new;
            *reinterpret_cast<float4 *>(vc + (size_t)pos * kvd + g * hd + lane * 4) = v_new;
        }
        /* scores: lane t ends up holding the score of position t */
        float sc = -3.402823466e+38f;
        for (int t0 = 0; t0 < n; t0 += 8) {
            float d[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int t = t0 + u;
                float4 kv = make_float4(0, 0, 0, 0);
                if (t < pos) { if (act) kv = __ldcg(reinterpret_cast<const float4 *>(kc + (size_t)t * kvd + g * hd) + lane); }
                else if (t == pos) kv = k_new;
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
            for (int u = 0; u < 8; u++) if (lane == t0 + u && t0 + u < n) sc = d[u] * scale;
        }
        const float mx = warp_max(sc);
        const float e = lane < n ? __expf(sc - mx) : 0.0f;
        const float tot = warp_sum(e);
        float4 o4 = make_float4(0, 0, 0, 0);
        for (int t0 = 0; t0 < n; t0 += 8) {
            float4 vv[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int t = t0 + u;
                vv[u] = make_float4(0, 0, 0, 0);
                if (t < pos) { if (act) vv[u] = __ldcg(reinterpret_cast<const float4 *>(vc + (size_t)t * kvd + g * hd) + lane); }
                else if (t == pos) vv[u] = v_new;
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const float p = __shfl_sync(0xffffffffu, e, (t0 + u) & 31);
                if (t0 + u < n) { o4.x = fmaf(p, vv[u].x, o4.x); o4.y = fmaf(p, vv[u].y, o4.y); o4.z = fmaf(p, vv[u].z, o4.z); o4.w = fmaf(p, vv[u].w, o4.w); }
            }
        }
        if (act) *reinterpret_cast<float4 *>(xs + h * hd + lane * 4) = make_float4(o4.x / tot, o4.y / tot, o4.z / tot, o4.w / tot);
        __syncwarp();
    }
}

__global__ void __launch_bounds__(RING_THREADS, 1) k_lm_ring(const __grid_constant__ FusedArgs a, const __grid_constant__ SubProg sp,
                                                              const __grid_constant__ LmDev lm) {
    extern __shared__ __align__(128) float fused_smem[];
    __shared__ float red[32];
    __shared__ int s_codes[QTC_MAX_GROUPS];
    __shared__ uint32_t s_rng;
    __shared__ __align__(8) uint64_t s_mbar[2 * RING_MAX_STAGES];
    const int cta = blockIdx.x, ncta = gridDim.x, tid = threadIdx.x;
    const int H = lm.hidden, L = lm.layers;
    float *sm = fused_smem, *att_sm = sm + a.xs_floats, *kvs = att_sm + a.att_floats;
    uint16_t *xa = reinterpret_cast<uint16_t *>(kvs);     /* split input vector of the phase in flight (the KV stage is idle then) */
    Ring rg;
    rg.part = kvs + a.kv_floats;
    rg.buf = reinterpret_cast<unsigned char *>(fused_smem) + a.ring_off;
    rg.full = s_mbar; rg.empty = s_mbar + RING_MAX_STAGES; rg.nst = a.nst;
    if (tid == 0) {
        for (int i = 0; i < rg.nst; i++) { mbar_init(&rg.full[i], 1); mbar_init(&rg.empty[i], RING_GW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid >= FUSED_THREADS) {            /* producer warp: never joins a block barrier again */
        ring_producer_main(rg, a, sp, lm, cta, ncta);
        return;
    }
    RingPos rp = {0, 0u};
    PhaseGeom pg;
    Seg3 s; int cols;
    const int pos_base = *a.pos_ptr + a.pos_add;
    unsigned long long target = 0;
    if (tid == 0) { const unsigned long long v = ld_acquire_u64(a.bar); target = v - v % (unsigned long long)ncta; }
    const int nsteps = sp.G > 0 ? sp.G : 1;
    if (sp.G > 0 && tid == 0) s_codes[0] = sp.codes[0];

    for (int g = 0; g < nsteps; g++) {
        const int pos = pos_base + g;
        const AttPlan plan = att_plan(lm, pos, cta, ncta, a.att_cap);
        const bool small = pos + 1 <= 32 && lm.head_dim <= 128 && lm.head_dim % 4 == 0 && (a.flags & 4);
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
                s.n = 1; s.w[0] = sp.in_w; s.out[0] = a.x; s.rows[0] = H;
                xa_build<false>(xa, sm, sp.Hin);
                pg = phase_geom(s, sp.Hin, 0, cta, ncta);
                rp = ring_consume(rg, rp, pg, xa);
                ring_finalize(rg, pg, s, sp.in_b);
            } else if (cta == 0) {
                for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = i < sp.Hin ? sm[i] : 0.0f;
            }
            grid_barrier(a.bar, target, ncta);
        }
#define PROF(i) do { if (a.prof && tid == 0 && l == a.prof_layer) a.prof[cta * 32 + (i)] = clock64(); } while (0)
        for (int l = 0; l < L; l++) {
            const LmLayerDev &w = lm.layer[l];
            /* P1: RMSNorm -> Q,K,V.  The KV rows of earlier positions start moving to shared memory now */
            PROF(0);
            stage_rmsnorm(sm, a.x, w.in_norm, H, lm.eps, red);
            xa_build<false>(xa, sm, H);
            PROF(1);
            phase_seg(a, lm, l, 0, s, cols);
            PROF(26);
            pg = phase_geom(s, cols, 0, cta, ncta);
            rp = ring_consume(rg, rp, pg, xa);
            PROF(16);
            ring_finalize(rg, pg, s, NULL);
            if (!small) att_stage_issue(a, lm, plan, l, pos, kvs);     /* every warp is past its last read of xa: the KV stage is free */
            PROF(2);
            grid_barrier(a.bar, target, ncta);
            PROF(3);
            if (small) {
                /* P2+P3: every CTA computes all heads itself -> O projection + residual */
                cta_attention_small(a, lm, l, pos, sm, kvs, cta, ncta);
                compute_sync();
                PROF(4);
                PROF(5);
            } else {
                /* P2: attention partials */
                cta_attention(a, lm, plan, l, pos, att_sm, kvs, red);
                PROF(4);
                grid_barrier(a.bar, target, ncta);
                PROF(5);
                /* P3: combine partials -> O projection + residual */
                if (plan.S == 1) stage_vec(sm, a.att, lm.heads * lm.head_dim);
                else stage_attention_out(sm, att_sm, kvs, a, lm, plan.S);
            }
            xa_build<false>(xa, sm, lm.heads * lm.head_dim);
            PROF(15);
            phase_seg(a, lm, l, 1, s, cols);
            pg = phase_geom(s, cols, 0, cta, ncta);
            rp = ring_consume(rg, rp, pg, xa);
            ring_finalize(rg, pg, s, a.x);
            PROF(6);
            grid_barrier(a.bar, target, ncta);
            PROF(7);
            /* P4: RMSNorm -> gate/up -> SwiGLU */
            stage_rmsnorm(sm, a.x, w.post_norm, H, lm.eps, red);
            xa_build<false>(xa, sm, H);
            PROF(8);
            phase_seg(a, lm, l, 2, s, cols);
            pg = phase_geom(s, cols, 1, cta, ncta);
            rp = ring_consume(rg, rp, pg, xa);
            ring_finalize(rg, pg, s, NULL);
            PROF(9);
            grid_barrier(a.bar, target, ncta);
            PROF(10);
            /* P5: down projection + residual; the K-slices of g go straight from L2 to registers */
            xa_build<true>(xa, a.g, lm.inter);
            PROF(11);
            phase_seg(a, lm, l, 3, s, cols);
            pg = phase_geom(s, cols, 0, cta, ncta);
            rp = ring_consume(rg, rp, pg, xa);
            ring_finalize(rg, pg, s, a.x);
            PROF(12);
            grid_barrier(a.bar, target, ncta);
            PROF(13);
        }
#undef PROF
        /* final norm (every CTA, into shared memory), then x is overwritten in place with the normed hidden */
        stage_rmsnorm(sm, a.x, lm.final_norm, H, lm.eps, red);
        const uint16_t *hw = ring_head_of(a, sp, g);
        if (sp.G == 0) {
            grid_barrier(a.bar, target, ncta);          /* everyone has read x before CTA 0 replaces it */
            if (cta == 0) {
                for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = sm[i];
                if (tid == 0 && a.st && !a.st->done) a.st->pos += 1;
            }
        }
        if (hw) {
            s.n = 1; s.w[0] = hw; s.out[0] = a.logits; s.rows[0] = a.head_rows;
            xa_build<false>(xa, sm, H);
            pg = phase_geom(s, H, 0, cta, ncta);
            rp = ring_consume(rg, rp, pg, xa);
            ring_finalize(rg, pg, s, NULL);
        }
        if (sp.G > 0) {
            /* (inside the sub-talker loop nobody reads the hidden state back.)  Every CTA samples for itself
             * from the same logits: same code everywhere, no broadcast needed */
            grid_barrier(a.bar, target, ncta);
            if (g >= 1) {
                const int V = sp.vocab;
                float *lg = kvs, *vv = kvs + V;
                SampleSmem &S = *reinterpret_cast<SampleSmem *>(kvs + 2 * ((V + 3) & ~3));
                for (int i = tid; i < V; i += FUSED_THREADS) lg[i] = __ldcg(a.logits + i);
                if (g == 1 && tid == 0) s_rng = sp.seed_bits;
                compute_sync();
                const int tok = block_sample_t<FUSED_THREADS>(lg, vv, V, sp.cfg.top_k, sp.cfg.top_p, sp.cfg.temperature, &s_rng, S,
                                                              sp.scratch_f, sp.scratch_i);
                if (tid == 0) { s_codes[g] = tok; if (cta == 0) sp.codes[g] = tok; }
            }
        }
    }
}
