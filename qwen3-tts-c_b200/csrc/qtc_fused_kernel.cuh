/*
 * qtc_fused_kernel.cuh - attention phase and main loop of the fused decode-step kernel (see qtc_fused.cuh).
 *
 * Latency engineering of the attention phase (it moves no weights, so every microsecond of it is exposed):
 *   - the KV rows a CTA will need (positions < pos, written by earlier steps) are copied into shared memory
 *     with cp.async BEFORE the grid barrier that ends the QKV phase, so they land while the barrier spins;
 *   - each (head, KV-range) CTA emits an un-normalised partial (max, sum, out[hd]); the O-projection phase
 *     folds the partials while it stages its input vector - no counter, no fence, no second pass here.
 */
#pragma once
#include "qtc_fused.cuh"

#define ATT_STAGE_MAX 64          /* positions per CTA that fit the shared-memory KV stage */

struct AttPlan { int S, h, sp, g, t0, t1, active, staged; };

__device__ __forceinline__ AttPlan att_plan(const LmDev &lm, int pos, int cta, int ncta) {
    AttPlan p;
    const int n = pos + 1;
    p.S = (n + 31) / 32;
    if (p.S > FUSED_MAX_SPLIT) p.S = FUSED_MAX_SPLIT;
    while (p.S > 1 && lm.heads * p.S > ncta) p.S--;
    p.active = cta < lm.heads * p.S;
    p.h = p.active ? cta / p.S : 0;
    p.sp = cta - p.h * p.S;
    p.g = p.h / (lm.heads / lm.kv_heads);
    const int chunk = (n + p.S - 1) / p.S;
    p.t0 = p.sp * chunk;
    p.t1 = p.t0 + chunk < n ? p.t0 + chunk : n;
    if (p.t1 < p.t0) p.t1 = p.t0;
    p.staged = chunk <= ATT_STAGE_MAX && lm.head_dim % 4 == 0;
    return p;
}

/* issue the cp.async copies of this CTA's K and V rows (t0 <= t < min(t1,pos)) into kvs = [2][ATT_STAGE_MAX][hd] */
__device__ __forceinline__ void att_stage_issue(const FusedArgs &a, const LmDev &lm, const AttPlan &p, int l, int pos, float *kvs) {
    if (!p.active || !p.staged) return;
    const int hd = lm.head_dim, kvd = lm.kv_heads * hd, q4 = hd / 4;
    const float *kc = a.kc + (size_t)l * a.kv_max * kvd + p.g * hd, *vc = a.vc + (size_t)l * a.kv_max * kvd + p.g * hd;
    const int te = p.t1 < pos ? p.t1 : pos, rows = te - p.t0;
    for (int i = threadIdx.x; i < rows * q4 * 2; i += FUSED_THREADS) {
        const int which = i >= rows * q4, j = which ? i - rows * q4 : i, r = j / q4, c = j - r * q4;
        const float *src = (which ? vc : kc) + (size_t)(p.t0 + r) * kvd + c * 4;
        float *dst = kvs + (size_t)which * ATT_STAGE_MAX * hd + r * hd + c * 4;
        const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src));
    }
    asm volatile("cp.async.commit_group;");
}

/* per-head QK-RMSNorm + RoPE + KV append + partial attention over [t0,t1).  sm: scratch, kvs: staged K/V */
__device__ __forceinline__ void cta_attention(const FusedArgs &a, const LmDev &lm, const AttPlan &p, int l, int pos, float *sm, float *kvs,
                                              float *red) {
    asm volatile("cp.async.wait_all;");
    if (!p.active) return;
    const int hd = lm.head_dim, half = hd >> 1, rep = lm.heads / lm.kv_heads, kvd = lm.kv_heads * hd;
    const int h = p.h, g = p.g, t0 = p.t0, t1 = p.t1;
    const int ngrp = FUSED_THREADS / hd > 0 ? FUSED_THREADS / hd : 1;
    float *qs = sm, *ks = sm + hd, *op = sm + 2 * hd, *sc = op + ngrp * hd;     /* sc: chunk floats */
    const LmLayerDev &w = lm.layer[l];
    const float *qr = a.q + h * hd, *kr = a.kk + g * hd, *vr = a.vv + g * hd;
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;
    float *kc = a.kc + (size_t)l * a.kv_max * kvd, *vc = a.vc + (size_t)l * a.kv_max * kvd;
    const float *Ks = kvs, *Vs = kvs + (size_t)ATT_STAGE_MAX * hd;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    /* q and k of this token: norm over the head, rotate-half.  One warp each (hd <= 512), no block barrier */
    if (warp < 2) {
        const float *src = warp == 0 ? qr : kr, *nw = warp == 0 ? w.q_norm : w.k_norm;
        float *dst = warp == 0 ? qs : ks;
        float ss = 0.0f;
        for (int i = lane; i < hd; i += 32) { const float v = __ldcg(src + i); dst[i] = v; ss = fmaf(v, v, ss); }
        ss = warp_sum(ss);
        const float inv = 1.0f / sqrtf(ss / (float)hd + lm.eps);
        __syncwarp();
        for (int i = lane; i < half; i += 32) {
            const float xa = dst[i] * inv * nw[i], xb = dst[i + half] * inv * nw[i + half];
            dst[i] = xa * cs[i] - xb * sn[i];
            dst[i + half] = xb * cs[i + half] + xa * sn[i + half];
        }
    }
    __syncthreads();
    if (p.sp == 0 && h % rep == 0)
        for (int i = tid; i < hd; i += FUSED_THREADS) { kc[(size_t)pos * kvd + g * hd + i] = ks[i]; vc[(size_t)pos * kvd + g * hd + i] = __ldcg(vr + i); }

    /* scores of my KV range: a warp per position */
    const float scale = 1.0f / sqrtf((float)hd);
    if (hd == 128) {
        const float4 q4 = *reinterpret_cast<const float4 *>(qs + lane * 4);
        for (int tb = t0 + warp; tb < t1; tb += 4 * FUSED_WARPS) {
            float4 kv[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int t = tb + j * FUSED_WARPS;
                if (t < t1 && t < pos)
                    kv[j] = p.staged ? *reinterpret_cast<const float4 *>(Ks + (size_t)(t - t0) * hd + lane * 4)
                                     : __ldcg(reinterpret_cast<const float4 *>(kc + (size_t)t * kvd + g * hd) + lane);
                else kv[j] = *reinterpret_cast<const float4 *>(ks + lane * 4);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int t = tb + j * FUSED_WARPS;
                float d = q4.x * kv[j].x;
                d = fmaf(q4.y, kv[j].y, d); d = fmaf(q4.z, kv[j].z, d); d = fmaf(q4.w, kv[j].w, d);
                d = warp_sum(d);
                if (lane == 0 && t < t1) sc[t - t0] = d * scale;
            }
        }
    } else {
        for (int t = t0 + warp; t < t1; t += FUSED_WARPS) {
            float d = 0.0f;
            if (t < pos) {
                const float *kt = p.staged ? Ks + (size_t)(t - t0) * hd : kc + (size_t)t * kvd + g * hd;
                for (int i = lane; i < hd; i += 32) d = fmaf(qs[i], p.staged ? kt[i] : __ldcg(kt + i), d);
            } else for (int i = lane; i < hd; i += 32) d = fmaf(qs[i], ks[i], d);
            d = warp_sum(d);
            if (lane == 0) sc[t - t0] = d * scale;
        }
    }
    __syncthreads();
    float mx = -3.402823466e+38f;
    for (int t = tid; t < t1 - t0; t += FUSED_THREADS) mx = fmaxf(mx, sc[t]);
    mx = block_max(mx, red);
    float tot = 0.0f;
    for (int t = tid; t < t1 - t0; t += FUSED_THREADS) { const float e = __expf(sc[t] - mx); sc[t] = e; tot += e; }
    tot = block_sum(tot, red);
    /* weighted V: thread groups of hd lanes take positions round-robin, then fold the groups */
    const int grp = tid / hd, dim = tid - grp * hd;
    if (grp < ngrp) {
        float o = 0.0f;
        for (int tb = t0 + grp; tb < t1; tb += 8 * ngrp) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int t = tb + j * ngrp;
                v[j] = 0.0f;
                if (t < t1) v[j] = t < pos ? (p.staged ? Vs[(size_t)(t - t0) * hd + dim] : __ldcg(vc + (size_t)t * kvd + g * hd + dim)) : __ldcg(vr + dim);
            }
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int t = tb + j * ngrp;
                if (t < t1) o = fmaf(sc[t - t0], v[j], o);
            }
        }
        op[grp * hd + dim] = o;
    }
    __syncthreads();
    float *mine = a.part + ((size_t)h * FUSED_MAX_SPLIT + p.sp) * (hd + 2);
    if (tid < hd) { float o = 0.0f; for (int k = 0; k < ngrp; k++) o += op[k * hd + tid]; mine[2 + tid] = o; }
    if (tid == 0) { mine[0] = mx; mine[1] = tot; }
}

/* xs[h*hd + i] = softmax-combine of the S partials of head h: the O projection's input vector */
__device__ __forceinline__ void stage_attention_out(float *xs, float *scr, const FusedArgs &a, const LmDev &lm, int S) {
    const int hd = lm.head_dim, heads = lm.heads, qd = heads * hd, tid = threadIdx.x;
    float *mb = scr, *lb = scr + heads * FUSED_MAX_SPLIT, *fac = lb + heads * FUSED_MAX_SPLIT;
    /* every load of the phase is independent of every other: issue them all, then fold */
    float v[4][FUSED_MAX_SPLIT];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int i = tid + r * FUSED_THREADS, h = i / hd, d = i - h * hd;
#pragma unroll
        for (int k = 0; k < FUSED_MAX_SPLIT; k++)
            v[r][k] = (i < qd && k < S) ? __ldcg(a.part + ((size_t)h * FUSED_MAX_SPLIT + k) * (hd + 2) + 2 + d) : 0.0f;
    }
    for (int idx = tid; idx < heads * S; idx += FUSED_THREADS) {
        const int h = idx / S, k = idx - h * S;
        const float *pp = a.part + ((size_t)h * FUSED_MAX_SPLIT + k) * (hd + 2);
        mb[h * FUSED_MAX_SPLIT + k] = __ldcg(pp);
        lb[h * FUSED_MAX_SPLIT + k] = __ldcg(pp + 1);
    }
    __syncthreads();
    for (int h = tid; h < heads; h += FUSED_THREADS) {
        float M = -3.402823466e+38f, L = 0.0f;
        for (int k = 0; k < S; k++) M = fmaxf(M, mb[h * FUSED_MAX_SPLIT + k]);
        for (int k = 0; k < S; k++) L = fmaf(lb[h * FUSED_MAX_SPLIT + k], __expf(mb[h * FUSED_MAX_SPLIT + k] - M), L);
        for (int k = 0; k < S; k++) fac[h * FUSED_MAX_SPLIT + k] = __expf(mb[h * FUSED_MAX_SPLIT + k] - M) / L;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int i = tid + r * FUSED_THREADS, h = i / hd;
        if (i < qd) {
            float acc = 0.0f;
#pragma unroll
            for (int k = 0; k < FUSED_MAX_SPLIT; k++) if (k < S) acc = fmaf(v[r][k], fac[h * FUSED_MAX_SPLIT + k], acc);
            xs[i] = acc;
        }
    }
    for (int i = tid + 4 * FUSED_THREADS; i < qd; i += FUSED_THREADS) {          /* qd > 2048: plain loop */
        const int h = i / hd, d = i - h * hd;
        float acc = 0.0f;
        for (int k = 0; k < S; k++) acc = fmaf(__ldcg(a.part + ((size_t)h * FUSED_MAX_SPLIT + k) * (hd + 2) + 2 + d), fac[h * FUSED_MAX_SPLIT + k], acc);
        xs[i] = acc;
    }
    __syncthreads();
}

/* dynamic smem: [xs_floats | attention scratch | KV stage 2*ATT_STAGE_MAX*hd] floats */
__global__ void __launch_bounds__(FUSED_THREADS, 1) k_lm_fused(const FusedArgs a) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[32];
    const LmDev &lm = *a.lm;
    const int cta = blockIdx.x, ncta = gridDim.x, tid = threadIdx.x;
    const int H = lm.hidden, I = lm.inter, L = lm.layers;
    const int pos = *a.pos_ptr + a.pos_add;
    float *att_sm = sm + a.xs_floats, *kvs = att_sm + a.att_floats;
    const AttPlan plan = att_plan(lm, pos, cta, ncta);
    unsigned long long target = 0;
    if (tid == 0) { const unsigned long long v = ld_acquire_u64(a.bar); target = v - v % (unsigned long long)ncta; }
    prefetch_phase(a, lm, 0, cta, ncta);
    prefetch_phase(a, lm, 1, cta, ncta);
    Seg3 s; int cols;
#define PROF(i) do { if (a.prof && tid == 0 && l == a.prof_layer) a.prof[cta * 16 + (i)] = clock64(); } while (0)
    for (int l = 0; l < L; l++) {
        const LmLayerDev &w = lm.layer[l];
        /* P1: RMSNorm -> Q,K,V */
        PROF(0);
        stage_rmsnorm(sm, a.x, w.in_norm, H, lm.eps, red);
        PROF(1);
        phase_seg(a, lm, l, 0, s, cols);
        cta_gemv(s, cols, sm, NULL, cta, ncta);
        att_stage_issue(a, lm, plan, l, pos, kvs);
        PROF(2);
        barrier_arrive(a.bar, target, ncta);
        prefetch_phase(a, lm, l * 4 + 2, cta, ncta);          /* issued while the barrier spins */
        barrier_wait(a.bar, target);
        PROF(3);
        /* P2: attention partials */
        cta_attention(a, lm, plan, l, pos, att_sm, kvs, red);
        PROF(4);
        barrier_arrive(a.bar, target, ncta);
        prefetch_phase(a, lm, l * 4 + 3, cta, ncta);
        barrier_wait(a.bar, target);
        PROF(5);
        /* P3: combine partials -> O projection + residual */
        stage_attention_out(sm, att_sm, a, lm, plan.S);
        PROF(15);
        phase_seg(a, lm, l, 1, s, cols);
        cta_gemv(s, cols, sm, a.x, cta, ncta);
        PROF(6);
        barrier_arrive(a.bar, target, ncta);
        prefetch_phase(a, lm, l * 4 + 4, cta, ncta);
        barrier_wait(a.bar, target);
        PROF(7);
        /* P4: RMSNorm -> gate/up -> SwiGLU */
        stage_rmsnorm(sm, a.x, w.post_norm, H, lm.eps, red);
        PROF(8);
        cta_swiglu(w.gate, w.up, a.g, I, H, sm, cta, ncta);
        PROF(9);
        barrier_arrive(a.bar, target, ncta);
        prefetch_phase(a, lm, l * 4 + 5, cta, ncta);
        barrier_wait(a.bar, target);
        PROF(10);
        /* P5: down projection + residual */
        stage_vec(sm, a.g, I);
        PROF(11);
        phase_seg(a, lm, l, 3, s, cols);
        cta_gemv(s, cols, sm, a.x, cta, ncta);
        PROF(12);
        grid_barrier(a.bar, target, ncta);
        PROF(13);
    }
#undef PROF
    /* final norm (every CTA, into shared memory), then x is overwritten in place with the normed hidden */
    stage_rmsnorm(sm, a.x, lm.final_norm, H, lm.eps, red);
    grid_barrier(a.bar, target, ncta);
    if (cta == 0) {
        for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = sm[i];
        if (tid == 0 && a.st && !a.st->done) a.st->pos += 1;
    }
    if (phase_seg(a, lm, L, 0, s, cols)) cta_gemv(s, cols, sm, NULL, cta, ncta);
}
