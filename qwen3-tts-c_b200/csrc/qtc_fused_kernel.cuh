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

struct AttPlan { int S, h, sp, g, t0, t1, active, staged, cap; };

/* land_words > 0 (mode 3): the partial records of all heads must fit a landing zone of that many 8-byte words */
__device__ __forceinline__ AttPlan att_plan(const LmDev &lm, int pos, int cta, int ncta, int cap = ATT_STAGE_MAX, int land_words = 0, int per = 32) {
    AttPlan p;
    p.cap = cap;
    const int n = pos + 1;
    p.S = (n + per - 1) / per;
    if (p.S > FUSED_MAX_SPLIT) p.S = FUSED_MAX_SPLIT;
    while (p.S > 1 && (lm.heads * p.S > ncta || lm.heads * p.S * (lm.head_dim + 4) > (land_words > 0 ? land_words : 2 * cap * lm.head_dim))) p.S--;
    p.active = cta < lm.heads * p.S;
    p.h = p.active ? cta / p.S : 0;
    p.sp = cta - p.h * p.S;
    p.g = p.h / (lm.heads / lm.kv_heads);
    const int chunk = (n + p.S - 1) / p.S;
    p.t0 = p.sp * chunk;
    p.t1 = p.t0 + chunk < n ? p.t0 + chunk : n;
    if (p.t1 < p.t0) p.t1 = p.t0;
    p.staged = chunk <= cap && lm.head_dim % 4 == 0;
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
        const float *src = (which ? vc : kc) + kv_off(a.table, p.t0 + r, kvd) + c * 4;
        float *dst = kvs + (size_t)which * p.cap * hd + r * hd + c * 4;
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
    const float *Ks = kvs, *Vs = kvs + (size_t)p.cap * hd;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    /* q and k of this token: norm over the head, rotate-half.  One warp each (hd <= 512), no block barrier */
    if (warp < 2) {
        const float *src = warp == 0 ? qr : kr, *nw = warp == 0 ? w.q_norm : w.k_norm;
        float *dst = warp == 0 ? qs : ks;
        float ss = 0.0f;
        if (hd == 128) {
            /* all twelve loads of a lane are independent: issue them before the reduction */
            float xa[2], xb[2], wa[2], wb[2], ca[2], cb[2], sa[2], sb[2];
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int i = lane + 32 * j;
                xa[j] = __ldcg(src + i); xb[j] = __ldcg(src + i + 64);
                wa[j] = nw[i]; wb[j] = nw[i + 64]; ca[j] = cs[i]; cb[j] = cs[i + 64]; sa[j] = sn[i]; sb[j] = sn[i + 64];
                ss = fmaf(xa[j], xa[j], ss); ss = fmaf(xb[j], xb[j], ss);
            }
            ss = warp_sum(ss);
            const float inv = 1.0f / sqrtf(ss / (float)hd + lm.eps);
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int i = lane + 32 * j;
                const float ya = xa[j] * inv * wa[j], yb = xb[j] * inv * wb[j];
                dst[i] = ya * ca[j] - yb * sa[j];
                dst[i + 64] = yb * cb[j] + ya * sb[j];
            }
        } else {
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
    }
    compute_sync();
    if (p.sp == 0 && h % rep == 0)
        for (int i = tid; i < hd; i += FUSED_THREADS) { kc[kv_off(a.table, pos, kvd) + g * hd + i] = ks[i]; vc[kv_off(a.table, pos, kvd) + g * hd + i] = __ldcg(vr + i); }

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
                                     : __ldcg(reinterpret_cast<const float4 *>(kc + kv_off(a.table, t, kvd) + g * hd) + lane);
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
                const float *kt = p.staged ? Ks + (size_t)(t - t0) * hd : kc + kv_off(a.table, t, kvd) + g * hd;
                for (int i = lane; i < hd; i += 32) d = fmaf(qs[i], p.staged ? kt[i] : __ldcg(kt + i), d);
            } else for (int i = lane; i < hd; i += 32) d = fmaf(qs[i], ks[i], d);
            d = warp_sum(d);
            if (lane == 0) sc[t - t0] = d * scale;
        }
    }
    compute_sync();
    float mx = -3.402823466e+38f, tot = 0.0f;
    const bool small = t1 - t0 <= 64;
    if (small) {
        /* <= 64 scores: every warp folds them for itself with shuffles - no block barriers on the critical path */
        const int n0 = t1 - t0;
        const float v0 = lane < n0 ? sc[lane] : -3.402823466e+38f, v1 = lane + 32 < n0 ? sc[lane + 32] : -3.402823466e+38f;
        mx = warp_max(fmaxf(v0, v1));
        tot = warp_sum((lane < n0 ? __expf(v0 - mx) : 0.0f) + (lane + 32 < n0 ? __expf(v1 - mx) : 0.0f));
    } else {
        for (int t = tid; t < t1 - t0; t += FUSED_THREADS) mx = fmaxf(mx, sc[t]);
        mx = cmax(mx, red);
        for (int t = tid; t < t1 - t0; t += FUSED_THREADS) { const float e = __expf(sc[t] - mx); sc[t] = e; tot += e; }
        tot = csum(tot, red);
    }
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
                if (t < t1) v[j] = t < pos ? (p.staged ? Vs[(size_t)(t - t0) * hd + dim] : __ldcg(vc + kv_off(a.table, t, kvd) + g * hd + dim)) : __ldcg(vr + dim);
            }
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int t = tb + j * ngrp;
                if (t < t1) o = fmaf(small ? __expf(sc[t - t0] - mx) : sc[t - t0], v[j], o);
            }
        }
        op[grp * hd + dim] = o;
    }
    compute_sync();
    if (p.S == 1) {                          /* no split: the head's output is final */
        if (tid < hd) { float o = 0.0f; for (int k = 0; k < ngrp; k++) o += op[k * hd + tid]; a.att[h * hd + tid] = o / tot; }
        return;
    }
    /* partial record: [max, sum, -, -, out[hd]]; records of one head are contiguous and 16-byte aligned */
    float *mine = a.part + ((size_t)h * p.S + p.sp) * (hd + 4);
    if (tid < hd) { float o = 0.0f; for (int k = 0; k < ngrp; k++) o += op[k * hd + tid]; mine[4 + tid] = o; }
    if (tid == 0) { mine[0] = mx; mine[1] = tot; }
}

/* Short contexts (every sub-talker step: n = pos + 1 <= 32, head_dim 128, one CTA per head, rows staged): the whole head in ONE
 * warp - lane i owns dims 4i..4i+3, q/k norm and RoPE through shuffles (the rotate-half partner of dim d is lane ^ 16), scores
 * eight rows at a time with interleaved butterfly reductions, lane t keeps score t, V rows weighted by shuffled probabilities.
 * No block barrier after the staged rows are visible; the step's latency chain is one L2 round trip plus ~1000 cycles. */
__device__ __forceinline__ void cta_attention_small(const FusedArgs &a, const LmDev &lm, const AttPlan &p, int l, int pos, const float *kvs) {
    asm volatile("cp.async.wait_all;" ::: "memory");
    if (!p.active) return;
    compute_sync();                                   /* the staged K/V rows were copied by every thread of the CTA */
    if (threadIdx.x >= 32) return;
    const int hd = 128, lane = threadIdx.x, h = p.h, g = p.g, rep = lm.heads / lm.kv_heads, kvd = lm.kv_heads * hd, n = pos + 1;
    const LmLayerDev &w = lm.layer[l];
    const float4 qv = __ldcg(reinterpret_cast<const float4 *>(a.q + h * hd) + lane);
    const float4 kv = __ldcg(reinterpret_cast<const float4 *>(a.kk + g * hd) + lane);
    const float4 v_new = __ldcg(reinterpret_cast<const float4 *>(a.vv + g * hd) + lane);
    const float4 qn = reinterpret_cast<const float4 *>(w.q_norm)[lane], kn = reinterpret_cast<const float4 *>(w.k_norm)[lane];
    const float4 c4 = reinterpret_cast<const float4 *>(a.cos + (size_t)pos * hd)[lane], s4 = reinterpret_cast<const float4 *>(a.sin + (size_t)pos * hd)[lane];
    float sq = qv.x * qv.x, sk = kv.x * kv.x;
    sq = fmaf(qv.y, qv.y, sq); sq = fmaf(qv.z, qv.z, sq); sq = fmaf(qv.w, qv.w, sq);
    sk = fmaf(kv.y, kv.y, sk); sk = fmaf(kv.z, kv.z, sk); sk = fmaf(kv.w, kv.w, sk);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { sq += __shfl_xor_sync(0xffffffffu, sq, o); sk += __shfl_xor_sync(0xffffffffu, sk, o); }
    const float iq = 1.0f / sqrtf(sq / (float)hd + lm.eps), ik = 1.0f / sqrtf(sk / (float)hd + lm.eps);
    const float4 qy = make_float4(qv.x * iq * qn.x, qv.y * iq * qn.y, qv.z * iq * qn.z, qv.w * iq * qn.w);
    const float4 ky = make_float4(kv.x * ik * kn.x, kv.y * ik * kn.y, kv.z * ik * kn.z, kv.w * ik * kn.w);
    const float sgn = lane < 16 ? -1.0f : 1.0f;       /* first half: y*cos - partner*sin; second half: y*cos + partner*sin */
    float4 q4, k_new;
    {
        const float px = __shfl_xor_sync(0xffffffffu, qy.x, 16), py = __shfl_xor_sync(0xffffffffu, qy.y, 16);
        const float pz = __shfl_xor_sync(0xffffffffu, qy.z, 16), pw = __shfl_xor_sync(0xffffffffu, qy.w, 16);
        q4 = lane < 16 ? make_float4(qy.x * c4.x - px * s4.x, qy.y * c4.y - py * s4.y, qy.z * c4.z - pz * s4.z, qy.w * c4.w - pw * s4.w)
                       : make_float4(qy.x * c4.x + px * s4.x, qy.y * c4.y + py * s4.y, qy.z * c4.z + pz * s4.z, qy.w * c4.w + pw * s4.w);
    }
    {
        const float px = __shfl_xor_sync(0xffffffffu, ky.x, 16), py = __shfl_xor_sync(0xffffffffu, ky.y, 16);
        const float pz = __shfl_xor_sync(0xffffffffu, ky.z, 16), pw = __shfl_xor_sync(0xffffffffu, ky.w, 16);
        k_new = lane < 16 ? make_float4(ky.x * c4.x - px * s4.x, ky.y * c4.y - py * s4.y, ky.z * c4.z - pz * s4.z, ky.w * c4.w - pw * s4.w)
                          : make_float4(ky.x * c4.x + px * s4.x, ky.y * c4.y + py * s4.y, ky.z * c4.z + pz * s4.z, ky.w * c4.w + pw * s4.w);
    }
    (void)sgn;
    if (h % rep == 0) {
        float *kc = a.kc + (size_t)l * a.kv_max * kvd, *vc = a.vc + (size_t)l * a.kv_max * kvd;
        *reinterpret_cast<float4 *>(kc + kv_off(a.table, pos, kvd) + g * hd + lane * 4) = k_new;
        *reinterpret_cast<float4 *>(vc + kv_off(a.table, pos, kvd) + g * hd + lane * 4) = v_new;
    }
    const float *Ks = kvs, *Vs = kvs + (size_t)p.cap * hd;
    const float scale = 1.0f / sqrtf((float)hd);
    float sc = -3.402823466e+38f;
    for (int t0 = 0; t0 < n; t0 += 8) {
        float d[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int t = t0 + u;
            float4 kr = make_float4(0, 0, 0, 0);
            if (t < pos) kr = *reinterpret_cast<const float4 *>(Ks + (size_t)t * hd + lane * 4);
            else if (t == pos) kr = k_new;
            d[u] = q4.x * kr.x;
            d[u] = fmaf(q4.y, kr.y, d[u]); d[u] = fmaf(q4.z, kr.z, d[u]); d[u] = fmaf(q4.w, kr.w, d[u]);
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
    for (int t = 0; t < n; t++) {
        const float pr = __shfl_sync(0xffffffffu, e, t);
        const float4 vr = t < pos ? *reinterpret_cast<const float4 *>(Vs + (size_t)t * hd + lane * 4) : v_new;
        o4.x = fmaf(pr, vr.x, o4.x); o4.y = fmaf(pr, vr.y, o4.y); o4.z = fmaf(pr, vr.z, o4.z); o4.w = fmaf(pr, vr.w, o4.w);
    }
    *reinterpret_cast<float4 *>(a.att + h * hd + lane * 4) = make_float4(o4.x / tot, o4.y / tot, o4.z / tot, o4.w / tot);
}

/* Split-KV attention with the block barriers taken out of the chain (head_dim 128, staged range of <= 64 positions): every warp
 * norms and rotates q for itself (lane i owns dims 4i..4i+3, rotate-half partner = lane ^ 16), takes up to four positions of the
 * CTA's range round-robin, reduces their four scores with interleaved butterflies, and keeps an un-normalised (max, sum, out)
 * triple in registers; ONE block barrier later 128 threads fold the 16 triples into the CTA's partial record
 * [max, sum, -, -, out[hd]] - also when the head is not split (the O projection's staging normalises).
 * sm: [wm 16 | wl 16 | wo 16 x hd] */
__device__ __forceinline__ void cta_attention_warps(const FusedArgs &a, const LmDev &lm, const AttPlan &p, int l, int pos, float *sm, const float *kvs) {
    asm volatile("cp.async.wait_all;" ::: "memory");
    if (!p.active) return;
    compute_sync();                                   /* the staged K/V rows were copied by every thread of the CTA */
    const int hd = 128, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int h = p.h, g = p.g, rep = lm.heads / lm.kv_heads, kvd = lm.kv_heads * hd, t0 = p.t0, t1 = p.t1;
    const LmLayerDev &w = lm.layer[l];
    float *wm = sm, *wl = sm + FUSED_WARPS, *wo = sm + 2 * FUSED_WARPS;
    const bool writer = p.sp == 0 && h % rep == 0 && warp == FUSED_WARPS - 1;
    const bool has_new = t1 == pos + 1 && ((pos - t0 - warp) % FUSED_WARPS == 0) && pos - t0 >= warp;      /* this warp's positions include t == pos */
    const float4 c4 = reinterpret_cast<const float4 *>(a.cos + (size_t)pos * hd)[lane], s4 = reinterpret_cast<const float4 *>(a.sin + (size_t)pos * hd)[lane];
    auto norm_rope = [&](const float *src, const float *nw) -> float4 {
        const float4 v = __ldcg(reinterpret_cast<const float4 *>(src) + lane), n4 = reinterpret_cast<const float4 *>(nw)[lane];
        float ss = v.x * v.x;
        ss = fmaf(v.y, v.y, ss); ss = fmaf(v.z, v.z, ss); ss = fmaf(v.w, v.w, ss);
        ss = warp_sum(ss);
        const float inv = 1.0f / sqrtf(ss / (float)hd + lm.eps);
        const float4 y = make_float4(v.x * inv * n4.x, v.y * inv * n4.y, v.z * inv * n4.z, v.w * inv * n4.w);
        const float px = __shfl_xor_sync(0xffffffffu, y.x, 16), py = __shfl_xor_sync(0xffffffffu, y.y, 16);
        const float pz = __shfl_xor_sync(0xffffffffu, y.z, 16), pw = __shfl_xor_sync(0xffffffffu, y.w, 16);
        return lane < 16 ? make_float4(y.x * c4.x - px * s4.x, y.y * c4.y - py * s4.y, y.z * c4.z - pz * s4.z, y.w * c4.w - pw * s4.w)
                         : make_float4(y.x * c4.x + px * s4.x, y.y * c4.y + py * s4.y, y.z * c4.z + pz * s4.z, y.w * c4.w + pw * s4.w);
    };
    const float4 q4 = norm_rope(a.q + h * hd, w.q_norm);
    float4 k_new = make_float4(0, 0, 0, 0), v_new = k_new;
    if (has_new || writer) {
        k_new = norm_rope(a.kk + g * hd, w.k_norm);
        v_new = __ldcg(reinterpret_cast<const float4 *>(a.vv + g * hd) + lane);
    }
    if (writer) {
        float *kc = a.kc + (size_t)l * a.kv_max * kvd, *vc = a.vc + (size_t)l * a.kv_max * kvd;
        *reinterpret_cast<float4 *>(kc + kv_off(a.table, pos, kvd) + g * hd + lane * 4) = k_new;
        *reinterpret_cast<float4 *>(vc + kv_off(a.table, pos, kvd) + g * hd + lane * 4) = v_new;
    }
    const float *Ks = kvs, *Vs = kvs + (size_t)p.cap * hd;
    const float scale = 1.0f / sqrtf((float)hd);
    float d[4];
    float4 vr[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int t = t0 + warp + FUSED_WARPS * j;
        float4 kr = make_float4(0, 0, 0, 0);
        vr[j] = kr;
        if (t < t1) {
            if (t < pos) { kr = *reinterpret_cast<const float4 *>(Ks + (size_t)(t - t0) * hd + lane * 4); vr[j] = *reinterpret_cast<const float4 *>(Vs + (size_t)(t - t0) * hd + lane * 4); }
            else { kr = k_new; vr[j] = v_new; }
        }
        d[j] = q4.x * kr.x;
        d[j] = fmaf(q4.y, kr.y, d[j]); d[j] = fmaf(q4.z, kr.z, d[j]); d[j] = fmaf(q4.w, kr.w, d[j]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float t4[4];
#pragma unroll
        for (int j = 0; j < 4; j++) t4[j] = __shfl_xor_sync(0xffffffffu, d[j], o);
#pragma unroll
        for (int j = 0; j < 4; j++) d[j] += t4[j];
    }
    float mx = -3.402823466e+38f;
#pragma unroll
    for (int j = 0; j < 4; j++) if (t0 + warp + FUSED_WARPS * j < t1) { d[j] *= scale; mx = fmaxf(mx, d[j]); }
    float tot = 0.0f;
    float4 o4 = make_float4(0, 0, 0, 0);
#pragma unroll
    for (int j = 0; j < 4; j++)
        if (t0 + warp + FUSED_WARPS * j < t1) {
            const float e = __expf(d[j] - mx);
            tot += e;
            o4.x = fmaf(e, vr[j].x, o4.x); o4.y = fmaf(e, vr[j].y, o4.y); o4.z = fmaf(e, vr[j].z, o4.z); o4.w = fmaf(e, vr[j].w, o4.w);
        }
    if (lane == 0) { wm[warp] = mx; wl[warp] = tot; }
    *reinterpret_cast<float4 *>(wo + warp * hd + lane * 4) = o4;
    compute_sync();
    if (tid < hd) {
        /* lanes k < 16 hold the warps' (max, sum); factors by butterfly, broadcast by shuffle */
        const float m = lane < FUSED_WARPS ? wm[lane] : -3.402823466e+38f, lw = lane < FUSED_WARPS ? wl[lane] : 0.0f;
        const float M = warp_max(m);
        const float f = lw > 0.0f ? __expf(m - M) : 0.0f;
        const float L = warp_sum(lw * f);
        float o = 0.0f;
#pragma unroll
        for (int k = 0; k < FUSED_WARPS; k++) o = fmaf(wo[k * hd + tid], __shfl_sync(0xffffffffu, f, k), o);
        float *mine = a.part + ((size_t)h * p.S + p.sp) * (hd + 4);
        mine[4 + tid] = o;
        if (tid == 0) { mine[0] = M; mine[1] = L; }
    }
}

/* xs[h*hd + i] = softmax-combine of the S partials of head h: the O projection's input vector.
 * The heads*S records (contiguous in HBM/L2) are copied into `land` (the KV stage, idle in this phase) with
 * cp.async - one round trip, no registers tied up - and folded from shared memory. */
__device__ __forceinline__ void stage_attention_out(float *xs, float *scr, float *land, const FusedArgs &a, const LmDev &lm, int S, long long *pp = NULL) {
    const int hd = lm.head_dim, heads = lm.heads, tid = threadIdx.x, rec = hd + 4;
    const int n16 = heads * S * rec / 4;
    for (int i = tid; i < n16; i += FUSED_THREADS) {
        const uint32_t sa = (uint32_t)__cvta_generic_to_shared(land + i * 4);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(a.part + i * 4));
    }
    asm volatile("cp.async.commit_group;");
    if (pp) pp[20] = clock64();
    asm volatile("cp.async.wait_all;" ::: "memory");
    if (pp) pp[21] = clock64();
    compute_sync();
    if (pp) pp[22] = clock64();
    /* a warp per head: lanes k < S hold the S (max, sum) pairs, the fold factors exp(m_k - M) / L come from two butterflies, then
     * each lane folds its dims (lane, lane + 32, ...) with the factors broadcast by shuffle.  (The first version - 16 threads
     * computing factors serially, then a generic per-output loop - cost 2.5 us of pure instruction issue per layer.) */
    (void)scr;
    const int warp = tid >> 5, lane = tid & 31;
    for (int h = warp; h < heads; h += FUSED_WARPS) {
        const float *r0 = land + (size_t)h * S * rec;
        const float m = lane < S ? r0[lane * rec] : -3.402823466e+38f, l = lane < S ? r0[lane * rec + 1] : 0.0f;
        const float M = warp_max(m);
        const float f = lane < S ? __expf(m - M) : 0.0f;
        const float L = warp_sum(l * f);
        const float fn = f / L;
        for (int d0 = 0; d0 < hd; d0 += 128) {
            float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            for (int k = 0; k < S; k++) {
                const float fk = __shfl_sync(0xffffffffu, fn, k);
                const float *ok = r0 + k * rec + 4 + d0 + lane;
#pragma unroll
                for (int j = 0; j < 4; j++) if (d0 + lane + 32 * j < hd) acc[j] = fmaf(ok[32 * j], fk, acc[j]);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) if (d0 + lane + 32 * j < hd) xs[h * hd + d0 + lane + 32 * j] = acc[j];
        }
    }
    if (pp) pp[23] = pp[24] = clock64();
    compute_sync();
}
