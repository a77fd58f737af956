/*
 * qtc_ll.cuh - k_lm_ll (mode 3): the fused decode-step kernel of qtc_fused_main.cuh WITHOUT grid barriers.
 * Same step (talker_attention_single x layers + final norm + head, c/qwen_tts_talker.c:119-248, 478-533; the
 * sub-talker's code-group loop :539-736), same row split, same weight loads - what changes is how the five
 * all-to-all exchanges of a layer are synchronised.
 *
 * Mode 1 pays per exchange: release-reduction on an L2 counter (waits for the CTA's stores to be acknowledged),
 * an acquire spin, then a separate L2 round trip to stage the vector.  Here every value that crosses CTAs travels
 * as ONE 8-byte word {fp32 value, 32-bit tag}; the tag is the sequence number of the exchange, unique over the
 * lifetime of the context.  A 64-bit store is single-copy atomic, so a consumer that reads the expected tag has
 * the value: it polls the data itself, with no barrier, no fence and no second round trip (the protocol NCCL calls
 * LL, applied inside one GPU).  The producer's store is fire-and-forget.
 *
 * Why no write-after-read hazard without barriers: every phase stages its WHOLE input vector before it produces
 * anything, so "CTA c produced phase p" implies "c consumed all of phase p-1".  A buffer is rewritten no earlier
 * than two exchanges after its last reader, and the rewriter has - transitively - waited for every CTA's output of
 * the phase after that reader.  The residual stream alternates between two buffers (O projection: xa -> xb, down
 * projection: xb -> xa) and a row is only ever read-modified-written by its owner.
 *
 * A poll that sees nothing for ~1 s raises a sticky error word in HBM and every later poll falls through: a
 * protocol bug shows up as an error (qwen_tts_cuda_last_error / wrong codes in the tests), never as a hung GPU.
 */
#pragma once
#include "qtc_fused_main.cuh"

typedef unsigned long long u64;

struct LLArgs {
    u64 *xa, *xb;          /* [H] residual stream, two versions */
    u64 *qkv;              /* [qd + 2 kvd] q | k | v of the token in flight */
    u64 *g;                /* [inter] SwiGLU output */
    u64 *part;             /* [heads, FUSED_MAX_SPLIT, hd + 4] attention partials */
    u64 *logits;           /* [vocab] (sub-talker: sampled inside the kernel) */
    uint32_t *epoch;       /* last tag handed out by the previous launch */
    int *err;              /* sticky: a poll timed out */
};

#define LL_SPIN_MAX (1 << 20)

__device__ __forceinline__ void ll_store(u64 *p, float v, uint32_t tag) {
    const u64 w = ((u64)tag << 32) | (u64)__float_as_uint(v);
    asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(w));       /* no "memory" clobber: the next rows' weight loads may be hoisted above it */
}
__device__ __forceinline__ u64 ll_load(const u64 *p) {
    u64 w;
    asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}
__device__ __forceinline__ void ll_load2(const u64 *p, u64 &a, u64 &b) {
    asm volatile("ld.relaxed.gpu.global.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ uint32_t ll_tag(u64 w) { return (uint32_t)(w >> 32); }
__device__ __forceinline__ float ll_val(u64 w) { return __uint_as_float((uint32_t)w); }

struct LLCtl { volatile int *dead; int *err; };
__device__ __forceinline__ bool ll_give_up(const LLCtl &c, int &spins, uint32_t tag) {
    if (*c.dead) return true;
    if (++spins > LL_SPIN_MAX) { *c.dead = 1; atomicExch(c.err, (int)tag | 1); return true; }
    return false;
}

/* one value */
__device__ __forceinline__ float ll_wait(const u64 *p, uint32_t tag, const LLCtl &c) {
    u64 w = ll_load(p);
    int spins = 0;
    while (ll_tag(w) != tag) {
        if (ll_give_up(c, spins, tag)) break;
        w = ll_load(p);
    }
    return ll_val(w);
}

/* Poll n (even) tagged values.  The words travel global -> shared memory through the async-copy path (cp.async.cg,
 * 16 bytes = two words, L1 bypassed), so a whole vector is in flight at once without holding a register - the
 * weight trip prefetched for this phase stays in registers while the CTA waits.  A thread re-requests only its
 * pairs whose tags are not there yet.  land: >= n words of shared memory.  No trailing barrier. */
__device__ __forceinline__ void ll_cp16(u64 *dst, const u64 *src) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void ll_lds2(const u64 *p, u64 &a, u64 &b) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("ld.shared.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(sa) : "memory");
}
__device__ __forceinline__ bool ll_pair_ready(const u64 *land, int q, uint32_t tag) {
    u64 w0, w1;
    ll_lds2(land + 2 * q, w0, w1);
    return ll_tag(w0) == tag && ll_tag(w1) == tag;
}
/* Round 0 asks for everything (data that is already there costs one round trip).  A thread whose pairs are not all
 * there then watches ONE of them (its lowest pending pair) with a short sleep between looks, and asks for the rest
 * again only when that one has arrived: CTAs that wait for a straggler do not flood L2 with whole-vector reads. */
__device__ __forceinline__ void ll_poll(u64 *land, const u64 *src, int n, uint32_t tag, const LLCtl &c) {
    const int npair = n >> 1, tid = threadIdx.x;
    unsigned pend = 0;
    for (int j = 0, q = tid; q < npair; q += FUSED_THREADS, j++) { ll_cp16(land + 2 * q, src + 2 * q); pend |= 1u << j; }
    asm volatile("cp.async.commit_group;" ::: "memory");
    int spins = 0;
    for (;;) {
        asm volatile("cp.async.wait_all;" ::: "memory");
        for (unsigned m = pend; m; m &= m - 1) {
            const int j = __ffs(m) - 1;
            if (ll_pair_ready(land, tid + FUSED_THREADS * j, tag)) pend &= ~(1u << j);
        }
        if (!pend) break;
        const int j0 = __ffs(pend) - 1, q0 = tid + FUSED_THREADS * j0;
        bool dead = false;
        for (;;) {
            ll_cp16(land + 2 * q0, src + 2 * q0);
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_all;" ::: "memory");
            if (ll_pair_ready(land, q0, tag)) break;
            if (ll_give_up(c, spins, tag)) { dead = true; break; }
            __nanosleep(40);
        }
        if (dead) break;
        pend &= ~(1u << j0);
        if (!pend) break;
        for (unsigned m = pend; m; m &= m - 1) { const int q = tid + FUSED_THREADS * (__ffs(m) - 1); ll_cp16(land + 2 * q, src + 2 * q); }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
}
/* poll + unpack into xs (floats); returns this thread's sum of squares.  No trailing barrier. */
__device__ __forceinline__ float ll_stage(float *xs, u64 *land, const u64 *src, int n, uint32_t tag, const LLCtl &c) {
    ll_poll(land, src, n, tag, c);
    float ss = 0.0f;
    const int npair = n >> 1;
    for (int q = threadIdx.x; q < npair; q += FUSED_THREADS) {
        u64 w0, w1;
        ll_lds2(land + 2 * q, w0, w1);
        const float v0 = ll_val(w0), v1 = ll_val(w1);
        *reinterpret_cast<float2 *>(xs + 2 * q) = make_float2(v0, v1);
        ss = fmaf(v0, v0, ss); ss = fmaf(v1, v1, ss);
    }
    return ss;
}

/* xs (staged raw, ss = per-thread sum of squares) -> rmsnorm(xs) * w in place */
__device__ __forceinline__ void ll_rms_finish(float *xs, float ss, const float *w, int n, float eps, float *red) {
    ss = csum(ss, red);
    const float inv = 1.0f / sqrtf(ss / (float)n + eps);
    for (int i = threadIdx.x * 4; i < n; i += FUSED_THREADS * 4) {
        const float4 v = *reinterpret_cast<const float4 *>(xs + i), g = *reinterpret_cast<const float4 *>(w + i);
        *reinterpret_cast<float4 *>(xs + i) = make_float4(v.x * inv * g.x, v.y * inv * g.y, v.z * inv * g.z, v.w * inv * g.w);
    }
    compute_sync();
}

/* out[r] = resid(r) + W[r,:] . xs for this CTA's rows (first pair of every warp already in registers) */
template <class R>
__device__ __forceinline__ void ll_gemv_run(const WTrip &t, const Seg3 &s, int cols, const float *xs, R resid, u64 *out, uint32_t tag, int cta, int ncta) {
    int r, rb, r1;
    if (!first_pair(s, cols, cta, ncta, r, rb, r1)) return;
    const int lane = threadIdx.x & 31;
    float *da, *db;
    const uint16_t *pa = seg_row(s, r, cols, da), *pb = seg_row(s, rb, cols, db);
    float va, vb;
    float ea = resid(r), eb = resid(rb);
    trip_finish(t, pa, pb, xs, cols, va, vb);
    if (lane == 0) {
        ll_store(out + r, ea + va, tag);
        if (rb != r) ll_store(out + rb, eb + vb, tag);
    }
    for (r += 2 * FUSED_WARPS; r < r1; r += 2 * FUSED_WARPS) {
        rb = r + FUSED_WARPS < r1 ? r + FUSED_WARPS : r;
        pa = seg_row(s, r, cols, da); pb = seg_row(s, rb, cols, db);
        ea = resid(r); eb = resid(rb);
        warp_dot2(pa, pb, xs, cols, va, vb);
        if (lane == 0) {
            ll_store(out + r, ea + va, tag);
            if (rb != r) ll_store(out + rb, eb + vb, tag);
        }
    }
}
/* the same without a prefetched trip (sub-talker input projection) */
__device__ __forceinline__ void ll_cta_gemv(const Seg3 &s, int cols, const float *xs, const float *bias, u64 *out, uint32_t tag, int cta, int ncta) {
    int r0, r1;
    row_range(s.rows[0], cta, ncta, r0, r1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = r0 + warp; r < r1; r += 2 * FUSED_WARPS) {
        const int rb = r + FUSED_WARPS < r1 ? r + FUSED_WARPS : r;
        float va, vb;
        warp_dot2(s.w[0] + (size_t)r * cols, s.w[0] + (size_t)rb * cols, xs, cols, va, vb);
        if (lane == 0) {
            ll_store(out + r, (bias ? bias[r] : 0.0f) + va, tag);
            if (rb != r) ll_store(out + rb, (bias ? bias[rb] : 0.0f) + vb, tag);
        }
    }
}
__device__ __forceinline__ void ll_swiglu_run(const WTrip &t, const uint16_t *gate, const uint16_t *up, u64 *g, uint32_t tag, int inter, int cols,
                                              const float *xs, int cta, int ncta) {
    int u0, u1;
    row_range(inter, cta, ncta, u0, u1);
    const int lane = threadIdx.x & 31;
    int u = u0 + (threadIdx.x >> 5);
    if (u >= u1) return;
    float va, vb;
    trip_finish(t, gate + (size_t)u * cols, up + (size_t)u * cols, xs, cols, va, vb);
    if (lane == 0) ll_store(g + u, swiglu_act<QTC_FAST>(va, vb), tag);
    for (u += FUSED_WARPS; u < u1; u += FUSED_WARPS) {
        warp_dot2(gate + (size_t)u * cols, up + (size_t)u * cols, xs, cols, va, vb);
        if (lane == 0) ll_store(g + u, swiglu_act<QTC_FAST>(va, vb), tag);
    }
}

/* cta_attention (qtc_fused_kernel.cuh) with tagged input (this head's q, its group's k and v) and a tagged
 * partial record [max, sum, 0, 0, out[hd]] as output - also when the head is not split.
 * sm: [qs hd | ks hd | raw 3 hd | op ngrp hd | sc chunk] */
__device__ __forceinline__ void ll_attention(const FusedArgs &a, const LLArgs &ll, const LmDev &lm, const AttPlan &p, int l, int pos, float *sm,
                                             float *kvs, float *red, uint32_t tin, uint32_t tout, const LLCtl &ctl) {
    asm volatile("cp.async.wait_all;" ::: "memory");
    if (!p.active) return;
    const int hd = lm.head_dim, half = hd >> 1, rep = lm.heads / lm.kv_heads, kvd = lm.kv_heads * hd, qd = lm.heads * hd;
    const int h = p.h, g = p.g, t0 = p.t0, t1 = p.t1;
    const int ngrp = FUSED_THREADS / hd > 0 ? FUSED_THREADS / hd : 1;
    float *qs = sm, *ks = sm + hd, *raw = sm + 2 * hd, *op = raw + 3 * hd, *sc = op + ngrp * hd;
    const LmLayerDev &w = lm.layer[l];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 3 * hd; i += FUSED_THREADS) {
        const int which = i / hd, d = i - which * hd;
        const u64 *src = ll.qkv + (which == 0 ? h * hd : which == 1 ? qd + g * hd : qd + kvd + g * hd) + d;
        raw[i] = ll_wait(src, tin, ctl);
    }
    compute_sync();
    const float *qr = raw, *kr = raw + hd, *vr = raw + 2 * hd;
    const float *cs = a.cos + (size_t)pos * hd, *sn = a.sin + (size_t)pos * hd;
    float *kc = a.kc + (size_t)l * a.kv_max * kvd, *vc = a.vc + (size_t)l * a.kv_max * kvd;
    const float *Ks = kvs, *Vs = kvs + (size_t)p.cap * hd;

    if (warp < 2) {
        const float *src = warp == 0 ? qr : kr, *nw = warp == 0 ? w.q_norm : w.k_norm;
        float *dst = warp == 0 ? qs : ks;
        float ss = 0.0f;
        for (int i = lane; i < hd; i += 32) { const float v = src[i]; ss = fmaf(v, v, ss); }
        ss = warp_sum(ss);
        const float inv = 1.0f / sqrtf(ss / (float)hd + lm.eps);
        for (int i = lane; i < half; i += 32) {
            const float xa = src[i] * inv * nw[i], xb = src[i + half] * inv * nw[i + half];
            dst[i] = xa * cs[i] - xb * sn[i];
            dst[i + half] = xb * cs[i + half] + xa * sn[i + half];
        }
    }
    compute_sync();
    const bool writer = p.sp == 0 && h % rep == 0;
    if (writer)
        for (int i = tid; i < hd; i += FUSED_THREADS) { kc[(size_t)pos * kvd + g * hd + i] = ks[i]; vc[(size_t)pos * kvd + g * hd + i] = vr[i]; }

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
    compute_sync();
    float mx = -3.402823466e+38f, tot = 0.0f;
    const bool small = t1 - t0 <= 64;
    if (small) {
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
    const int grp = tid / hd, dim = tid - grp * hd;
    if (grp < ngrp) {
        float o = 0.0f;
        for (int tb = t0 + grp; tb < t1; tb += 8 * ngrp) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int t = tb + j * ngrp;
                v[j] = 0.0f;
                if (t < t1) v[j] = t < pos ? (p.staged ? Vs[(size_t)(t - t0) * hd + dim] : __ldcg(vc + (size_t)t * kvd + g * hd + dim)) : vr[dim];
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
    u64 *mine = ll.part + ((size_t)h * p.S + p.sp) * (hd + 4);
    for (int i = tid; i < hd; i += FUSED_THREADS) { float o = 0.0f; for (int k = 0; k < ngrp; k++) o += op[k * hd + i]; ll_store(mine + 4 + i, o, tout); }
    if (tid == 0) { ll_store(mine, mx, tout); ll_store(mine + 1, tot, tout); ll_store(mine + 2, 0.0f, tout); ll_store(mine + 3, 0.0f, tout); }
    /* the appended KV row is read by other CTAs in later steps of the same launch (sub-talker): push it out */
    if (writer) __threadfence();
}

/* xs[h*hd + i] = softmax-combine of the S partial records of head h, polled into `land` */
__device__ __forceinline__ void ll_stage_attention_out(float *xs, float *scr, u64 *land, const LLArgs &ll, const LmDev &lm, int S, uint32_t tag,
                                                       const LLCtl &ctl) {
    const int hd = lm.head_dim, heads = lm.heads, qd = heads * hd, tid = threadIdx.x, rec = hd + 4;
    ll_poll(land, ll.part, heads * S * rec, tag, ctl);
    compute_sync();
    float *fac = scr;
    for (int h = tid; h < heads; h += FUSED_THREADS) {
        float M = -3.402823466e+38f, L = 0.0f;
        for (int k = 0; k < S; k++) M = fmaxf(M, ll_val(land[(h * S + k) * rec]));
        for (int k = 0; k < S; k++) L = fmaf(ll_val(land[(h * S + k) * rec + 1]), __expf(ll_val(land[(h * S + k) * rec]) - M), L);
        for (int k = 0; k < S; k++) fac[h * FUSED_MAX_SPLIT + k] = __expf(ll_val(land[(h * S + k) * rec]) - M) / L;
    }
    compute_sync();
    for (int i = tid; i < qd; i += FUSED_THREADS) {
        const int h = i / hd, d = i - h * hd;
        float acc = 0.0f;
        for (int k = 0; k < S; k++) acc = fmaf(ll_val(land[(h * S + k) * rec + 4 + d]), fac[h * FUSED_MAX_SPLIT + k], acc);
        xs[i] = acc;
    }
    compute_sync();
}

/* k_lm_ll(a0, sp, ll): sp.G == 0 -> one stack step (talker: + codec head + KV length);
 *                      sp.G  > 0 -> the sub-talker's G steps with gathers, heads and sampling inside. */
__global__ void __launch_bounds__(FUSED_BLOCK, 1) k_lm_ll(const FusedArgs a0, const SubProg sp, const LLArgs ll) {
    extern __shared__ __align__(128) float fused_smem[];
    float *sm = fused_smem;
    __shared__ float red[32];
    __shared__ SampleSmem S_samp;
    __shared__ int s_codes[QTC_MAX_GROUPS];
    __shared__ uint32_t s_rng;
    __shared__ int s_dead;
    FusedArgs a = a0;
    const LmDev &lm = *a.lm;
    const int cta = blockIdx.x, ncta = gridDim.x, tid = threadIdx.x;
    const int H = lm.hidden, I = lm.inter, L = lm.layers;
    float *att_sm = sm + a.xs_floats, *kvs = att_sm + a.att_floats;
    u64 *land = reinterpret_cast<u64 *>(kvs + a.kv_floats);      /* a.ring_off: its size in words */
    const int pos_base = *a.pos_ptr + a.pos_add;
    const uint32_t base = *ll.epoch;
    uint32_t seq = 0;
    if (tid == 0) s_dead = *ll.err != 0;
    const int nsteps = sp.G > 0 ? sp.G : 1;
    if (sp.G > 0 && tid == 0) s_codes[0] = sp.codes[0];
    __syncthreads();
    LLCtl ctl;
    ctl.dead = &s_dead; ctl.err = ll.err;
    Seg3 s; int cols;
    WTrip tp;
    const u64 *xcur = NULL;          /* NULL: the residual stream is still the plain input vector a.x (talker entry) */
    uint32_t xtag = 0;
    auto no_resid = [](int) -> float { return 0.0f; };

    for (int g = 0; g < nsteps; g++) {
        const int pos = pos_base + g;
        const AttPlan plan = att_plan(lm, pos, cta, ncta, ATT_STAGE_MAX, a.ring_off);
        if (sp.G > 0) {
            /* ST_PROJECT_INPUT: gather (every CTA for itself) -> input projection rows -> xa */
            compute_sync();
            if (g == 0) stage_vec(sm, sp.hidden, sp.Hin);
            else {
                const uint16_t *row = (g == 1 ? sp.emb0 : sp.emb[g - 2]) + (size_t)s_codes[g - 1] * sp.Hin;
                for (int i = tid; i < sp.Hin; i += FUSED_THREADS) sm[i] = bf16_to_f32(row[i]);
                compute_sync();
            }
            const uint32_t t0 = base + (++seq);
            if (sp.in_w) {
                s.n = 1; s.w[0] = sp.in_w; s.out[0] = NULL; s.rows[0] = H; cols = sp.Hin;
                ll_cta_gemv(s, cols, sm, sp.in_b, ll.xa, t0, cta, ncta);
            } else {
                int r0, r1;
                row_range(H, cta, ncta, r0, r1);
                for (int r = r0 + tid; r < r1; r += FUSED_THREADS) ll_store(ll.xa + r, r < sp.Hin ? sm[r] : 0.0f, t0);
            }
            xcur = ll.xa; xtag = t0;
            a.head_w = g >= 1 ? sp.heads[g - 1] : NULL;
            a.head_rows = sp.vocab;
        }
        phase_seg(a, lm, 0, 0, s, cols);
        gemv_prefetch(tp, s, cols, cta, ncta);
        compute_sync();
#define PROF(i) do { if (a.prof && tid == 0 && l == a.prof_layer) a.prof[cta * 32 + (i)] = clock64(); } while (0)
        for (int l = 0; l < L; l++) {
            const LmLayerDev &w = lm.layer[l];
            /* P1: (poll x) RMSNorm -> Q,K,V rows */
            PROF(0);
            if (xcur) { const float ss = ll_stage(sm, land, xcur, H, xtag, ctl); ll_rms_finish(sm, ss, w.in_norm, H, lm.eps, red); }
            else stage_rmsnorm(sm, a.x, w.in_norm, H, lm.eps, red);
            PROF(1);
            phase_seg(a, lm, l, 0, s, cols);
            const uint32_t t1 = base + (++seq);
            ll_gemv_run(tp, s, cols, sm, no_resid, ll.qkv, t1, cta, ncta);
            att_stage_issue(a, lm, plan, l, pos, kvs);
            PROF(2); PROF(3);
            /* P2: (poll q,k,v of my head) attention partial */
            const uint32_t t2 = base + (++seq);
            ll_attention(a, ll, lm, plan, l, pos, att_sm, kvs, red, t1, t2, ctl);
            PROF(4); PROF(5);
            /* P3: (poll partials) combine -> O projection + residual: xcur -> xb */
            phase_seg(a, lm, l, 1, s, cols);
            gemv_prefetch(tp, s, cols, cta, ncta);
            ll_stage_attention_out(sm, att_sm, land, ll, lm, plan.S, t2, ctl);
            PROF(15);
            const uint32_t t3 = base + (++seq);
            if (xcur) { const u64 *xc = xcur; ll_gemv_run(tp, s, cols, sm, [xc](int r) -> float { return __ldcg(reinterpret_cast<const float *>(xc + r)); }   /* validated when it was staged: a plain L2 read of the value half */, ll.xb, t3, cta, ncta); }
            else { const float *xp = a.x; ll_gemv_run(tp, s, cols, sm, [xp](int r) -> float { return __ldcg(xp + r); }, ll.xb, t3, cta, ncta); }
            PROF(6); PROF(7);
            /* P4: (poll xb) RMSNorm -> gate/up -> SwiGLU */
            swiglu_prefetch(tp, w.gate, w.up, I, H, cta, ncta);
            compute_sync();
            { const float ss = ll_stage(sm, land, ll.xb, H, t3, ctl); ll_rms_finish(sm, ss, w.post_norm, H, lm.eps, red); }
            PROF(8);
            const uint32_t t4 = base + (++seq);
            ll_swiglu_run(tp, w.gate, w.up, ll.g, t4, I, H, sm, cta, ncta);
            PROF(9); PROF(10);
            /* P5: (poll g) down projection + residual: xb -> xa */
            phase_seg(a, lm, l, 3, s, cols);
            gemv_prefetch(tp, s, cols, cta, ncta);
            compute_sync();
            ll_stage(sm, land, ll.g, I, t4, ctl);
            compute_sync();
            PROF(11);
            const uint32_t t5 = base + (++seq);
            { const u64 *xc = ll.xb; ll_gemv_run(tp, s, cols, sm, [xc](int r) -> float { return __ldcg(reinterpret_cast<const float *>(xc + r)); }   /* validated when it was staged: a plain L2 read of the value half */, ll.xa, t5, cta, ncta); }
            xcur = ll.xa; xtag = t5;
            PROF(12); PROF(13);
            if (phase_seg(a, lm, l + 1, 0, s, cols)) gemv_prefetch(tp, s, cols, cta, ncta);   /* next layer's QKV, or the head */
            compute_sync();
        }
#undef PROF
        const bool want_head = a.head_w != NULL;
        if (sp.G == 0 || want_head) {
            /* final norm (every CTA, into shared memory) */
            if (xcur) { const float ss = ll_stage(sm, land, xcur, H, xtag, ctl); ll_rms_finish(sm, ss, lm.final_norm, H, lm.eps, red); }
            else stage_rmsnorm(sm, a.x, lm.final_norm, H, lm.eps, red);
            if (cta == 0 && sp.G == 0) {          /* ctx->tk_x: the plain vector now holds the normed hidden state */
                for (int i = tid; i < H; i += FUSED_THREADS) a.x[i] = sm[i];
                if (tid == 0 && a.st && !a.st->done) a.st->pos += 1;
            }
            if (want_head) {
                phase_seg(a, lm, L, 0, s, cols);
                if (sp.G > 0) {
                    const uint32_t tl = base + (++seq);
                    ll_gemv_run(tp, s, cols, sm, no_resid, ll.logits, tl, cta, ncta);
                    /* every CTA samples for itself from the same logits: same code everywhere, no broadcast needed */
                    const int V = sp.vocab;
                    float *lg = sm, *vv = sm + V;
                    compute_sync();
                    ll_stage(lg, land, ll.logits, V, tl, ctl);
                    if (g == 1 && tid == 0) s_rng = sp.seed_bits;
                    __syncthreads();
                    const int tok = block_sample(lg, vv, V, sp.cfg.top_k, sp.cfg.top_p, sp.cfg.temperature, &s_rng, S_samp, sp.scratch_f, sp.scratch_i);
                    if (tid == 0) { s_codes[g] = tok; if (cta == 0) sp.codes[g] = tok; }
                } else {
                    gemv_run(tp, s, cols, sm, NULL, cta, ncta);        /* talker logits: plain floats for the next kernel */
                }
            }
        }
    }
    if (cta == 0 && tid == 0) *ll.epoch = base + seq;
}
