/*
 * qtc_sample.cuh - on-device token sampling with the reference's exact semantics.
 *
 * Replaces, per decision, c/qwen_tts.c:1302-1321 (suppress list, repetition
 * penalty, sample, fixed-length EOS resample) and kernel_sample_top_k
 * (c/qwen_tts_kernels.c:407-558) + rand_uniform (:384-393).
 *
 * One CTA of 1024 threads.  The reference keeps a descending insertion list of
 * the k best (value, index) pairs with ties resolved to the lower index; that is
 * the first k elements of the order (value desc, index asc).  Here: an 8-bit
 * radix select finds the k-th largest value, ties at the threshold are admitted
 * in index order, the <=k survivors are ranked by counting, and one thread
 * walks the inverse CDF with the same fp32 operation order (expf via the
 * glibc algorithm, division not reciprocal, sequential sums).
 * top_p < 1 or k outside (0, vocab) takes the reference's slow path, restated
 * with a parallel stable rank in place of its O(V^2) insertion sort.
 */
#pragma once
#include "qtc_device.cuh"
#include "qtc_internal.h"

#define SAMPLE_THREADS 1024
#define SAMPLE_MAX_K 1024

__device__ __forceinline__ uint32_t f32_order_key(float v) {
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct SampleSmem {
    uint32_t hist[256];
    uint32_t prefix, remaining, nsel, eq_total;
    int pick; int flag;
    float sel_v[SAMPLE_MAX_K]; int sel_i[SAMPLE_MAX_K];
    float ord_v[SAMPLE_MAX_K]; int ord_i[SAMPLE_MAX_K];
    float red[32];
};

/* lg: logits in shared memory [V] (already suppressed / penalised).  Returns the pick in every thread.
 * scratch_f/scratch_i: global scratch of V entries each (slow path only). */
/* The sampler runs either as a whole CTA (NTF == 0: __syncthreads, blockDim.x threads) or as the first NTF
 * threads of a larger CTA that synchronise on named barrier 1 (the fused kernels keep a producer warp apart). */
#define SSYNC() do { if (NTF) asm volatile("bar.sync 1, %0;" ::"n"(NTF ? NTF : 32) : "memory"); else __syncthreads(); } while (0)
template <int NTF> __device__ __forceinline__ float sblock_sum(float v, float *red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = ((NTF ? NTF : (int)blockDim.x) + 31) >> 5;
    v = warp_sum(v);
    SSYNC();
    if (lane == 0) red[w] = v;
    SSYNC();
    return warp_sum(lane < nw ? red[lane] : 0.0f);
}
template <int NTF> __device__ __forceinline__ float sblock_max(float v, float *red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = ((NTF ? NTF : (int)blockDim.x) + 31) >> 5;
    v = warp_max(v);
    SSYNC();
    if (lane == 0) red[w] = v;
    SSYNC();
    return warp_max(lane < nw ? red[lane] : -3.402823466e+38f);
}
template <int NTF>
__device__ int block_sample_t(const float *lg, float *vv, int V, int top_k, float top_p, float temperature,
                            uint32_t *rng, SampleSmem &S, float *scratch_f, int *scratch_i) {
    const int tid = threadIdx.x;
    const int NT = NTF ? NTF : (int)blockDim.x;
    if (temperature <= 0.0f) temperature = 1e-5f;
    SSYNC();
    if (top_p >= 1.0f && top_k == 1 && V > 1) {
        /* greedy: the list holds one entry - the first strict maximum above -FLT_MAX; its probability is
         * expf(0) = 1, the walk always lands on it, and the RNG still advances once (:451-474) */
        float bv = -3.402823466e+38f; int bi = -1;
        for (int i = tid; i < V; i += NT) { const float v = __fdiv_rn(lg[i], temperature); if (v > bv) { bv = v; bi = i; } }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (oi >= 0 && (ov > bv || (ov == bv && (bi < 0 || oi < bi)))) { bv = ov; bi = oi; }
        }
        if ((tid & 31) == 0) { S.sel_v[tid >> 5] = bv; S.sel_i[tid >> 5] = bi; }
        SSYNC();
        if (tid == 0) {
            const int nw = (NT + 31) >> 5;
            for (int w = 1; w < nw; w++) {
                const float ov = S.sel_v[w]; const int oi = S.sel_i[w];
                if (oi >= 0 && (ov > bv || (ov == bv && (bi < 0 || oi < bi)))) { bv = ov; bi = oi; }
            }
            if (bi >= 0) { uint32_t st = *rng; (void)rng_next(&st); *rng = st; }
            S.pick = bi >= 0 ? bi : 0;
        }
        SSYNC();
        return S.pick;
    }
    if (top_p >= 1.0f && top_k > 0 && top_k < V && top_k <= SAMPLE_MAX_K) {
        for (int i = tid; i < V; i += NT) vv[i] = __fdiv_rn(lg[i], temperature);
        if (tid == 0) { S.prefix = 0; S.nsel = 0; S.pick = 0; }
        SSYNC();
        /* candidates: v > -FLT_MAX (the list is initialised with -FLT_MAX and admits only strictly greater) */
        const uint32_t floor_key = f32_order_key(-3.402823466e+38f);
        uint32_t nvalid = 0;
        for (int i = tid; i < V; i += NT) nvalid += f32_order_key(vv[i]) > floor_key && vv[i] == vv[i];
        nvalid = (uint32_t)(sblock_sum<NTF>((float)nvalid, S.red) + 0.5f);
        const uint32_t kk = nvalid < (uint32_t)top_k ? nvalid : (uint32_t)top_k;
        if (kk == 0) return 0;                    /* sum == 0 branch: idx[0] < 0 -> token 0, RNG untouched */
        if (tid == 0) S.remaining = kk;
        SSYNC();
        /* radix select of the kk-th largest key, 8 bits per pass from the top */
        for (int shift = 24; shift >= 0; shift -= 8) {
            for (int i = tid; i < 256; i += NT) S.hist[i] = 0;
            SSYNC();
            const uint32_t prefix = S.prefix, himask = shift == 24 ? 0u : (0xffffffffu << (shift + 8));
            for (int i = tid; i < V; i += NT) {
                const uint32_t key = f32_order_key(vv[i]);
                if (key > floor_key && vv[i] == vv[i] && (key & himask) == prefix) atomicAdd(&S.hist[(key >> shift) & 255u], 1u);
            }
            SSYNC();
            if (tid < 32) {                        /* suffix counts: lane l owns digits 8l..8l+7 */
                uint32_t c[8], tot = 0;
#pragma unroll
                for (int j = 0; j < 8; j++) { c[j] = S.hist[tid * 8 + j]; tot += c[j]; }
                uint32_t sfx = tot;                /* inclusive suffix sum over lanes >= tid */
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_down_sync(0xffffffffu, sfx, o); if (tid + o < 32) sfx += t; }
                uint32_t above = sfx - tot;        /* elements whose digit lies in a higher lane */
                const uint32_t rem = S.remaining;
                __syncwarp();
#pragma unroll
                for (int j = 7; j >= 0; j--) {
                    if (above < rem && above + c[j] >= rem) {
                        S.prefix = prefix | ((uint32_t)(tid * 8 + j) << shift);
                        S.remaining = rem - above;
                        S.eq_total = c[j];
                    }
                    above += c[j];
                }
            }
            SSYNC();
        }
        const uint32_t tkey = S.prefix, need_eq = S.remaining, eq_total = S.eq_total;
        /* admit keys above the threshold, and ties at the threshold in index order */
        for (int i = tid; i < V; i += NT) {
            const uint32_t key = f32_order_key(vv[i]);
            if (vv[i] == vv[i] && key > tkey) { const uint32_t s = atomicAdd(&S.nsel, 1u); S.sel_v[s] = vv[i]; S.sel_i[s] = i; }
            else if (key == tkey && eq_total == need_eq) { const uint32_t s = atomicAdd(&S.nsel, 1u); S.sel_v[s] = vv[i]; S.sel_i[s] = i; }
        }
        if (eq_total != need_eq && tid < 32) {
            uint32_t base = 0;
            for (int c0 = 0; c0 < V && base < need_eq; c0 += 32) {
                const int i = c0 + tid;
                const bool eq = i < V && f32_order_key(vv[i]) == tkey;
                const uint32_t m = __ballot_sync(0xffffffffu, eq);
                const uint32_t rank = base + __popc(m & ((1u << tid) - 1u));
                if (eq && rank < need_eq) { const uint32_t s = atomicAdd(&S.nsel, 1u); S.sel_v[s] = vv[i]; S.sel_i[s] = i; }
                base += __popc(m);
            }
        }
        SSYNC();
        /* rank by (value desc, index asc) */
        for (uint32_t j = tid; j < kk; j += NT) {
            const float v = S.sel_v[j]; const int id = S.sel_i[j];
            uint32_t r = 0;                                /* branch-free and unrolled: with || / && this loop was one shared-memory round trip plus two branches per
                                                             * element (measured 140 cycles each: 7.4 of the 11.8 us a draw took inside the decode kernel) */
#pragma unroll 8
            for (uint32_t m = 0; m < kk; m++) { const float vm = S.sel_v[m]; const int im = S.sel_i[m]; r += (uint32_t)((vm > v) | ((vm == v) & (im < id))); }
            S.ord_v[r] = v; S.ord_i[r] = id;
        }
        SSYNC();
        const float mx = S.ord_v[0];
        for (uint32_t j = tid; j < kk; j += NT) S.sel_v[j] = expf_libm(__fsub_rn(S.ord_v[j], mx));
        SSYNC();
        if (tid == 0) {
            float tot = 0.0f;
            for (uint32_t j = 0; j < kk; j++) tot = __fadd_rn(tot, S.sel_v[j]);
            int pick = 0;
            if (tot > 0.0f) {
                uint32_t st = *rng;
                const float r = __fmul_rn(rng_next(&st), tot);
                *rng = st;
                float run = 0.0f;
                for (uint32_t j = 0; j < kk; j++) { run = __fadd_rn(run, S.sel_v[j]); if (run >= r) { pick = S.ord_i[j]; break; } }
            } else {
                pick = S.ord_i[0];
            }
            S.pick = pick;
        }
        SSYNC();
        return S.pick;
    }

    if (top_p >= 1.0f && top_k > 0 && top_k < V) {
        /* k beyond the parallel path's capacity: the literal insertion list, one thread (:419-484) */
        if (tid == 0) {
            float *val = scratch_f; int *idx = scratch_i;
            const int k = top_k;
            for (int j = 0; j < k; j++) { idx[j] = -1; val[j] = -3.402823466e+38f; }
            for (int i = 0; i < V; i++) {
                const float v = __fdiv_rn(lg[i], temperature);
                if (!(v > val[k - 1])) continue;
                int q = k - 1;
                for (; q > 0 && v > val[q - 1]; q--) { val[q] = val[q - 1]; idx[q] = idx[q - 1]; }
                val[q] = v; idx[q] = i;
            }
            const float mx0 = val[0];
            float tot = 0.0f;
            for (int j = 0; j < k; j++) {
                if (idx[j] < 0) { val[j] = 0.0f; continue; }
                val[j] = expf_libm(__fsub_rn(val[j], mx0)); tot = __fadd_rn(tot, val[j]);
            }
            int pick = 0;
            if (tot > 0.0f) {
                uint32_t st = *rng;
                const float r = __fmul_rn(rng_next(&st), tot);
                *rng = st;
                float run = 0.0f;
                for (int j = 0; j < k; j++) { run = __fadd_rn(run, val[j]); if (run >= r) { pick = idx[j] >= 0 ? idx[j] : 0; break; } }
            } else pick = idx[0] >= 0 ? idx[0] : 0;
            S.pick = pick;
        }
        SSYNC();
        return S.pick;
    }
    /* ---- slow path, c/qwen_tts_kernels.c:486-557 ---- */
    float *p = scratch_f; int *ord = scratch_i;
    for (int i = tid; i < V; i += NT) p[i] = __fdiv_rn(lg[i], temperature);
    SSYNC();
    float mx = -3.402823466e+38f;
    for (int i = tid; i < V; i += NT) mx = fmaxf(mx, p[i]);
    mx = sblock_max<NTF>(mx, S.red);
    for (int i = tid; i < V; i += NT) p[i] = expf_libm(__fsub_rn(p[i], mx));
    SSYNC();
    if (tid == 0) {
        float tot = 0.0f;
        for (int i = 0; i < V; i++) tot = __fadd_rn(tot, p[i]);
        S.red[0] = __fdiv_rn(1.0f, tot);
    }
    SSYNC();
    { const float inv = S.red[0]; SSYNC(); for (int i = tid; i < V; i += NT) p[i] = __fmul_rn(p[i], inv); }
    SSYNC();
    const bool use_k = top_k > 0 && top_k < V, use_p = top_p < 1.0f;
    if (use_k || use_p) {                               /* stable descending order == the insertion sort */
        for (int i = tid; i < V; i += NT) {
            const float v = p[i];
            int r = 0;
            for (int m = 0; m < V; m++) r += (p[m] > v) || (p[m] == v && m < i);
            ord[r] = i;
        }
        SSYNC();
    }
    if (use_k) {
        const float thr = p[ord[top_k - 1]];
        SSYNC();
        for (int i = tid; i < V; i += NT) if (p[i] < thr) p[i] = 0.0f;
        SSYNC();
    }
    if (tid == 0) {
        if (use_p) {
            /* zeroed entries sort behind every kept one and add nothing, so the pre-zeroing order serves */
            float run = 0.0f; int cut = V;
            for (int i = 0; i < V; i++) { run = __fadd_rn(run, p[ord[i]]); if (run >= top_p) { cut = i + 1; break; } }
            for (int i = cut; i < V; i++) p[ord[i]] = 0.0f;
        }
        float tot = 0.0f;
        for (int i = 0; i < V; i++) tot = __fadd_rn(tot, p[i]);
        const float inv = tot > 0.0f ? __fdiv_rn(1.0f, tot) : 1.0f;
        uint32_t st = *rng;
        const float r = rng_next(&st);
        *rng = st;
        float run = 0.0f; int pick = 0;
        for (int i = 0; i < V; i++) {
            const float q = tot > 0.0f ? __fmul_rn(p[i], inv) : p[i];
            run = __fadd_rn(run, q);
            if (run >= r) { pick = i; break; }
        }
        S.pick = pick;
    }
    SSYNC();
    return S.pick;
}
#undef SSYNC
__device__ __forceinline__ int block_sample(const float *lg, float *vv, int V, int top_k, float top_p, float temperature,
                                            uint32_t *rng, SampleSmem &S, float *scratch_f, int *scratch_i) {
    return block_sample_t<0>(lg, vv, V, top_k, top_p, temperature, rng, S, scratch_f, scratch_i);
}

/* The same draw with a third of the block barriers, for the common case top_p >= 1, 1 < k <= 64, whole CTA (mode 2's in-kernel
 * sampler: 15 draws per frame, each on the critical path of every CTA).  Instead of a 4-pass radix select (12 barriers) the k-th
 * value is bracketed from the data: every element >= mean + c sd is a candidate (one pass, candidates appended through one shared
 * counter); if at least k and at most 512 qualified, the k largest of ALL elements are among them - ties with the k-th included -
 * and ranking the candidates by (value desc, index asc) gives exactly the reference's list; otherwise c is lowered, and after three
 * tries (or for any other parameter set) the radix path above decides.  From the ordered list on, the arithmetic is the same
 * statement by statement (glibc-algorithm expf, sequential sums, the same RNG draw). */
__device__ int block_sample_thr(const float *lg, float *vv, int V, int top_k, float top_p, float temperature,
                                uint32_t *rng, SampleSmem &S, float *scratch_f, int *scratch_i, long long *dbg = NULL) {
#define TPROF(i) do { if (dbg && threadIdx.x == 0) dbg[i] = clock64(); } while (0)
    const int tid = threadIdx.x, NT = blockDim.x, lane = tid & 31, warp = tid >> 5, nw = NT >> 5;
    if (!(top_p >= 1.0f && top_k > 1 && top_k <= 64 && top_k * 8 <= V && nw <= 32 && (NT & 31) == 0))
        return block_sample(lg, vv, V, top_k, top_p, temperature, rng, S, scratch_f, scratch_i);
    if (temperature <= 0.0f) temperature = 1e-5f;
    __syncthreads();
    TPROF(0);
    float s1 = 0.0f, s2 = 0.0f, cnt = 0.0f;
    for (int i = tid; i < V; i += NT) {
        const float v = __fdiv_rn(lg[i], temperature);
        vv[i] = v;
        if (v > -1e8f && v < 1e30f) { s1 += v; s2 = fmaf(v, v, s2); cnt += 1.0f; }      /* suppressed logits (-1e9) stay out of the moments: they only bracket the k-th value */
    }
    s1 = warp_sum(s1); s2 = warp_sum(s2); cnt = warp_sum(cnt);
    if (lane == 0) { S.ord_v[warp] = s1; S.ord_v[32 + warp] = s2; S.ord_v[64 + warp] = cnt; }
    if (tid == 0) S.nsel = 0;
    __syncthreads();
    s1 = warp_sum(lane < nw ? S.ord_v[lane] : 0.0f); s2 = warp_sum(lane < nw ? S.ord_v[32 + lane] : 0.0f); cnt = warp_sum(lane < nw ? S.ord_v[64 + lane] : 0.0f);
    TPROF(1);
    const float mean = cnt > 0.0f ? s1 / cnt : 0.0f;
    const float var = cnt > 0.0f ? s2 / cnt - mean * mean : 0.0f;
    const float sd = sqrtf(var > 0.0f ? var : 0.0f);
    uint32_t C = 0;
    bool ok = false;
    for (int attempt = 0; attempt < 3 && !ok; attempt++) {
        const float T0 = mean + (1.6f - 0.6f * (float)attempt) * sd;
        for (int i = tid; i < V; i += NT) {
            const float v = vv[i];
            if (v >= T0) { const uint32_t q = atomicAdd(&S.nsel, 1u); if (q < 512u) { S.sel_v[q] = v; S.sel_i[q] = i; } }
        }
        __syncthreads();
        C = S.nsel;
        ok = C >= (uint32_t)top_k && C <= 512u;
        if (!ok) {
            __syncthreads();                               /* every thread has read the count */
            if (tid == 0) S.nsel = 0;
            __syncthreads();
        }
    }
    if (!ok) return block_sample(lg, vv, V, top_k, top_p, temperature, rng, S, scratch_f, scratch_i);
    TPROF(2);
    const uint32_t kk = (uint32_t)top_k;
    for (uint32_t j0 = 0; j0 < C; j0 += (uint32_t)NT >> 2) {         /* four threads per candidate, a quarter of the list each */
        const uint32_t jj = j0 + ((uint32_t)tid >> 2), part = (uint32_t)tid & 3u;
        const bool on = jj < C;
        const float v = on ? S.sel_v[jj] : 0.0f; const int id = on ? S.sel_i[jj] : 0;
        uint32_t r = 0;
#pragma unroll 8
        for (uint32_t m = part; m < C; m += 4) { const float vm = S.sel_v[m]; const int im = S.sel_i[m]; r += (uint32_t)((vm > v) | ((vm == v) & (im < id))); }
        r += __shfl_xor_sync(0xffffffffu, r, 1);
        r += __shfl_xor_sync(0xffffffffu, r, 2);
        if (on && part == 0 && r < kk) { S.ord_v[r] = v; S.ord_i[r] = id; }
    }
    __syncthreads();
    TPROF(3);
    const float mx = S.ord_v[0];
    float ev = 0.0f;
    if ((uint32_t)tid < kk) ev = expf_libm(__fsub_rn(S.ord_v[tid], mx));
    __syncthreads();                                       /* the candidates in sel_v have been ranked by everyone */
    if ((uint32_t)tid < kk) S.sel_v[tid] = ev;
    __syncthreads();
    TPROF(4);
    if (tid == 0) {
        float tot = 0.0f;
        for (uint32_t j = 0; j < kk; j++) tot = __fadd_rn(tot, S.sel_v[j]);
        int pick = 0;
        if (tot > 0.0f) {
            uint32_t st = *rng;
            const float r = __fmul_rn(rng_next(&st), tot);
            *rng = st;
            float run = 0.0f;
            for (uint32_t j = 0; j < kk; j++) { run = __fadd_rn(run, S.sel_v[j]); if (run >= r) { pick = S.ord_i[j]; break; } }
        } else {
            pick = S.ord_i[0];
        }
        S.pick = pick;
    }
    TPROF(5);
    __syncthreads();
    TPROF(6);
    if (dbg && threadIdx.x == 0) dbg[7] = (long long)C;
    return S.pick;
#undef TPROF
}

struct SampleArgs {
    const float *logits;
    SampleCfg cfg;
    uint32_t *rng;
    int *hist_count;       /* talker only */
    int *out_token;        /* where the pick goes (codes[...]) or NULL */
    GenState *st;          /* talker role: EOS / fixed-length policy */
    int *codes; int G;     /* talker role: codes[0] = token (the frame's code row) */
    int role;              /* 0 talker frame decision, 1 sub-talker / bare sample */
    uint32_t seed_bits; int reseed;   /* sub-talker: re-seed before sampling (first code group of a frame) */
    float *scratch_f; int *scratch_i;
    /* batched launch (grid = utterances): element strides between utterances; all 0 for a single utterance */
    int bs_logits, bs_rng, bs_hist, bs_out, bs_codes, bs_scratch;
    const uint32_t *seed_tab;   /* batched re-seed: seed bits per utterance (NULL: seed_bits for all) */
};

/* dynamic smem: 2*V floats.  blockIdx.x = utterance */
__global__ void __launch_bounds__(SAMPLE_THREADS) k_sample(const SampleArgs a0) {
    pdl_enter();
    extern __shared__ __align__(16) float dyn[];
    __shared__ SampleSmem S;
    SampleArgs a = a0;
    {
        const size_t b = blockIdx.x;
        a.logits += b * a.bs_logits;
        if (a.rng) a.rng += b * a.bs_rng;
        if (a.hist_count) a.hist_count += b * a.bs_hist;
        if (a.out_token) a.out_token += b * a.bs_out;
        if (a.codes) a.codes += b * a.bs_codes;
        if (a.st) a.st += b;
        if (a.scratch_f) { a.scratch_f += b * a.bs_scratch; a.scratch_i += b * a.bs_scratch; }
    }
    const int V = a.cfg.vocab, tid = threadIdx.x;
    float *lg = dyn, *vv = dyn + V;
    if (a.role == 0 && a.st->done) return;
    if (a.reseed && tid == 0) *a.rng = a.seed_tab ? a.seed_tab[blockIdx.x] : a.seed_bits;
    for (int i = tid; i < V; i += blockDim.x) {
        float l = a.logits[i];
        if (a.cfg.suppress_from >= 0 && i >= a.cfg.suppress_from && i != a.cfg.eos_id) l = -1e9f;
        if (a.hist_count && a.cfg.rep_penalty != 1.0f) {
            const int c = a.hist_count[i];
            for (int j = 0; j < c; j++) l = l > 0.0f ? __fdiv_rn(l, a.cfg.rep_penalty) : __fmul_rn(l, a.cfg.rep_penalty);
        }
        lg[i] = l;
    }
    __syncthreads();
    /* the bracketing select (a third of the radix select's block barriers; falls back to it when its conditions do not hold) */
    int tok = block_sample_thr(lg, vv, V, a.cfg.top_k, a.cfg.top_p, a.cfg.temperature, a.rng, S, a.scratch_f, a.scratch_i);
    if (a.role == 0) {
        const int eos = a.cfg.eos_id;
        if (a.st->fixed && tok == eos) {                       /* c/qwen_tts.c:1315-1321: second draw */
            __syncthreads();
            if (tid == 0) lg[eos] = -1e9f;
            __syncthreads();
            tok = block_sample(lg, vv, V, a.cfg.top_k, a.cfg.top_p, a.cfg.temperature, a.rng, S, a.scratch_f, a.scratch_i);
        }
        if (tid == 0) {
            if (!a.st->fixed && tok == eos) { a.st->done = 1; a.st->stop_step = a.st->step; }
            else {
                a.st->token = tok;
                a.hist_count[tok] += 1;
                a.codes[0] = tok;
            }
        }
    } else if (tid == 0 && a.out_token) {
        *a.out_token = tok;
    }
}
