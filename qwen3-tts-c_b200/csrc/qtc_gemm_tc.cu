/*
 * qtc_gemm_tc.cu - the batched (lock-step, B <= 64 utterances) projections of the talker / sub-talker
 * as tcgen05 GEMMs:  Y[b, n] += sum_k X[b, k] * W[n, k]      (kernel_matvec_bf16, c/qwen_tts_kernels.c:95-149,
 * once per utterance in the reference; here one pass over W serves the whole batch).
 *
 * Mapping (sm_100a): the bf16 weight matrix is the A operand - a CTA owns 128 weight rows (UMMA M = 128) and one
 * K range (split-K so that tiles x splits ~ the 148 SMs: the op is weight-bandwidth bound at batch 64, every SM
 * has to pull HBM); the activations, rounded to bf16, are the B operand (UMMA N = 64 = batch, K-major).  TMA
 * (cp.async.bulk.tensor, SWIZZLE_128B) fills a 6-stage shared-memory pipeline of [128 x 64] + [64 x 64] tiles,
 * one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (4 per stage), the fp32 accumulator [128 x 64]
 * lives in TMEM (64 columns), four epilogue warps read it back with tcgen05.ld (32 lanes x 32 bit x 16) and add it
 * to Y with fp32 reductions in L2 (red.global.add.f32) - Y is pre-loaded with zero, a bias or the residual stream,
 * so split-K needs no second pass and "x += W.h" costs nothing extra.
 *
 * Warp roles: 0 TMA producer, 1 MMA issuer (+ TMEM alloc/free), 2..5 epilogue (TMEM lane quarter = warp % 4).
 */
#include "qtc_internal.h"
#include "qtc_device.cuh"
#include "qtc_tc.cuh"

#include <cuda.h>
#include <unordered_map>


#define TC_BN 64            /* batch columns per piece; UMMA N = 64 x pieces */
#define TC_BK 64            /* K per stage: 64 bf16 = one 128-byte swizzle row */
#define TC_A_BYTES (TC_BM * TC_BK * 2)
#define TC_THREADS 192
/* NP = activation pieces: 1 = activations rounded to bf16; 3 = exact split x = hi + mid + lo (three bf16 pieces stacked
 * as 192 B-operand rows), which makes every product exact and leaves only fp32 summation-order differences */
/* ST = pipeline stages; 0 = as many as one CTA per SM holds (6 / 5).  ST = 2 (81 KB at three pieces) lets TWO CTAs share an SM: the
 * CTAs of the next launch become resident - barrier init, TMEM allocation, first weight tiles under way - while this one's are still
 * streaming or draining their epilogue, which a 206 KB CTA that owns the SM rules out.  The split-K accumulate path uses ST = 2
 * (QTC_GEMM_STAGES=5 brings the deep pipeline back for A/B); K splits are then sized for two CTAs per SM. */
template <int NP, int ST = 0> struct TcCfg {
    static constexpr int N = TC_BN * NP;                       /* UMMA N */
    static constexpr int B_BYTES = N * TC_BK * 2;
    static constexpr int STAGE_BYTES = TC_A_BYTES + B_BYTES;
    static constexpr int STAGES = ST > 0 ? ST : (NP == 1 ? 6 : 5);
    static constexpr int SMEM = STAGES * STAGE_BYTES + 1024;
    static constexpr int TMEM_COLS = NP == 1 ? 64 : 256;       /* power of two >= N */
};

/* grid (row tiles, K splits).  Y [batch, ldy] fp32, pre-initialised; rows >= N and batch columns >= nbatch are dropped */
/* EPI 0: Y += acc (fp32 reductions in L2, any K split).
 * EPI 1 (one K split only): the weight rows are gate/up interleaved (row 2u = gate u, 2u+1 = up u); the epilogue forms
 *        silu(gate.x) * (up.x) (c/qwen_tts_kernels.c:229-232) and stores it as bf16 pieces - the next GEMM's input - at
 *        xo[(p*64 + b) * (N/2) + u]: SwiGLU costs no kernel and no fp32 round trip.
 * EPI 2: K split s STORES its partial sums at Y + s * split_stride (plain stores, no pre-initialised Y): the consumer adds the
 *        splits in a fixed order, so the result is bit-reproducible run to run (the batch-1 prefill, qtc_prefill.cuh). */
template <int NP, int EPI, int ST = 0>
__global__ void __launch_bounds__(TC_THREADS, ST == 2 ? 2 : 1) k_gemm_tc(const __grid_constant__ CUtensorMap mapW, const __grid_constant__ CUtensorMap mapX,
                                                            float *Y, int N, int ldy, int kb_per_split, int nbatch, uint16_t *xo, int xo_pieces, int split_stride) {
    extern __shared__ __align__(1024) unsigned char tc_smem_raw[];
    typedef TcCfg<NP, ST> Cfg;
    constexpr int TC_STAGES = Cfg::STAGES, TC_STAGE_BYTES = Cfg::STAGE_BYTES;
    __shared__ __align__(8) uint64_t full[TC_STAGES], empty[TC_STAGES], acc_full;
    __shared__ uint32_t tmem_base_s;
    unsigned char *tiles = reinterpret_cast<unsigned char *>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * TC_BM, kb0 = blockIdx.y * kb_per_split;

    if (threadIdx.x == 0) {
        for (int i = 0; i < TC_STAGES; i++) { tc_mbar_init(&full[i], 1); tc_mbar_init(&empty[i], 1); }
        tc_mbar_init(&acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapW) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapX) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(&tmem_base_s)), "r"(Cfg::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;

    /* programmatic dependent launch: everything above ran while the previous kernel was still finishing; the weights do not
     * depend on it either, so the producer fills the pipeline's weight tiles BEFORE it waits - only the activation tiles and
     * the accumulation into Y are ordered after the previous kernel. */
    pdl_trigger();
    if (warp == 0) {
        if (lane == 0) {
            const int npre = kb_per_split < TC_STAGES ? kb_per_split : TC_STAGES;
            for (int kb = 0; kb < npre; kb++) {
                unsigned char *sa = tiles + (size_t)kb * TC_STAGE_BYTES;
                tc_mbar_expect_tx(&full[kb], TC_STAGE_BYTES);
                tc_tma_2d(sa, &mapW, (kb0 + kb) * TC_BK, row0, &full[kb]);
            }
            pdl_wait();
            for (int kb = 0; kb < npre; kb++) {
                unsigned char *sb = tiles + (size_t)kb * TC_STAGE_BYTES + TC_A_BYTES;
                tc_tma_2d(sb, &mapX, (kb0 + kb) * TC_BK, 0, &full[kb]);
            }
            for (int kb = npre; kb < kb_per_split; kb++) {
                const int st = kb % TC_STAGES, ph = (kb / TC_STAGES) & 1;
                tc_mbar_wait(&empty[st], ph ^ 1);
                unsigned char *sa = tiles + (size_t)st * TC_STAGE_BYTES, *sb = sa + TC_A_BYTES;
                tc_mbar_expect_tx(&full[st], TC_STAGE_BYTES);
                tc_tma_2d(sa, &mapW, (kb0 + kb) * TC_BK, row0, &full[st]);
                tc_tma_2d(sb, &mapX, (kb0 + kb) * TC_BK, 0, &full[st]);
            }
        }
    } else if (warp == 1) {
        pdl_wait();
        if (lane == 0) {
            const uint32_t idesc = tc_idesc(Cfg::N);
            for (int kb = 0; kb < kb_per_split; kb++) {
                const int st = kb % TC_STAGES, ph = (kb / TC_STAGES) & 1;
                tc_mbar_wait(&full[st], ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = tc_smem_u32(tiles + (size_t)st * TC_STAGE_BYTES), sb = sa + TC_A_BYTES;
                const uint64_t ad = tc_smem_desc(sa), bd = tc_smem_desc(sb);
#pragma unroll
                for (int k = 0; k < TC_BK / 16; k++)          /* UMMA K = 16 bf16 = 32 bytes inside the swizzle row */
                    tc_mma(tmem_d, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc, (kb | k) != 0);
                tc_commit(&empty[st]);                         /* frees the stage when these MMAs have read it */
            }
            tc_commit(&acc_full);
        }
    } else {
        /* epilogue: warp w may touch TMEM lanes [32 (w % 4), +32); thread = one weight row, 64 batch columns */
        const int q = warp & 3, row = row0 + q * 32 + lane;
        pdl_wait();
        tc_mbar_wait(&acc_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int c0 = 0; c0 < TC_BN; c0 += 16) {
            float acc[16];
#pragma unroll
            for (int j = 0; j < 16; j++) acc[j] = 0.0f;
#pragma unroll
            for (int p = 0; p < NP; p++) {
                uint32_t v[16];
                const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(p * TC_BN + c0);
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                               "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                             : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 16; j++) acc[j] += __uint_as_float(v[j]);
            }
            if (EPI == 0) {
                if (row < N) {
#pragma unroll
                    for (int j = 0; j < 16; j++)
                        if (c0 + j < nbatch) atomicAdd(Y + (size_t)(c0 + j) * ldy + row, acc[j]);
                }
            } else if (EPI == 2) {
                if (row < N) {
                    float *Ys = Y + (size_t)blockIdx.y * split_stride;
#pragma unroll
                    for (int j = 0; j < 16; j++)
                        if (c0 + j < nbatch) Ys[(size_t)(c0 + j) * ldy + row] = acc[j];
                }
            } else {
                const int I = N >> 1, u = row >> 1;
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const float up = __shfl_xor_sync(0xffffffffu, acc[j], 1);       /* odd lane = the up row of my unit */
                    if (!(lane & 1) && row < N && c0 + j < nbatch) {
                        float v = (acc[j] / (1.0f + __expf(-acc[j]))) * up;
                        for (int p = 0; p < xo_pieces; p++) {
                            const uint32_t bits = __float_as_uint(v), h = (bits + 0x7fffu + ((bits >> 16) & 1u)) >> 16;
                            xo[((size_t)p * TC_BN + (c0 + j)) * I + u] = (uint16_t)h;
                            v -= __uint_as_float(h << 16);
                        }
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(Cfg::TMEM_COLS) : "memory");
    }
}

/* ---- host side ---------------------------------------------------------------------------------------- */
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = NULL;
static int get_encode() {
    if (g_encode) return 0;
    void *fn = NULL;
    cudaDriverEntryPointQueryResult qr;
    QTC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    if (!fn || qr != cudaDriverEntryPointSuccess) { qtc_set_error("cuTensorMapEncodeTiled not available from the driver"); return -1; }
    g_encode = (EncodeTiledFn)fn;
    return 0;
}
/* bf16 [rows, cols] row-major -> boxes of [box_rows x 64 columns], 128-byte swizzle, zero fill out of bounds */
static int make_map(CUtensorMap *m, const void *ptr, int rows, int cols, int box_rows) {
    if (get_encode()) return -1;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows}, strides[1] = {(cuuint64_t)cols * 2};
    const cuuint32_t box[2] = {TC_BK, (cuuint32_t)box_rows}, es[2] = {1, 1};
    const CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void *)ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { qtc_set_error("cuTensorMapEncodeTiled failed (%d) for a [%d, %d] bf16 matrix", (int)r, rows, cols); return -1; }
    return 0;
}

struct MapKey { const void *p; int rows, cols, box; bool operator==(const MapKey &o) const { return p == o.p && rows == o.rows && cols == o.cols && box == o.box; } };
struct MapHash { size_t operator()(const MapKey &k) const { return std::hash<const void *>()(k.p) ^ ((size_t)k.rows * 1315423911u) ^ ((size_t)k.cols << 20) ^ (size_t)k.box; } };
static std::unordered_map<MapKey, CUtensorMap, MapHash> g_maps;
const CUtensorMap *qtc_tmap_bf16(const void *ptr, int rows, int cols, int box) {
    const MapKey k = {ptr, rows, cols, box};
    auto it = g_maps.find(k);
    if (it != g_maps.end()) return &it->second;
    CUtensorMap m;
    if (make_map(&m, ptr, rows, cols, box)) return NULL;
    return &g_maps.emplace(k, m).first->second;
}
void qtc_gemm_tc_forget(const void *ptr) {          /* a buffer was freed: drop its maps */
    for (auto it = g_maps.begin(); it != g_maps.end();) { if (it->first.p == ptr) it = g_maps.erase(it); else ++it; }
}

/* K split: the divisor of the k-block count that brings tiles x splits closest to the SM count without passing 2 waves */
static int pick_splits(int tiles, int kblocks, int sms) {
    int best = 1; double bs = 1e30;
    for (int s = 1; s <= kblocks; s++) {
        if (kblocks % s) continue;
        const int ctas = tiles * s, waves = (ctas + sms - 1) / sms;
        /* time ~ waves x (work per CTA + fixed pipeline fill of ~3 k-blocks) */
        const double cost = waves * ((double)kblocks / s + 3.0);
        if (cost < bs - 1e-9) { bs = cost; best = s; }
    }
    return best;
}

static int g_tc_attr = 0;
/* Y[b, n] += sum_k X[b, k] W[n, k];  X: bf16 [64 x pieces, K] on device (piece p of utterance b in row 64 p + b; rows of
 * absent utterances zero or ignored), W: bf16 [N, K].  pieces = 1 or 3. */
int qtc_gemm_tc_ex(qtc_ctx *c, float *Y, int ldy, const uint16_t *X, const uint16_t *W, int nbatch, int N, int K, int pieces, uint16_t *swiglu_out);
int qtc_gemm_tc(qtc_ctx *c, float *Y, int ldy, const uint16_t *X, const uint16_t *W, int nbatch, int N, int K, int pieces) {
    return qtc_gemm_tc_ex(c, Y, ldy, X, W, nbatch, N, K, pieces, NULL);
}
/* swiglu_out != NULL: W is gate/up interleaved [N = 2I, K]; the result goes to swiglu_out as bf16 pieces [pieces*64, I] (Y unused) */
int qtc_gemm_tc_ex(qtc_ctx *c, float *Y, int ldy, const uint16_t *X, const uint16_t *W, int nbatch, int N, int K, int pieces, uint16_t *swiglu_out) {
    if (K % TC_BK || nbatch < 1 || nbatch > TC_BN || (pieces != 1 && pieces != 3)) {
        qtc_set_error("gemm_tc: K %% 64 != 0, batch outside 1..64 or pieces not 1/3 (K=%d, batch=%d, pieces=%d)", K, nbatch, pieces); return -1;
    }
    if (!g_tc_attr) {
        QTC_CUDA(cudaFuncSetAttribute(k_gemm_tc<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<1>::SMEM));
        QTC_CUDA(cudaFuncSetAttribute(k_gemm_tc<3, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<3>::SMEM));
        QTC_CUDA(cudaFuncSetAttribute(k_gemm_tc<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<1>::SMEM));
        QTC_CUDA(cudaFuncSetAttribute(k_gemm_tc<3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<3>::SMEM));
        g_tc_attr = 1;
    }
    const CUtensorMap *mw = qtc_tmap_bf16(W, N, K, TC_BM), *mx = qtc_tmap_bf16(X, TC_BN * pieces, K, TC_BN * pieces);
    if (!mw || !mx) return -1;
    static int st2 = -1;
    if (st2 < 0) {
        const char *e = getenv("QTC_GEMM_STAGES");
        st2 = !(e && atoi(e) != 2);                                   /* default since the A/B on config 4: 5.89 -> 5.64 ms per 64-utterance frame */
        if (st2) {
            QTC_CUDA(cudaFuncSetAttribute(k_gemm_tc<1, 0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<1, 2>::SMEM));
            QTC_CUDA(cudaFuncSetAttribute(k_gemm_tc<3, 0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<3, 2>::SMEM));
        }
    }
    const int tiles = (N + TC_BM - 1) / TC_BM, kblocks = K / TC_BK, splits = swiglu_out ? 1 : pick_splits(tiles, kblocks, st2 ? 2 * c->num_sms : c->num_sms);
    if (swiglu_out && (N & 1)) { qtc_set_error("gemm_tc: interleaved gate/up needs an even row count"); return -1; }
    const dim3 grid(tiles, splits);
    if (st2 && !swiglu_out) {
        if (pieces == 1) qtc_launch_pdl(c->stream, k_gemm_tc<1, 0, 2>, grid, dim3(TC_THREADS), (size_t)TcCfg<1, 2>::SMEM, *mw, *mx, Y, N, ldy, kblocks / splits, nbatch, (uint16_t *)NULL, 0, 0);
        else qtc_launch_pdl(c->stream, k_gemm_tc<3, 0, 2>, grid, dim3(TC_THREADS), (size_t)TcCfg<3, 2>::SMEM, *mw, *mx, Y, N, ldy, kblocks / splits, nbatch, (uint16_t *)NULL, 0, 0);
    } else if (swiglu_out) {
        if (pieces == 1) qtc_launch_pdl(c->stream, k_gemm_tc<1, 1>, grid, dim3(TC_THREADS), (size_t)TcCfg<1>::SMEM, *mw, *mx, Y, N, ldy, kblocks, nbatch, swiglu_out, pieces, 0);
        else qtc_launch_pdl(c->stream, k_gemm_tc<3, 1>, grid, dim3(TC_THREADS), (size_t)TcCfg<3>::SMEM, *mw, *mx, Y, N, ldy, kblocks, nbatch, swiglu_out, pieces, 0);
    } else {
        if (pieces == 1) qtc_launch_pdl(c->stream, k_gemm_tc<1, 0>, grid, dim3(TC_THREADS), (size_t)TcCfg<1>::SMEM, *mw, *mx, Y, N, ldy, kblocks / splits, nbatch, (uint16_t *)NULL, 0, 0);
        else qtc_launch_pdl(c->stream, k_gemm_tc<3, 0>, grid, dim3(TC_THREADS), (size_t)TcCfg<3>::SMEM, *mw, *mx, Y, N, ldy, kblocks / splits, nbatch, (uint16_t *)NULL, 0, 0);
    }
    c->launches++;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { qtc_set_error("gemm_tc launch -> %s", cudaGetErrorString(e)); return -1; }
    return 0;
}

/* K-split count qtc_gemm_tc_parts will use for a [N, K] weight matrix */
int qtc_gemm_tc_splits(const qtc_ctx *c, int N, int K) { return pick_splits((N + TC_BM - 1) / TC_BM, K / TC_BK, c->num_sms); }
/* parts[s][b, n] = sum over K split s of X[b, k] W[n, k] for s < qtc_gemm_tc_splits(c, N, K): split s at parts + s * split_stride,
 * rows of ldy floats; three exact activation pieces; nothing is accumulated - every element of every split is stored. */
int qtc_gemm_tc_parts(qtc_ctx *c, float *parts, int ldy, int split_stride, const uint16_t *X, const uint16_t *W, int nbatch, int N, int K) {
    if (K % TC_BK || nbatch < 1 || nbatch > TC_BN) { qtc_set_error("gemm_tc_parts: K %% 64 != 0 or batch outside 1..64 (K=%d, batch=%d)", K, nbatch); return -1; }
    static int attr = 0;
    if (!attr) { QTC_CUDA(cudaFuncSetAttribute(k_gemm_tc<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<3>::SMEM)); attr = 1; }
    const CUtensorMap *mw = qtc_tmap_bf16(W, N, K, TC_BM), *mx = qtc_tmap_bf16(X, TC_BN * 3, K, TC_BN * 3);
    if (!mw || !mx) return -1;
    const int tiles = (N + TC_BM - 1) / TC_BM, kblocks = K / TC_BK, splits = pick_splits(tiles, kblocks, c->num_sms);
    qtc_launch_pdl(c->stream, k_gemm_tc<3, 2>, dim3(tiles, splits), dim3(TC_THREADS), (size_t)TcCfg<3>::SMEM, *mw, *mx, parts, N, ldy, kblocks / splits, nbatch, (uint16_t *)NULL, 0, split_stride);
    c->launches++;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { qtc_set_error("gemm_tc_parts launch -> %s", cudaGetErrorString(e)); return -1; }
    return 0;
}

/* fp32 [rows, cols] -> bf16 pieces [pieces x 64, cols]: piece p of row r at row 64 p + r; absent rows zero.
 * pieces = 1: round to nearest even; 3: exact split v = hi + mid + lo */
__global__ void k_to_bf16_rows(uint16_t *out, const float *in, int rows, int cols, int pieces) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, n = (size_t)TC_BN * cols;
    if (i >= n) return;
    const size_t r = i / cols;
    float v = r < (size_t)rows ? in[i] : 0.0f;
    for (int p = 0; p < pieces; p++) {
        const uint32_t u = __float_as_uint(v), h = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;
        out[(size_t)p * n + i] = (uint16_t)h;
        v -= __uint_as_float(h << 16);
    }
}

/* test surface: C[M, N] = A[M, K] (fp32, rounded to bf16) . W[N, K]^T (bf16), host pointers, M <= 64 */
extern "C" int qwen_tts_cuda_gemm_bf16_tc(float *C, const float *A, const uint16_t *W, int M, int N, int K, int pieces) {
    if (!C || !A || !W || M < 1 || M > TC_BN || N < 1 || K < TC_BK || K % TC_BK || (pieces != 1 && pieces != 3)) { qtc_set_error("gemm_bf16_tc: bad shape"); return -1; }
    qtc_ctx tmp;                                   /* only stream / num_sms / launches are used */
    memset((void *)&tmp, 0, offsetof(qtc_ctx, allocs));
    int dev = 0;
    QTC_CUDA(cudaGetDevice(&dev));
    QTC_CUDA(cudaDeviceGetAttribute(&tmp.num_sms, cudaDevAttrMultiProcessorCount, dev));
    float *dA = NULL, *dC = NULL; uint16_t *dW = NULL, *dX = NULL;
    QTC_CUDA(cudaMalloc(&dA, (size_t)M * K * 4)); QTC_CUDA(cudaMalloc(&dC, (size_t)M * N * 4));
    QTC_CUDA(cudaMalloc(&dW, (size_t)N * K * 2)); QTC_CUDA(cudaMalloc(&dX, (size_t)TC_BN * pieces * K * 2));
    QTC_CUDA(cudaMemcpy(dA, A, (size_t)M * K * 4, cudaMemcpyHostToDevice));
    QTC_CUDA(cudaMemcpy(dW, W, (size_t)N * K * 2, cudaMemcpyHostToDevice));
    QTC_CUDA(cudaMemset(dC, 0, (size_t)M * N * 4));
    const size_t n = (size_t)TC_BN * K;
    k_to_bf16_rows<<<(unsigned)((n + 255) / 256), 256>>>(dX, dA, M, K, pieces);
    int rc = qtc_gemm_tc(&tmp, dC, N, dX, dW, M, N, K, pieces);
    if (rc == 0 && cudaDeviceSynchronize() != cudaSuccess) { qtc_set_error("gemm_bf16_tc: %s", cudaGetErrorString(cudaGetLastError())); rc = -1; }
    if (rc == 0) QTC_CUDA(cudaMemcpy(C, dC, (size_t)M * N * 4, cudaMemcpyDeviceToHost));
    qtc_gemm_tc_forget(dW); qtc_gemm_tc_forget(dX);
    cudaFree(dA); cudaFree(dC); cudaFree(dW); cudaFree(dX);
    return rc;
}
