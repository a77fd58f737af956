/*
 * qtc_internal.h - host-side view of the device-resident model and state.
 * Layout in HBM is described in DESIGN.md section 3.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "../../include/qwen_tts_cuda.h"

/* ---- error plumbing -------------------------------------------------------- */
void qtc_set_error(const char *fmt, ...);
#define QTC_CUDA(call)                                                                         \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            qtc_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return -1;                                                                         \
        }                                                                                      \
    } while (0)
#define QTC_CUDA_NULL(call)                                                                    \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            qtc_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return NULL;                                                                       \
        }                                                                                      \
    } while (0)

/* launch with the programmatic-stream-serialization attribute (kernels that begin with pdl_enter(), qtc_device.cuh);
 * QTC_PDL=0 in the environment launches them the plain way */
#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
static inline cudaError_t qtc_launch_pdl(cudaStream_t st, void (*k)(KArgs...), dim3 g, dim3 b, size_t smem, Args &&...args) {
    static int on = -1;
    if (on < 0) { const char *e = getenv("QTC_PDL"); on = e ? atoi(e) : 1; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = g; cfg.blockDim = b; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    memset(at, 0, sizeof(at));
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = on ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, k, static_cast<KArgs>(args)...);
}
#endif

/* ---- device-side descriptors (plain structs, passed by value or by pointer) - */
struct LmLayerDev {
    const uint16_t *wq, *wk, *wv, *wo, *gate, *up, *down;   /* bf16, row-major as in the reference */
    const float *q_norm, *k_norm, *in_norm, *post_norm;
    /* mode 2 (qtc_cluster.cuh), built at upload when the geometry fits: wo_t [kv_heads][H][256] = the O projection's columns
     * of one KV group made contiguous; down_t [I/8][H][8] = the down projection in 16-byte pieces, column chunk major */
    const uint16_t *wo_t, *down_t;
};
struct LmDev {
    int vocab, hidden, inter, layers, heads, kv_heads, head_dim;
    float eps, theta;
    LmLayerDev layer[QTC_MAX_LM_LAYERS];
    const float *final_norm;
};

/* per-utterance generation state, resident in HBM so that a captured CUDA graph
 * can be replayed without host-side argument patching */
struct GenState {
    int pos;          /* talker KV length (ctx->talker_kv_len)                      */
    int step;         /* frames generated so far (n_generated)                      */
    int done;         /* 0 running, 1 EOS, 2 budget reached                          */
    int stop_step;
    uint32_t rng;     /* talker RNG: bit pattern of (float)seed, c/qwen_tts.c:1270   */
    int token;        /* last sampled codebook-0 token                               */
    int limit;        /* frame budget of this call                                   */
    int fixed;        /* fixed_codec_tokens > 0                                      */
    int n_trailing;   /* rows in trailing_text                                       */
    int pad_[7];
};

struct SampleCfg {
    int vocab, top_k;
    float top_p, temperature;
    int suppress_from, eos_id;        /* suppress [suppress_from, vocab) except eos_id; -1 = none */
    float rep_penalty;                /* 1.0 = off */
};

/* the talker's paged KV pool (qtc_device.cuh: kv_off): allocated once at upload, never grown or copied.  Slots are
 * sequences (slot 0: the batch-1 utterance and the stage API; slots 0..B-1: the rows of a lock-step batch). */
#define QTC_KV_PAGE_POS 64
struct KvPool {
    int pages = 0;                     /* pages per layer */
    int slots = 0, max_seq_pages = 0;
    int *tables = NULL;                /* device [slots][max_seq_pages] */
    int *tables_host = NULL;           /* pinned mirror */
    std::vector<int> free_pages;       /* stack of free page ids */
    std::vector<int> used;             /* pages held by each slot */
    long long page_allocs = 0;         /* pages handed out so far */
};

/* one decoder stack's mutable buffers */
struct LmBuffers {
    float *k, *v;                 /* [layers, kv_max, kv_heads*head_dim] f32: talker = the pool (kv_max = pool positions per layer) */
    const int *table;             /* talker: page table of slot 0 (device); sub-talker: NULL (contiguous) */
    int kv_max;
    float *cos, *sin;             /* [kv_max, head_dim], host-computed with the oracle's expression */
    float *x, *xn, *q, *kk, *vv, *att, *g, *logits;
    long long *xacc;              /* mode 2: [3][H] fixed-point residual accumulators (zero at upload) */
    int *cl_phase;                /* mode 2: which accumulator is current (device int, carried across launches) */
};

struct CodecDev;   /* qtc_codec.cu */
struct PfBuf;      /* qtc_prefill.cuh: scratch of the batched prefill */
struct BatchBuf;   /* qtc_batch.cu */

struct qtc_ctx {
    qtc_model_desc_t host;        /* dims + host pointers as given (pointers not used after upload) */
    LmDev talker, sub;
    int H, G;
    /* device copies of non-layer tensors */
    const uint16_t *codec_embedding, *text_embedding, *text_fc1_w, *text_fc2_w, *codec_head;
    const float *text_fc1_b, *text_fc2_b;
    const uint16_t *sub_emb[QTC_MAX_GROUPS - 1], *sub_head[QTC_MAX_GROUPS - 1];
    const uint16_t *sub_in_w;
    const float *sub_in_b;
    const float *sub_ptab[QTC_MAX_GROUPS - 1];   /* mode 2: projected sub-talker inputs per code, fp32 [vocab, sub.hidden] per position g >= 1 (qtc_lm.cu) */
    LmBuffers tk, st;
    KvPool kv;
    long long dev_mallocs;        /* cudaMalloc calls made for this context after upload returned (steady state: none) */
    /* generation */
    GenState *state;              /* device */
    GenState *state_host;         /* pinned mirror */
    int *hist_count;              /* [talker.vocab] occurrences of each token in the history */
    int *frame_codes;             /* [QTC_MAX_GROUPS] device: codes of the frame in flight */
    int *codes;                   /* [codes_cap, G] device */
    int codes_cap;
    int *codes_host;              /* pinned */
    float *next_embed;            /* [H] */
    float *trailing;              /* [trailing_cap, H] */
    int trailing_cap;
    float *tts_pad;               /* [H] */
    float *prefill;               /* [prefill_cap, H] */
    int prefill_cap;
    float *st_embed, *st_proj;    /* sub-talker input staging */
    float *text_tmp0, *text_tmp1; /* text projection staging [text_hidden] */
    uint32_t *st_rng;             /* sub-talker RNG (re-seeded every frame) */
    int *zero_pos;                /* device int, always 0: position base of the sub-talker */
    /* fused persistent decode kernel (mode 1) */
    LmDev *talker_dev, *sub_dev;  /* device copies of the stack descriptors */
    unsigned long long *bar;      /* grid barrier counter */
    unsigned *head_cnt;           /* [heads] */
    float *attn_part;             /* [heads, 8, head_dim + 2] */
    int num_sms;
    int prof_sub;                      /* 1: stamp the sub-talker's steps instead of the talker's */
    long long *prof; int prof_layer;   /* optional per-phase clock stamps of the fused kernel (debug) */
    int *sample_scratch;          /* slow-path sampler scratch */
    float *sample_scratch_f;
    float *host_stage;            /* pinned staging for stage-API vectors */
    size_t host_stage_bytes;
    CodecDev *codec;
    PfBuf *pf;
    qtc_gen_params_t params;
    qtc_perf_t perf;
    int mode;                     /* 0 per-op kernels, 1 fused persistent kernel (148 CTAs, 5 exchanges per layer), 2 cluster kernel (default where it fits) */
    int cluster_ok;               /* 8 clusters of 16 CTAs can be co-resident on this device */
    int *cl_err;                  /* device flag: a mode-2 barrier wait gave up */
    int arith;                    /* QTC_FAST / QTC_STRICT */
    cudaStream_t stream;
    cudaStream_t cstream;         /* stream the codec kernels are enqueued on: == stream, or stream2 while a generate overlaps decode and frame loop */
    cudaStream_t stream2;
    cudaEvent_t ev0, ev1, ev2, ev3;   /* ev2..ev3: device span of a whole utterance (prompt build .. last codec kernel) */
    std::vector<void *> allocs;   /* everything cudaMalloc'd, for release() */
    long long launches;
    /* streamed decode inside generate: PCM of the whole utterance on the device, chunk schedule, per-frame / per-chunk events */
    float *pcm_dev; size_t pcm_cap;          /* [frames * samples_per_frame] */
    const float *last_pcm_dev; int last_pcm_n;   /* PCM of the last one-shot decode, still in the codec workspace (qwen_tts_cuda_last_pcm_s16) */
    int16_t *s16_dev; size_t s16_cap;
    float *first_pcm_host;                   /* pinned landing buffer of the streamed PCM (pcm_cap bytes); chunks are handed on from it */
    int stream_first, stream_chunk;          /* frames in the first chunk / in later chunks; stream_chunk <= 0: decode once at the end */
    qtc_pcm_callback_t stream_cb; void *stream_user;
    int codec_tc;                            /* 1 (default): vocoder convolutions on tcgen05 where the decoder's shapes tile; 0: FFMA kernels */
    std::vector<cudaEvent_t> frame_ev, chunk_ev;
    /* CUDA graphs of the frame loop (built lazily, keyed by params) */
    cudaGraphExec_t g_front, g_talker;
    /* mode 2: the talker step's shared-memory K/V stage is sized for the positions the step can see, not for the pages the utterance
     * holds (shared memory is L1 taken away: a stage of 32 instead of 64 positions per CTA is 0.9 % of the step at 190 positions):
     * one captured graph per stage size, 128 positions apart, picked by the frame loop from the frame index */
    cudaGraphExec_t g_talker_b[4];
    int tk_reach_hint;            /* > 0: positions the next talker step(s) can see (set around a capture / a stage-API call) */
    int graph_key;
    int max_batch;
    BatchBuf *bb;                 /* lock-step batch state (qtc_batch.cu), allocated on first use */
    int batch_pieces;             /* activation pieces of the batched GEMMs: 3 (exact split, default) or 1 (bf16) */
};

/* ---- launchers implemented in qtc_lm.cu (all enqueue on ctx->stream) ------- */
struct GemvSeg { const uint16_t *w; float *out; int rows; };
enum { EPI_STORE = 0, EPI_BIAS = 1, EPI_RESID = 2 };
/* out[r] = epi(sum_c W[r,c]*x[c]); up to 3 row segments share x (q/k/v).  aux: bias or residual */
int qtc_gemv(qtc_ctx *c, const GemvSeg *segs, int nseg, const float *x, int cols, int epi, const float *aux);
int qtc_gemv_swiglu(qtc_ctx *c, float *out, const uint16_t *gate, const uint16_t *up, const float *x, int inter, int cols);
int qtc_rmsnorm(qtc_ctx *c, float *out, const float *x, const float *w, int n, float eps);
int qtc_lm_step(qtc_ctx *c, const LmDev *lm, LmBuffers *b, const int *pos_ptr, int pos_add);
int qtc_embed_row(qtc_ctx *c, float *out, const uint16_t *table, const int *idx_ptr, int idx_imm, int n);
int qtc_sample(qtc_ctx *c, const float *logits, const SampleCfg &cfg, uint32_t *rng, int *hist_count,
               int *out_token, GenState *st, int role);
int qtc_ensure_kv(qtc_ctx *c, LmBuffers *b, const LmDev *lm, int need);
int qtc_prefill_gemm(qtc_ctx *c, const float *embeds_dev, int seq_len, int base);   /* qtc_prefill.cuh */
bool qtc_prefill_gemm_ok(const qtc_ctx *c);
void qtc_prefill_release(qtc_ctx *c);
int qtc_kv_pool_init(qtc_ctx *c, int max_batch);
int qtc_cluster_prepare(qtc_ctx *c, LmDev *d);
int qtc_build_proj_tables(qtc_ctx *c);              /* sub-talker input tables for mode 2 (after the embeddings are uploaded) */      /* mode 2 weight layouts (qtc_lm.cu); 0 also when the geometry does not fit */
int qtc_kv_reserve(qtc_ctx *c, int slot, int positions);      /* the slot holds pages for at least `positions` positions */
void qtc_kv_release_all(qtc_ctx *c);
#define QTC_COUNT_MALLOC(c) ((c)->dev_mallocs++)
int qtc_talker_forward_dev(qtc_ctx *c, const float *embed_dev);   /* one decode step + head -> tk.logits */
int qtc_subtalker_dev(qtc_ctx *c, const float *hidden_dev, const int *first_code_ptr, int *codes_out_dev);
int qtc_sample_talker(qtc_ctx *c);
int qtc_frame_front(qtc_ctx *c);    /* sample + sub-talker + next embedding, device resident */
struct StreamOut {                 /* streamed decode of one utterance */
    float *host; int spf, samples;     /* the caller's PCM buffer (malloc), samples per frame, samples written */
    int decoded = 0, nchunks = 0, delivered = 0;
    std::vector<int> chunk_end;        /* frame index each chunk ends at */
};
int qtc_generate_frames_resident(qtc_ctx *c, int prefill_len, int n_trailing, int *out_codes, int limit, StreamOut *so);
int qtc_build_prompt_dev(qtc_ctx *c, const int *tok, int ntok, int spk, int lang, float *pre, int *plen_out, float *trailing, int *ntr_out, float *pad);
void qtc_batch_release(qtc_ctx *c);
void qtc_batch_invalidate_graphs(qtc_ctx *c);
void qtc_gemm_tc_forget(const void *ptr);
void *qtc_dmalloc(qtc_ctx *c, size_t bytes);
template <class T> T *qtc_upload(qtc_ctx *c, const T *host, size_t n);

/* ---- vocoder convolutions on tcgen05 (qtc_conv_tc.cu) ------------------------ */
struct ConvTcW { const uint16_t *hi, *lo; int cin, cout, taps, phases; };   /* weight planes [phases*taps*cout, cin] bf16 */
bool qtc_conv_tc_ok(int cin, int cout);
int qtc_conv_tc_prep(qtc_ctx *c, ConvTcW *o, const float *w, int cout, int cin, int K);
int qtc_tconv_tc_prep(qtc_ctx *c, ConvTcW *o, const float *w, int cin, int cout, int k, int s);
int qtc_conv_tc(qtc_ctx *c, const ConvTcW *w, const uint16_t *in_hi, const uint16_t *in_lo, int L, int dil, const float *bias, const float *res,
                float *out_raw, uint16_t *out_hi, uint16_t *out_lo, const float *alpha, const float *inv_beta, const float *scale = NULL, int gelu = 0,
                const uint16_t *in_x = NULL);   /* in_x: third activation plane (exact 3-piece split; bf16 weights only) */
int qtc_rows_to_planes(qtc_ctx *c, uint16_t *hi, uint16_t *lo, const float *src, size_t n);
int qtc_dwconv_tm(qtc_ctx *c, float *out, const float *in, const float *w, const float *bias, int C, int K, int L);
int qtc_cf_to_planes(qtc_ctx *c, uint16_t *hi, uint16_t *lo, const float *src, int C, int L, const float *alpha, const float *inv_beta);
int qtc_mono_from_planes(qtc_ctx *c, float *out, const uint16_t *hi, const uint16_t *lo, const float *w, const float *bias, int C, int K, int L);

/* ---- codec (qtc_codec.cu) --------------------------------------------------- */
CodecDev *qtc_codec_upload(qtc_ctx *c, const qtc_codec_t *k);
void qtc_codec_release(CodecDev *cd);
float *qtc_codec_decode_dev(qtc_ctx *c, const int *codes_dev, int T, int *out_samples, float **dev_out);
int qtc_codec_stream_begin(qtc_ctx *c, int max_frames);
int qtc_codec_stream_reserve(qtc_ctx *c, int chunk);
int qtc_codec_stream_push(qtc_ctx *c, const int *codes_dev, int n, float **pcm_dev);

/* small launchers (qtc_lm.cu) */
int qtc_copy(qtc_ctx *c, float *dst, const float *src, int n);
int qtc_silu(qtc_ctx *c, float *x, int n);
int qtc_add_row(qtc_ctx *c, float *out, const float *a, const uint16_t *table, int idx, int n);
int qtc_set_int(qtc_ctx *c, int *p, int v);
int qtc_bump_pos(qtc_ctx *c);
int qtc_sample_bare(qtc_ctx *c, const float *logits_dev, int vocab, int top_k, float top_p, float temperature,
                    uint32_t *rng_dev, int *out_dev);
extern "C" int qwen_tts_cuda_set_arith(qtc_ctx_t *c, int strict);
