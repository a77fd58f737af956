/*
 * qwen_tts_cuda.h - C ABI of the B200 (sm_100a) backend for Qwen3-TTS-C's
 * per-frame generation hot path.
 *
 * This is the drop-in boundary: plain C, plain pointers and sizes, no C++ or
 * torch types.  Every entry point names the reference interface it replaces
 * (paths are relative to the reference repository, muonium-ai/Qwen3-TTS-C).
 *
 * Shape of the integration (see INTEGRATION.md for the reference-side stub):
 *   kernel_init()            (c/qwen_tts_kernels.c:978, called c/qwen_tts.c:914)  -> qwen_tts_cuda_init()
 *   end of qwen_tts_load()   (c/qwen_tts.c:862-921)                                -> qwen_tts_cuda_upload()
 *   qwen_tts_free()          (c/qwen_tts.c:927)                                    -> qwen_tts_cuda_release()
 *   the four stage functions (c/qwen_tts.h:483-502)                                -> qwen_tts_cuda_talker_prefill/
 *                                                                                    _talker_forward/_subtalker_generate/_codec_decode
 *   the per-frame loop       (c/qwen_tts.c:1282-1373)                              -> qwen_tts_cuda_generate_frames()
 *   qwen_tts_generate()      (c/qwen_tts.c:1059-1443)                              -> qwen_tts_cuda_generate()
 *
 * All pointers in qtc_model_desc_t are HOST pointers owned by the caller (in
 * the reference they point into the safetensors mmap or into malloc'd f32
 * copies, c/qwen_tts.h:150-350).  upload() copies everything to HBM; the host
 * memory is not referenced afterwards.
 *
 * Conventions: "bf16" is the raw upper half of an IEEE f32 (c/qwen_tts.h:475-480).
 * All matrices are row-major [rows, cols].  int returns are 0 on success, -1 on
 * failure (message on stderr), like metal_init() in metal/qwen_tts_metal.h:23.
 */
#ifndef QWEN_TTS_CUDA_H
#define QWEN_TTS_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QTC_MAX_LM_LAYERS     32   /* QWEN_TTS_MAX_TALKER_LAYERS, c/qwen_tts.h:59 */
#define QTC_MAX_CODEC_LAYERS  12   /* QWEN_TTS_MAX_CODEC_LAYERS,  c/qwen_tts.h:61 */
#define QTC_MAX_GROUPS        32   /* QWEN_TTS_NUM_CODE_GROUPS,   c/qwen_tts.h:35 */
#define QTC_MAX_QUANTIZERS    16   /* QWEN_TTS_CODEC_NUM_QUANTIZERS, c/qwen_tts.h:47 */
#define QTC_MAX_NAMES         64

/* ---- one decoder layer of the talker or the sub-talker --------------------
 * mirrors qwen_tts_talker_layer_t / qwen_tts_subtalker_layer_t (c/qwen_tts.h:150-172,199-212).
 * gate/up are the two halves of the reference's gate_up_fused_bf16 [2I,H]. */
typedef struct {
    const uint16_t *wq;             /* bf16 [heads*head_dim, hidden]    */
    const uint16_t *wk;             /* bf16 [kv_heads*head_dim, hidden] */
    const uint16_t *wv;             /* bf16 [kv_heads*head_dim, hidden] */
    const uint16_t *wo;             /* bf16 [hidden, heads*head_dim]    */
    const float    *q_norm;         /* f32 [head_dim] */
    const float    *k_norm;         /* f32 [head_dim] */
    const float    *input_norm;     /* f32 [hidden]   */
    const float    *post_attn_norm; /* f32 [hidden]   */
    const uint16_t *gate;           /* bf16 [intermediate, hidden] */
    const uint16_t *up;             /* bf16 [intermediate, hidden] */
    const uint16_t *down;           /* bf16 [hidden, intermediate] */
} qtc_lm_layer_t;

/* A decoder stack: the talker (c/qwen_tts.h:174-193) or the sub-talker (:214-230). */
typedef struct {
    int   vocab, hidden, intermediate, layers, heads, kv_heads, head_dim;
    float rms_eps;       /* both stacks use talker_rms_norm_eps (c/qwen_tts_talker.c:555) */
    float rope_theta;    /* both stacks use talker_rope_theta  (c/qwen_tts_talker.c:619) */
    qtc_lm_layer_t layer[QTC_MAX_LM_LAYERS];
    const float *final_norm;    /* f32 [hidden] */
} qtc_lm_t;

/* ---- codec decoder (c/qwen_tts.h:236-350); everything f32 ---------------- */
typedef struct {
    const float *input_norm, *post_attn_norm;        /* [hidden] */
    const float *attn_layer_scale, *mlp_layer_scale; /* [hidden] or NULL */
    const float *wq, *wk, *wv, *wo;                  /* [hidden,hidden]-shaped, see c/qwen_tts.h:264-268 */
    const float *gate, *up, *down;                   /* [inter,hidden],[inter,hidden],[hidden,inter] */
} qtc_codec_layer_t;

typedef struct {
    const float *dwconv_w, *dwconv_b;   /* [dim,1,7], [dim] */
    const float *norm_w, *norm_b;       /* [dim] */
    const float *pw1_w, *pw1_b;         /* [4dim,dim], [4dim] */
    const float *pw2_w, *pw2_b;         /* [dim,4dim], [dim] */
    const float *gamma;                 /* [dim] */
} qtc_convnext_t;

typedef struct {
    /* alpha/beta are the PRE-PROCESSED values the reference keeps after load:
     * alpha=expf(a_log), beta=1/(expf(b_log)+1e-9)  (c/qwen_tts.c:596-602) */
    const float *act1_alpha, *act1_beta;  /* [dim] */
    const float *conv1_w, *conv1_b;       /* [dim,dim,7], [dim] */
    const float *act2_alpha, *act2_beta;  /* [dim] */
    const float *conv2_w, *conv2_b;       /* [dim,dim,1], [dim] */
} qtc_resunit_t;

typedef struct {
    const float *act_alpha, *act_beta;    /* [in_dim], pre-processed as above */
    const float *tconv_w, *tconv_b;       /* [in_dim,out_dim,2*rate], [out_dim] */
    qtc_resunit_t res[3];                 /* dilations 1,3,9 */
} qtc_vocoder_block_t;

typedef struct {
    int num_quantizers, codebook_size, codebook_dim, hidden, latent, layers;
    int heads, kv_heads, intermediate, sliding_window, decoder_dim;
    float rms_eps;
    int upsample_rates[4];      /* vocoder strides, default 8,5,4,3 */
    int upsampling_ratios[2];   /* ConvNeXt stage strides, default 2,2 */

    /* codebooks[q] = embedding_sum / max(usage,1e-5), [codebook_size, codebook_dim/2]
     * (qwen_tts_codebook_t.embeddings, c/qwen_tts.c:577-594). q=0 semantic, 1.. acoustic */
    const float *codebooks[QTC_MAX_QUANTIZERS];
    const float *semantic_out_proj;   /* [latent/2, codebook_dim/2] */
    const float *acoustic_out_proj;   /* [latent/2, codebook_dim/2] */
    const float *pre_conv_w, *pre_conv_b;       /* [latent, codebook_dim, 3], [latent] */
    const float *in_proj_w, *in_proj_b;         /* [hidden, latent], [hidden] */
    const float *out_proj_w, *out_proj_b;       /* [latent, hidden], [latent] */
    qtc_codec_layer_t layer[QTC_MAX_CODEC_LAYERS];
    const float *final_norm;                    /* [hidden] */
    const float *up_tconv_w[2], *up_tconv_b[2]; /* [latent,latent,r], [latent] */
    qtc_convnext_t convnext[2];
    const float *voc_pre_w, *voc_pre_b;         /* [decoder_dim, latent, 7], [decoder_dim] */
    qtc_vocoder_block_t block[4];
    const float *final_alpha, *final_beta;      /* [decoder_dim/16], pre-processed */
    const float *final_conv_w, *final_conv_b;   /* [1, decoder_dim/16, 7], [1] */
} qtc_codec_t;

/* ---- whole model: the subset of qwen_tts_ctx_t the hot path reads --------- */
typedef struct {
    qtc_lm_t talker;
    int   text_hidden, text_vocab, num_code_groups;
    int   mrope_section[3];
    const uint16_t *codec_embedding;    /* bf16 [talker.vocab, talker.hidden] */
    const uint16_t *text_embedding;     /* bf16 [text_vocab, text_hidden] */
    const uint16_t *text_fc1_w;         /* bf16 [text_hidden, text_hidden] */
    const float    *text_fc1_b;         /* f32 [text_hidden] or NULL */
    const uint16_t *text_fc2_w;         /* bf16 [talker.hidden, text_hidden] */
    const float    *text_fc2_b;         /* f32 [talker.hidden] or NULL */
    const uint16_t *codec_head;         /* bf16 [talker.vocab, talker.hidden] */

    qtc_lm_t sub;
    const uint16_t *sub_codec_embeddings[QTC_MAX_GROUPS - 1]; /* bf16 [sub.vocab, talker.hidden] each */
    const uint16_t *sub_input_proj_w;   /* bf16 [sub.hidden, talker.hidden] or NULL */
    const float    *sub_input_proj_b;   /* f32 [sub.hidden] or NULL */
    const uint16_t *sub_lm_heads[QTC_MAX_GROUPS - 1];         /* bf16 [sub.vocab, sub.hidden] each */

    qtc_codec_t codec;

    /* codec-domain special ids (qwen_tts_config_t, c/qwen_tts.h:136-143) */
    int codec_pad_id, codec_bos_id, codec_eos_id, codec_nothink_id;
    int codec_think_id, codec_think_bos_id, codec_think_eos_id;
    /* text-domain special ids (c/qwen_tts.h:64-69) */
    int tts_pad_id, tts_bos_id, tts_eos_id;

    /* speaker / language maps (c/qwen_tts.h:128-134); names compared case-insensitively */
    int n_speakers;  const char *speaker_names[QTC_MAX_NAMES];  int speaker_ids[QTC_MAX_NAMES];
    int n_languages; const char *language_names[QTC_MAX_NAMES]; int language_ids[QTC_MAX_NAMES];
} qtc_model_desc_t;

/* Generation parameters: the public fields of qwen_tts_ctx_t that main.c pokes
 * (c/qwen_tts.h:420-430, defaults c/qwen_tts.c:871-880). */
typedef struct {
    float temperature, subtalker_temperature;
    int   top_k, subtalker_top_k;
    float top_p, subtalker_top_p;
    float repetition_penalty;
    int   max_new_tokens, fixed_codec_tokens, sample_seed;
} qtc_gen_params_t;

/* perf_* fields of qwen_tts_ctx_t (c/qwen_tts.h:436-440) plus device-side detail */
typedef struct {
    double total_ms, talker_ms, codec_ms, prefill_ms;
    int    codec_tokens;
    int    stop_reason;   /* 1 eos, 2 max_tokens  (c/qwen_tts.c:1264) */
    int    stop_step;
    long long kernel_launches;  /* kernels of this library launched so far on this context */
    double device_ms;           /* whole utterance on the device: prompt build .. last codec kernel (CUDA events) */
    /* latency of the last qwen_tts_cuda_generate (device clock, from the start of the prompt build) */
    double ttff_ms;             /* time to the first frame's 16 codes (prompt build + prefill + first frame) */
    double ttfa_ms;             /* time to first audio: first PCM chunk in host memory (0 if the decode was not streamed) */
    double frame_ms_p50, frame_ms_p99;   /* frame-to-frame interval (includes codec chunks that interleave) */
    /* memory: the talker's paged KV pool (pages of 64 positions, allocated once at upload) and every cudaMalloc after upload */
    long long device_mallocs;   /* cudaMalloc calls since upload returned: workspaces grown on first use; 0 in steady state */
    long long kv_page_allocs;   /* pages handed out by the pool's free list so far */
    int    kv_pool_pages, kv_pages_in_use;
    double upload_ms, upload_bytes;   /* qwen_tts_cuda_upload: wall time and bytes copied host -> device (pinned staging) */
    int    queue_admitted;      /* qwen_tts_cuda_generate_queue: utterances admitted into a row (the first `rows` at the start, the others into freed rows) */
} qtc_perf_t;

/* streamed PCM: called as chunks land in host memory (see qwen_tts_cuda_set_stream_callback) */
typedef void (*qtc_pcm_callback_t)(const float *pcm, int n_samples, int first_frame, void *user);

typedef struct qtc_ctx qtc_ctx_t;   /* opaque device-side model + state */

/* ---- lifecycle ------------------------------------------------------------ */
int  qwen_tts_cuda_init(int device);         /* kernel_init hook; -1 if no sm_100 device */
int  qwen_tts_cuda_is_available(void);       /* cf. metal_is_available() */
void qwen_tts_cuda_shutdown(void);
void qwen_tts_cuda_print_info(void);
void qwen_tts_cuda_default_params(qtc_gen_params_t *p);   /* c/qwen_tts.c:871-880 */

qtc_ctx_t *qwen_tts_cuda_upload(const qtc_model_desc_t *desc, int max_batch);  /* NULL on failure */
void       qwen_tts_cuda_release(qtc_ctx_t *ctx);
void       qwen_tts_cuda_set_params(qtc_ctx_t *ctx, const qtc_gen_params_t *p);
void       qwen_tts_cuda_get_perf(const qtc_ctx_t *ctx, qtc_perf_t *out);
/* 0: separate kernels per op (baseline)
 * 1: one persistent 148-CTA kernel per decoder-stack step, five grid-wide exchanges per layer (round 1's path; fallback)
 * 2 (default): the cluster kernel - 15 clusters of 8 CTAs, two grid-wide exchanges per layer - where the stack's geometry fits
 *    (16 query / 8 KV heads of 128, hidden 1024 or 2048, enough 8-CTA clusters on the device); other stacks run mode 1 */
int        qwen_tts_cuda_set_mode(qtc_ctx_t *ctx, int mode);
/* Arithmetic of the talker / sub-talker path.
 * 0 (default): fp32 with FMA and tree reductions - weight-streaming speed; differs from the C path by
 *    fp32 summation order only (logits within ~1e-5 relative).
 * 1: "strict" - every sum in the reference's order (k ascending, multiply then add, no FMA; expf by
 *    glibc's algorithm): logits, hidden states and therefore sampled tokens are BIT-IDENTICAL to the
 *    non-BLAS C build.  Slower: one dependent add per weight column. */
int        qwen_tts_cuda_set_arith(qtc_ctx_t *ctx, int strict);
/* the same switch for the kernel-level entry points below */
int        qwen_tts_cuda_kernel_arith(int strict);

/* ---- stage API: same host-visible semantics as c/qwen_tts.h:483-502 ------- */
/* resets nothing: caller sets kv_len=0 via _reset() like c/qwen_tts.c:1249 */
void qwen_tts_cuda_reset(qtc_ctx_t *ctx);
int  qwen_tts_cuda_talker_prefill(qtc_ctx_t *ctx, const float *input_embeds, int seq_len);
int  qwen_tts_cuda_talker_forward(qtc_ctx_t *ctx, const float *input_embed, float *logits);
int  qwen_tts_cuda_subtalker_generate(qtc_ctx_t *ctx, const float *talker_hidden,
                                      int first_code, int *out_codes);
float *qwen_tts_cuda_codec_decode(qtc_ctx_t *ctx, const int *codes, int time_steps,
                                  int *out_samples);   /* malloc'd, caller frees */
/* ctx side effects the reference exposes as fields: ctx->tk_x, ctx->talker_kv_len */
int  qwen_tts_cuda_get_hidden(qtc_ctx_t *ctx, float *tk_x_out);   /* f32 [talker.hidden] */
int  qwen_tts_cuda_get_kv_len(const qtc_ctx_t *ctx);
/* logits of "step 0": codec_head . tk_x  (c/qwen_tts.c:1284-1296) */
int  qwen_tts_cuda_head_logits(qtc_ctx_t *ctx, float *logits);

/* ---- prompt side (c/qwen_tts.c:823-856): text token -> projected embedding - */
int  qwen_tts_cuda_embed_text_tokens(qtc_ctx_t *ctx, const int *token_ids, int n, float *out /*[n,hidden]*/);

/* ---- fused device-resident frame loop (c/qwen_tts.c:1258-1373) -------------
 * prefill_embeds [prefill_len,hidden], trailing_text [n_trailing,hidden],
 * tts_pad [hidden]: what qwen_tts_generate builds at c/qwen_tts.c:1147-1243.
 * out_codes [max_tokens, num_code_groups] int32.  Returns n_generated or -1. */
int  qwen_tts_cuda_generate_frames(qtc_ctx_t *ctx,
                                   const float *prefill_embeds, int prefill_len,
                                   const float *trailing_text, int n_trailing,
                                   const float *tts_pad,
                                   int *out_codes, int max_tokens);

/* ---- whole utterance: qwen_tts_generate (c/qwen_tts.c:1059-1443) ------------
 * text = comma-separated token ids (>= 8 ids), speaker/language names or NULL.
 * Returns malloc'd f32 PCM @24 kHz (caller frees) or NULL with *out_samples=0.
 * If out_codes != NULL it receives up to max_codes*num_code_groups ints. */
float *qwen_tts_cuda_generate(qtc_ctx_t *ctx, const char *text, const char *speaker,
                              const char *language, int *out_samples,
                              int *out_codes, int max_codes, int *out_n_frames);

/* ---- streaming decode ---------------------------------
 * The C reference decodes once after the last frame (c/qwen_tts.c:1399-1420); the Python package streams 300-frame chunks
 * with 25 frames of left context (qwen_tts/core/tokenizer_12hz/modeling_qwen3_tts_tokenizer_v2.py:886-896).  Here every
 * layer's causality is used for an EXACT chunked decode: streamed PCM equals the one-shot PCM bit for bit.
 * With qwen_tts_cuda_set_streaming(ctx, first, chunk), qwen_tts_cuda_generate decodes chunks of finished frames between
 * two frames of the loop (first chunk `first` frames, later ones `chunk`) and copies their PCM to the host while the loop
 * goes on: time to first audio is prefill + `first` frames + one small decode instead of the whole utterance.  Every chunk
 * pays the decoder's fixed cost once more, so throughput is a few per cent lower than with the default, chunk_frames <= 0:
 * the reference's decode-at-the-end order. */
int  qwen_tts_cuda_set_streaming(qtc_ctx_t *ctx, int first_frames, int chunk_frames);
/* vocoder convolutions (c/qwen_tts_kernels.c:659-972): 1 (default) tcgen05 implicit GEMMs, 0 fp32 FFMA kernels */
int  qwen_tts_cuda_set_codec_tc(qtc_ctx_t *ctx, int on);
void qwen_tts_cuda_set_stream_callback(qtc_ctx_t *ctx, qtc_pcm_callback_t cb, void *user);
/* stage-level: a decode session of up to max_frames frames fed chunk by chunk.  push: codes [n, num_quantizers] (host) are
 * the NEXT n frames, pcm_out (host) receives exactly their n * 1920 samples; returns the sample count or -1. */
int  qwen_tts_cuda_codec_stream_begin(qtc_ctx_t *ctx, int max_frames);
int  qwen_tts_cuda_codec_stream_push(qtc_ctx_t *ctx, const int *codes, int n, float *pcm_out);

/* ---- batched, request-sharded variant (BASELINE configs 4-5) ---------------
 * batch <= 64 independent utterances decoded in lock-step on this GPU (more: call again, or shard over GPUs);
 * projections are tcgen05 GEMMs (one pass over the weights per step for the whole batch).  Utterance b uses seed
 * params.sample_seed + b.  token_ids is a ragged list: offsets[batch+1]; speaker_ids / language_ids are codec ids
 * (qwen_tts_config_t.spk_id / codec_language_id values) or -1, or NULL.
 * out_codes [batch, max_tokens, groups]; out_n_frames [batch] (optional); pcm_out (optional) [batch, max_tokens*1920]. */
int  qwen_tts_cuda_generate_batch(qtc_ctx_t *ctx, int batch,
                                  const int *token_ids, const int *offsets,
                                  const int *speaker_ids, const int *language_ids,
                                  int *out_codes, int *out_n_frames,
                                  float *pcm_out, int max_tokens);
/* Continuous batching over the same lock-step rows: n_utt utterances queue for `rows` <= 64 rows (<= the max_batch of
 * qwen_tts_cuda_upload).  The reference stops an utterance at EOS (c/qwen_tts.c:1324-1330) and is batch 1; in a lock-step
 * batch utterances end at different frames, so a row whose utterance has ended is retired at the next poll (every 8 frames;
 * with fixed_codec_tokens exactly at the budget): its codes are saved, its KV pages return to the pool's free list, and the
 * next queued utterance is admitted into the row (pages reserved again, prompt built and prefilled on its own, per-row
 * state / RNG / repetition history reset) while the other rows are mid-utterance.  Utterance u uses seed
 * params.sample_seed + u and produces what it produces alone.  Buffers as for generate_batch with n_utt in place of batch.
 * frame_budgets (optional, [n_utt]): per-utterance frame limit (the reference's max_new_tokens / fixed_codec_tokens are per
 * context, c/qwen_tts.h:420-430; a queue of requests has one each), capped by the call's own limit.
 * Needs fast arithmetic (the default). */
int  qwen_tts_cuda_generate_queue(qtc_ctx_t *ctx, int n_utt, int rows,
                                  const int *token_ids, const int *offsets,
                                  const int *speaker_ids, const int *language_ids, const int *frame_budgets,
                                  int *out_codes, int *out_n_frames, float *pcm_out, int max_tokens);

/* activations of the batched GEMMs: 3 = exact split of every fp32 value into three bf16 pieces (default: results differ
 * from the batch-1 fast path by summation order only), 1 = rounded to bf16 (logits within 1e-2 relative) */
int  qwen_tts_cuda_set_batch_pieces(qtc_ctx_t *ctx, int pieces);
/* stage API of the batched path (host buffers), the batched forms of c/qwen_tts.h:486-494 */
int  qwen_tts_cuda_batch_reset(qtc_ctx_t *ctx, int batch, int kv_max);
int  qwen_tts_cuda_batch_forward(qtc_ctx_t *ctx, int batch, const float *embeds /*[B,hidden]*/,
                                 float *logits /*[B,vocab] or NULL*/, float *hidden /*[B,hidden] or NULL*/);
int  qwen_tts_cuda_batch_subtalker(qtc_ctx_t *ctx, int batch, const float *hidden /*[B,hidden]*/,
                                   const int *first_codes /*[B]*/, int *out_codes /*[B,groups]*/);

/* ---- kernel-level test surface: signatures of c/qwen_tts_kernels.h:27-131 ---
 * host pointers in/out (one H2D/D2H per call: for parity tests, not for speed) */
int qwen_tts_cuda_matvec_bf16(float *out, const uint16_t *A_bf16, const float *x, int rows, int cols);
int qwen_tts_cuda_swiglu_matvec_bf16(float *out, const uint16_t *gate_up_fused, const float *x,
                                     int intermediate, int hidden);
int qwen_tts_cuda_rms_norm(float *out, const float *x, const float *w, int dim, float eps);
int qwen_tts_cuda_sample_top_k(const float *logits, int vocab, int top_k, float top_p,
                               float temperature, float *rng_state);   /* returns token */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  The same sampling kernel as
 * qwen_tts_cuda_sample_top_k, allocating nothing: logits_dev [vocab] f32, rng_state_dev (1 x f32 holding the xorshift
 * bits, advanced in place) and token_dev (1 x int32, receives the token) are DEVICE pointers of the caller; the launch is
 * queued on cuda_stream (a cudaStream_t; NULL = the default stream) and NOT synchronised.  The top-p / large-k branches'
 * scratch is a static device array of QTC_NOALLOC_MAX_VOCAB (16384) entries = 128 KB per GPU, held until exit: vocab
 * beyond that is rejected with -1 and a message in qwen_tts_cuda_last_error().  Every call shares the array, so calls
 * that overlap on the GPU race - serialise them.  Returns 0 when the launch was queued. */
#define QTC_NOALLOC_MAX_VOCAB 16384
int qwen_tts_cuda_sample_top_k_noalloc(const float *logits_dev, int vocab, int top_k, float top_p, float temperature,
                                       float *rng_state_dev, int *token_dev, void *cuda_stream);
int qwen_tts_cuda_causal_conv1d(float *out, const float *in, const float *w, const float *b,
                                int cin, int cout, int k, int length, int dilation, int groups);
int qwen_tts_cuda_transposed_conv1d(float *out, const float *in, const float *w, const float *b,
                                    int cin, int cout, int k, int stride, int length, int *out_length);
int qwen_tts_cuda_snake_beta(float *out, const float *x, const float *alpha, const float *beta,
                             int channels, int length);
int qwen_tts_cuda_matmul_f32(float *C, const float *A, const float *B, int M, int N, int K);
int qwen_tts_cuda_matmul_bf16(float *C, const float *A, const uint16_t *B_bf16, int M, int N, int K);
/* the batched path's projection kernel on its own (tcgen05 / TMEM / TMA): C[M,N] = A[M,K] . W[N,K]^T, M <= 64, K % 64 == 0.
 * pieces = 1: A rounded to bf16; pieces = 3: A split exactly into three bf16 pieces (fp32-accurate result) */
int qwen_tts_cuda_gemm_bf16_tc(float *C, const float *A, const uint16_t *W_bf16, int M, int N, int K, int pieces);

/* ---- measurement hooks ------------------------------------------------------ */
/* time `iters` talker decode steps at KV position `pos` with CUDA events on the
 * launch stream; returns mean ms per step (inputs resident in HBM). */
double qwen_tts_cuda_bench_talker_step(qtc_ctx_t *ctx, int pos, int iters, int flush_l2);
double qwen_tts_cuda_bench_subtalker(qtc_ctx_t *ctx, int iters, int flush_l2);
/* algorithmic bytes the talker step / sub-talker frame read */
double qwen_tts_cuda_talker_step_bytes(const qtc_ctx_t *ctx, int pos);
double qwen_tts_cuda_subtalker_frame_bytes(const qtc_ctx_t *ctx);
/* per-phase SM-clock stamps of the fused talker step (layer `layer`, KV position `pos`): 16 slots per CTA,
 * out[num_sms*16]; returns the CTA count.  Evidence for profiles/, not part of the reference's interface. */
int qwen_tts_cuda_debug_phase_clocks(qtc_ctx_t *ctx, int pos, int layer, long long *out);
/* 16-bit PCM of the last decode (qwen_tts_cuda_generate / codec_decode), converted on the device exactly as the reference's WAV
 * writer does on the host (c/qwen_tts_audio.c:70-82: clamp to [-1, 1], times 32767.0f, truncate) - half the device-to-host bytes
 * of the float PCM.  Returns the samples written (<= max_samples) or -1. */
int qwen_tts_cuda_last_pcm_s16(qtc_ctx_t *ctx, int16_t *out, int max_samples);
/* mode 2 (cluster kernel) diagnostics: *capacity = 8-CTA clusters the device holds at once (-1 before the first mode-2 step;
 * mode 2 needs kv_heads of them), *gave_up = 1 if a grid-barrier wait of that kernel ever hit its spin bound.  No reference counterpart. */
int qwen_tts_cuda_cluster_status(qtc_ctx_t *ctx, int *capacity, int *gave_up);
const char *qwen_tts_cuda_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* QWEN_TTS_CUDA_H */
