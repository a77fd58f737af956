/*
 * ref_shim.c - thin test shim over the UNMODIFIED reference (TEST INFRASTRUCTURE ONLY).
 * Linked with the reference's own c/*.c into oracle/_ref/libqtts_ref.so by
 * oracle/Makefile.  Exposes the stage API (c/qwen_tts.h:482-502) and the ctx
 * fields a test needs, so Python can drive it through ctypes.
 *
 * The reference CLI never prints codebooks 1..15 (c/qwen_tts.c:1391-1397), so
 * the Makefile compiles c/qwen_tts.c with -Dqwen_tts_codec_decode=ref_hook_codec_decode:
 * the call at c/qwen_tts.c:1421 then lands here, where the codes are recorded
 * before the real decoder runs.  No reference source is edited.
 */
#include "qwen_tts.h"
#include "qwen_tts_cuda.h"
#include <stdlib.h>
#include <string.h>

void qwen_tts_cuda_fill_desc(const qwen_tts_ctx_t *ctx, qtc_model_desc_t *d);

static int *g_codes = NULL; static int g_frames = 0, g_groups = 0;

float *ref_hook_codec_decode(qwen_tts_ctx_t *ctx, const int *codes, int T, int *out_samples) {
    free(g_codes);
    g_groups = ctx->config.codec_num_quantizers; g_frames = T;
    g_codes = (int *)malloc((size_t)T * g_groups * sizeof(int));
    memcpy(g_codes, codes, (size_t)T * g_groups * sizeof(int));
    return qwen_tts_codec_decode(ctx, codes, T, out_samples);
}
int ref_last_frames(void) { return g_frames; }
int ref_last_groups(void) { return g_groups; }
void ref_last_codes(int *dst) { memcpy(dst, g_codes, (size_t)g_frames * g_groups * sizeof(int)); }

void *ref_load(const char *dir, int verbose) { qwen_tts_verbose = verbose; return qwen_tts_load(dir); }
void ref_free(void *c) { qwen_tts_free((qwen_tts_ctx_t *)c); }
void ref_fill_desc(void *c, qtc_model_desc_t *d) { qwen_tts_cuda_fill_desc((qwen_tts_ctx_t *)c, d); }
void ref_set_params(void *c, const qtc_gen_params_t *p) {
    qwen_tts_ctx_t *x = (qwen_tts_ctx_t *)c;
    x->temperature = p->temperature; x->subtalker_temperature = p->subtalker_temperature;
    x->top_k = p->top_k; x->subtalker_top_k = p->subtalker_top_k;
    x->top_p = p->top_p; x->subtalker_top_p = p->subtalker_top_p;
    x->repetition_penalty = p->repetition_penalty; x->max_new_tokens = p->max_new_tokens;
    x->fixed_codec_tokens = p->fixed_codec_tokens; x->sample_seed = p->sample_seed;
}
void ref_reset(void *c) { ((qwen_tts_ctx_t *)c)->talker_kv_len = 0; }
void ref_prefill(void *c, const float *e, int n) { qwen_tts_talker_prefill((qwen_tts_ctx_t *)c, e, n); }
void ref_forward(void *c, const float *e, float *logits) { qwen_tts_talker_forward((qwen_tts_ctx_t *)c, e, logits); }
void ref_subtalker(void *c, const float *h, int code, int *out) { qwen_tts_subtalker_generate((qwen_tts_ctx_t *)c, h, code, out); }
float *ref_codec_decode(void *c, const int *codes, int T, int *n) { return qwen_tts_codec_decode((qwen_tts_ctx_t *)c, codes, T, n); }
const float *ref_hidden(void *c) { return ((qwen_tts_ctx_t *)c)->tk_x; }
int ref_kv_len(void *c) { return ((qwen_tts_ctx_t *)c)->talker_kv_len; }
float *ref_generate(void *c, const char *csv, const char *spk, const char *lang, int *n) {
    return qwen_tts_generate((qwen_tts_ctx_t *)c, csv, spk, lang, n);
}
void ref_perf(void *c, double *out4) {
    qwen_tts_ctx_t *x = (qwen_tts_ctx_t *)c;
    out4[0] = x->perf_total_ms; out4[1] = x->perf_talker_ms; out4[2] = x->perf_codec_ms; out4[3] = x->perf_codec_tokens;
}
void ref_free_buf(void *p) { free(p); }
