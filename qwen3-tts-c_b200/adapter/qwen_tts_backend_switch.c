/*
 * qwen_tts_backend_switch.c - reference-side glue: the outer API of c/qwen_tts.h:448-465
 * (qwen_tts_load / qwen_tts_free / qwen_tts_generate) with a backend switch, WITHOUT editing a
 * reference source ("switch with a backend flag").
 *
 * How it links (oracle/Makefile, target _ref/qwen-tts-cuda): the reference's c/qwen_tts.c is compiled
 * with -Dqwen_tts_load=qwen_tts_load_cpu -Dqwen_tts_free=qwen_tts_free_cpu
 * -Dqwen_tts_generate=qwen_tts_generate_cpu, so its own definitions keep working under those names;
 * this file provides the public names.  The reference's unmodified CLI (c/main.c) and every other
 * caller then reach the B200 backend when QWEN_TTS_BACKEND=cuda is in the environment (read once per
 * load, like metal/qwen_tts_talker.c:274-283 does for its own switch) and the CPU path otherwise -
 * also when the device or the upload is unavailable (CPU fallback lives HERE, on the reference's
 * side of the boundary; the library itself has none).
 *
 * Same contract as the CPU path: `text` is the comma-separated token-id list, PCM is malloc'd and
 * the caller frees it, NULL and *out_samples = 0 on error, tunables read from the ctx fields
 * (c/qwen_tts.h:420-430), perf_* written back (:436-440), the -v stderr lines the scripts parse
 * (c/qwen_tts.c:1386-1397, 1433-1438) are printed in the same format.
 *
 * In a maintained fork this is three #ifdef USE_CUDA branches inside c/qwen_tts.c (INTEGRATION.md
 * section 2) and the device handle is a ctx field; here it is a one-entry table because the struct
 * cannot grow without touching the reference's header.  One ctx per process, not thread-safe - the
 * reference's own rule.
 */
#include "qwen_tts.h"          /* the reference's header, -I<reference>/c */
#include "qwen_tts_cuda.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

qwen_tts_ctx_t *qwen_tts_load_cpu(const char *model_dir);
void qwen_tts_free_cpu(qwen_tts_ctx_t *ctx);
float *qwen_tts_generate_cpu(qwen_tts_ctx_t *ctx, const char *text, const char *speaker, const char *language, int *out_samples);
void qwen_tts_cuda_fill_desc(const qwen_tts_ctx_t *ctx, qtc_model_desc_t *d);

static qwen_tts_ctx_t *g_owner = NULL;
static qtc_ctx_t *g_dev = NULL;

static int want_cuda(void) {
    const char *b = getenv("QWEN_TTS_BACKEND");
    return b && strcmp(b, "cuda") == 0;
}

qwen_tts_ctx_t *qwen_tts_load(const char *model_dir) {
    qwen_tts_ctx_t *ctx = qwen_tts_load_cpu(model_dir);
    if (!ctx || !want_cuda()) return ctx;
    if (qwen_tts_cuda_init(0) != 0) {
        fprintf(stderr, "CUDA backend unavailable (%s), using CPU\n", qwen_tts_cuda_last_error());
        return ctx;
    }
    qtc_model_desc_t desc;
    qwen_tts_cuda_fill_desc(ctx, &desc);
    qtc_ctx_t *dev = qwen_tts_cuda_upload(&desc, 1);
    if (!dev) {
        fprintf(stderr, "CUDA upload failed (%s), using CPU\n", qwen_tts_cuda_last_error());
        return ctx;
    }
    if (g_dev) qwen_tts_cuda_release(g_dev);
    g_dev = dev; g_owner = ctx;
    if (qwen_tts_verbose >= 1) { fprintf(stderr, "Backend: cuda\n"); qwen_tts_cuda_print_info(); }
    return ctx;
}

void qwen_tts_free(qwen_tts_ctx_t *ctx) {
    if (ctx && ctx == g_owner) { qwen_tts_cuda_release(g_dev); g_dev = NULL; g_owner = NULL; }
    qwen_tts_free_cpu(ctx);
}

float *qwen_tts_generate(qwen_tts_ctx_t *ctx, const char *text, const char *speaker, const char *language, int *out_samples) {
    if (!ctx || ctx != g_owner || !g_dev) return qwen_tts_generate_cpu(ctx, text, speaker, language, out_samples);
    qtc_gen_params_t p;
    p.temperature = ctx->temperature; p.subtalker_temperature = ctx->subtalker_temperature;
    p.top_k = ctx->top_k; p.subtalker_top_k = ctx->subtalker_top_k;
    p.top_p = ctx->top_p; p.subtalker_top_p = ctx->subtalker_top_p;
    p.repetition_penalty = ctx->repetition_penalty;
    p.max_new_tokens = ctx->max_new_tokens; p.fixed_codec_tokens = ctx->fixed_codec_tokens; p.sample_seed = ctx->sample_seed;
    qwen_tts_cuda_set_params(g_dev, &p);
    int cap = ctx->fixed_codec_tokens > 0 ? ctx->fixed_codec_tokens : ctx->max_new_tokens;
    if (cap < 1) cap = 1;
    /* the library writes codes with stride num_code_groups (c/qwen_tts.c:1261,1339), not codec_num_quantizers */
    const int G = ctx->config.num_code_groups > 0 ? ctx->config.num_code_groups : 16;
    int *codes = (int *)malloc((size_t)cap * G * sizeof(int));
    int n_frames = 0;
    if (out_samples) *out_samples = 0;
    float *pcm = codes ? qwen_tts_cuda_generate(g_dev, text, speaker, language, out_samples, codes, cap, &n_frames) : NULL;
    qtc_perf_t pf;
    qwen_tts_cuda_get_perf(g_dev, &pf);
    ctx->perf_total_ms = pf.total_ms; ctx->perf_talker_ms = pf.talker_ms;
    ctx->perf_codec_ms = pf.codec_ms; ctx->perf_codec_tokens = pf.codec_tokens;
    if (!pcm) {
        /* EOS at step 0 is not an error: the reference returns NULL / 0 samples silently (c/qwen_tts.c:1324-1330, 1388-1392) */
        if (!(n_frames == 0 && pf.stop_reason == 1)) fprintf(stderr, "Error: %s\n", qwen_tts_cuda_last_error());
        free(codes);
        return NULL;
    }
    if (qwen_tts_verbose >= 1) {
        fprintf(stderr, "Generated %d codec tokens in %.1f ms (%.1f ms/token)\n", n_frames, pf.talker_ms,
                n_frames > 0 ? pf.talker_ms / n_frames : 0.0);
        fprintf(stderr, "Stop: %s at step %d\n", pf.stop_reason == 1 ? "eos" : "max_tokens", pf.stop_step);
        if (qwen_tts_verbose >= 2) {
            fprintf(stderr, "Token trace:");
            for (int i = 0; i < n_frames; i++) fprintf(stderr, "%s%d", i == 0 ? " " : ",", codes[(size_t)i * G]);
            fprintf(stderr, "\n");
        }
        fprintf(stderr, "Codec decode: %d samples in %.1f ms\n", *out_samples, pf.codec_ms);
        fprintf(stderr, "Total: %.1f ms (%.2f s audio, %.2fx realtime)\n", pf.total_ms, (float)*out_samples / QWEN_TTS_SAMPLE_RATE,
                ((float)*out_samples / QWEN_TTS_SAMPLE_RATE) / (pf.total_ms / 1000.0));
    }
    free(codes);
    return pcm;
}
