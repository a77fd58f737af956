/*
 * qtc_generate.cu - stage API with host buffers (c/qwen_tts.h:483-502), prompt
 * side (c/qwen_tts.c:823-856, 1147-1243), the device-resident frame loop
 * (c/qwen_tts.c:1258-1373, replayed as two CUDA graphs per frame) and the whole
 * utterance (qwen_tts_generate, c/qwen_tts.c:1059-1443).
 */
#include "qtc_internal.h"

#include <algorithm>
#include <strings.h>
#include <sys/time.h>

static double now_ms() { struct timeval tv; gettimeofday(&tv, NULL); return tv.tv_sec * 1000.0 + tv.tv_usec / 1000.0; }

static int need_lm(qtc_ctx *c) {
    if (!c) { qtc_set_error("null context"); return -1; }
    if (c->talker.layers <= 0) { qtc_set_error("this context was uploaded without talker weights"); return -1; }
    return 0;
}

static int push_state(qtc_ctx *c, const GenState &s) {
    *c->state_host = s;
    QTC_CUDA(cudaMemcpyAsync(c->state, c->state_host, sizeof(GenState), cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
static int pull_state(qtc_ctx *c) {
    QTC_CUDA(cudaMemcpyAsync(c->state_host, c->state, sizeof(GenState), cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" void qwen_tts_cuda_reset(qtc_ctx_t *c) {      /* ctx->talker_kv_len = 0, c/qwen_tts.c:1249 */
    if (!c) return;
    GenState s;
    memset(&s, 0, sizeof(s));
    float f = (float)c->params.sample_seed;
    memcpy(&s.rng, &f, 4);
    s.limit = 1 << 30; s.fixed = c->params.fixed_codec_tokens > 0;
    push_state(c, s);
    if (c->hist_count) cudaMemsetAsync(c->hist_count, 0, (size_t)c->talker.vocab * 4, c->stream);
}

extern "C" int qwen_tts_cuda_get_kv_len(const qtc_ctx_t *c) {
    if (!c) return -1;
    if (pull_state((qtc_ctx *)c)) return -1;
    return c->state_host->pos;
}

/* prefill as consecutive decode steps: per token the reference's batched prefill performs exactly the
 * operations of a decode step, in the same order (oracle/qtts_oracle.c:orc_lm_step, pinned to the reference) */
static int prefill_dev(qtc_ctx *c, const float *embeds_dev, int seq_len) {
    if (qtc_ensure_kv(c, &c->tk, &c->talker, c->state_host->pos + seq_len + 1)) return -1;
    /* fast arithmetic: ONE batched pass, every projection a tensor-core GEMM over the seq_len rows (qtc_prefill.cuh);
     * strict arithmetic (and QTC_PREFILL_GEMM=0): token by token, bit-identical to the reference */
    if (seq_len >= 2 && qtc_prefill_gemm_ok(c)) return qtc_prefill_gemm(c, embeds_dev, seq_len, c->state_host->pos);
    for (int t = 0; t < seq_len; t++) {
        if (qtc_copy(c, c->tk.x, embeds_dev + (size_t)t * c->H, c->H)) return -1;
        if (qtc_lm_step(c, &c->talker, &c->tk, &c->state->pos, 0)) return -1;
        if (qtc_bump_pos(c)) return -1;
    }
    return 0;
}

static int ensure_cap(qtc_ctx *c, float **buf, int *cap, int need_rows, int H) {
    if (need_rows <= *cap) return 0;
    cudaFree(*buf);
    *buf = NULL;
    const int n = need_rows + 64;
    QTC_CUDA(cudaMalloc(buf, (size_t)n * H * 4));
    QTC_COUNT_MALLOC(c);
    *cap = n;
    return 0;
}

extern "C" int qwen_tts_cuda_talker_prefill(qtc_ctx_t *c, const float *input_embeds, int seq_len) {
    if (need_lm(c)) return -1;
    if (!input_embeds || seq_len <= 0) { qtc_set_error("prefill: empty input"); return -1; }
    if (pull_state(c)) return -1;
    if (ensure_cap(c, &c->prefill, &c->prefill_cap, seq_len, c->H)) return -1;
    c->graph_key = -1;
    QTC_CUDA(cudaMemcpyAsync(c->prefill, input_embeds, (size_t)seq_len * c->H * 4, cudaMemcpyHostToDevice, c->stream));
    if (prefill_dev(c, c->prefill, seq_len)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int qwen_tts_cuda_head_logits(qtc_ctx_t *c, float *logits) {
    if (need_lm(c)) return -1;
    GemvSeg h = {c->codec_head, c->tk.logits, c->talker.vocab};
    if (qtc_gemv(c, &h, 1, c->tk.x, c->H, EPI_STORE, NULL)) return -1;
    if (logits) QTC_CUDA(cudaMemcpyAsync(logits, c->tk.logits, (size_t)c->talker.vocab * 4, cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int qwen_tts_cuda_talker_forward(qtc_ctx_t *c, const float *input_embed, float *logits) {
    if (need_lm(c)) return -1;
    if (!input_embed) { qtc_set_error("forward: null embedding"); return -1; }
    if (pull_state(c)) return -1;
    if (qtc_ensure_kv(c, &c->tk, &c->talker, c->state_host->pos + 2)) return -1;
    QTC_CUDA(cudaMemcpyAsync(c->next_embed, input_embed, (size_t)c->H * 4, cudaMemcpyHostToDevice, c->stream));
    c->tk_reach_hint = ((c->state_host->pos + 1 + 127) / 128) * 128;       /* the stage sized as the frame loop sizes it at this position */
    const int frc = qtc_talker_forward_dev(c, c->next_embed);
    c->tk_reach_hint = 0;
    if (frc) return -1;
    if (logits) QTC_CUDA(cudaMemcpyAsync(logits, c->tk.logits, (size_t)c->talker.vocab * 4, cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int qwen_tts_cuda_get_hidden(qtc_ctx_t *c, float *out) {
    if (need_lm(c) || !out) return -1;
    QTC_CUDA(cudaMemcpyAsync(out, c->tk.x, (size_t)c->H * 4, cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int qwen_tts_cuda_subtalker_generate(qtc_ctx_t *c, const float *talker_hidden, int first_code, int *out_codes) {
    if (need_lm(c)) return -1;
    if (!talker_hidden || !out_codes) { qtc_set_error("subtalker: null argument"); return -1; }
    if (first_code < 0 || first_code >= c->talker.vocab) { qtc_set_error("subtalker: first_code %d out of range", first_code); return -1; }
    QTC_CUDA(cudaMemcpyAsync(c->st_proj, talker_hidden, (size_t)c->H * 4, cudaMemcpyHostToDevice, c->stream));
    if (qtc_set_int(c, c->frame_codes, first_code)) return -1;
    if (qtc_subtalker_dev(c, c->st_proj, c->frame_codes, c->frame_codes)) return -1;
    QTC_CUDA(cudaMemcpyAsync(out_codes, c->frame_codes, (size_t)c->G * 4, cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

/* embed_text_token, c/qwen_tts.c:823-847: bf16 row -> fc1 + bias -> SiLU -> fc2 + bias */
static int embed_text_dev(qtc_ctx *c, int token, float *out_dev) {
    const int TH = c->host.text_hidden, H = c->H;
    if (token < 0 || token >= c->host.text_vocab) { qtc_set_error("text token id %d out of range", token); return -1; }
    if (qtc_embed_row(c, c->text_tmp0, c->text_embedding, NULL, token, TH)) return -1;
    GemvSeg f1 = {c->text_fc1_w, c->text_tmp1, TH};
    if (qtc_gemv(c, &f1, 1, c->text_tmp0, TH, c->text_fc1_b ? EPI_BIAS : EPI_STORE, c->text_fc1_b)) return -1;
    if (qtc_silu(c, c->text_tmp1, TH)) return -1;
    GemvSeg f2 = {c->text_fc2_w, out_dev, H};
    return qtc_gemv(c, &f2, 1, c->text_tmp1, TH, c->text_fc2_b ? EPI_BIAS : EPI_STORE, c->text_fc2_b);
}

extern "C" int qwen_tts_cuda_embed_text_tokens(qtc_ctx_t *c, const int *token_ids, int n, float *out) {
    if (need_lm(c)) return -1;
    if (!token_ids || !out || n <= 0) { qtc_set_error("embed_text_tokens: empty input"); return -1; }
    if (ensure_cap(c, &c->prefill, &c->prefill_cap, n, c->H)) return -1;
    for (int i = 0; i < n; i++) if (embed_text_dev(c, token_ids[i], c->prefill + (size_t)i * c->H)) return -1;
    QTC_CUDA(cudaMemcpyAsync(out, c->prefill, (size_t)n * c->H * 4, cudaMemcpyDeviceToHost, c->stream));
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

/* ---- frame loop -------------------------------------------------------------- */
static int params_key(const qtc_ctx *c) {
    uint32_t h = 2166136261u;
    const unsigned char *p = (const unsigned char *)&c->params;
    for (size_t i = 0; i < sizeof(c->params); i++) h = (h ^ p[i]) * 16777619u;
    h = (h ^ (uint32_t)c->mode) * 16777619u;
    h = (h ^ (uint32_t)c->arith) * 16777619u;
    if (c->mode == 0 || c->mode == 2 || c->arith) h = (h ^ (uint32_t)c->kv.used[0]) * 16777619u;   /* per-op attention and the cluster kernel size their shared memory by the capacity */
    return (int)(h & 0x7fffffff);
}

static int capture(qtc_ctx *c, cudaGraphExec_t *out, int (*body)(qtc_ctx *)) {
    cudaGraph_t g = NULL;
    QTC_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const long long before = c->launches;
    const int rc = body(c);
    const cudaError_t e = cudaStreamEndCapture(c->stream, &g);
    c->launches = before;                       /* captured, not executed */
    if (rc != 0 || e != cudaSuccess) { if (g) cudaGraphDestroy(g); if (e != cudaSuccess) qtc_set_error("graph capture failed: %s", cudaGetErrorString(e)); return -1; }
    if (*out) { cudaGraphExecDestroy(*out); *out = NULL; }
    const cudaError_t ei = cudaGraphInstantiate(out, g, 0);
    cudaGraphDestroy(g);
    if (ei != cudaSuccess) { qtc_set_error("graph instantiate failed: %s", cudaGetErrorString(ei)); return -1; }
    return 0;
}
static int body_front(qtc_ctx *c) { return qtc_frame_front(c); }
static int body_talker(qtc_ctx *c) { return qtc_talker_forward_dev(c, c->next_embed); }
static int ensure_graphs(qtc_ctx *c) {
    const int key = params_key(c);
    if (c->graph_key == key && c->g_front && c->g_talker) return 0;
    if (capture(c, &c->g_front, body_front)) return -1;
    if (capture(c, &c->g_talker, body_talker)) return -1;
    for (int i = 0; i < 4; i++) if (c->g_talker_b[i]) { cudaGraphExecDestroy(c->g_talker_b[i]); c->g_talker_b[i] = NULL; }   /* captured under the old key */
    c->graph_key = key;
    return 0;
}
/* mode 2: the talker graph for steps that see at most 128 (bk + 1) positions (stage of 16 (bk + 1) positions per CTA); other modes
 * and longer contexts: the one graph sized by the pages the utterance holds */
static cudaGraphExec_t talker_graph_for(qtc_ctx *c, int positions) {
    const int bk = (positions - 1) / 128;
    if (c->mode != 2 || c->arith || bk < 0 || bk >= 3) return c->g_talker;
    if (!c->g_talker_b[bk]) {
        c->tk_reach_hint = 128 * (bk + 1);
        const int rc = capture(c, &c->g_talker_b[bk], body_talker);
        c->tk_reach_hint = 0;
        if (rc) { c->g_talker_b[bk] = NULL; return c->g_talker; }
    }
    return c->g_talker_b[bk];
}

/* kernels per replay of each graph: run the body under a throwaway capture and read the launch counter */
static int count_launches(qtc_ctx *c, int (*body)(qtc_ctx *), long long *out) {
    cudaGraph_t g = NULL;
    QTC_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const long long before = c->launches;
    const int rc = body(c);
    *out = c->launches - before;
    c->launches = before;
    const cudaError_t e = cudaStreamEndCapture(c->stream, &g);
    if (g) cudaGraphDestroy(g);
    return (rc == 0 && e == cudaSuccess) ? 0 : -1;
}

static int samples_per_frame(const qtc_ctx *c) {
    const qtc_codec_t *k = &c->host.codec;
    int s = 1;
    for (int i = 0; i < 2; i++) s *= k->upsampling_ratios[i];
    for (int i = 0; i < 4; i++) s *= k->upsample_rates[i];
    return s;
}

/* streamed decode of frames [so->decoded, upto): the chunk's kernels are enqueued on the frame loop's OWN stream, between two
 * frames.  The loop's cooperative kernels occupy every SM, so nothing can run beside them; on a second stream a chunk's ~400
 * small dependent kernels got one kernel per gap between two of the loop's kernels, whatever the stream priorities (the next
 * chunk kernel becomes ready a few microseconds after its predecessor ends, and by then the loop's next kernel - always ready -
 * has taken the SMs): time to first audio 220 ms and a 20 % slower loop (profiles/r02_bench_stream_2streams_*.json).  Inline, a
 * chunk costs its own run time once.  Only the PCM copy to the (page-locked) host buffer goes to the second stream. */
static int push_chunk(qtc_ctx *c, StreamOut *so, int upto) {
    const int n = upto - so->decoded, G = c->G;
    if (n <= 0) return 0;
    if (so->nchunks > 0) QTC_CUDA(cudaStreamWaitEvent(c->stream, c->chunk_ev[so->nchunks - 1], 0));   /* the previous chunk's PCM has left the workspace */
    float *wav = NULL;
    const int ns = qtc_codec_stream_push(c, c->codes + (size_t)so->decoded * G, n, &wav);
    if (ns != n * so->spf) return -1;
    if ((size_t)so->nchunks >= c->chunk_ev.size()) { cudaEvent_t e; QTC_CUDA(cudaEventCreate(&e)); c->chunk_ev.push_back(e); }
    QTC_CUDA(cudaEventRecord(c->ev3, c->stream));                                   /* last codec kernel so far */
    QTC_CUDA(cudaStreamWaitEvent(c->stream2, c->ev3, 0));
    float *dst = c->first_pcm_host + (size_t)so->decoded * so->spf;          /* pinned landing buffer of the utterance */
    QTC_CUDA(cudaMemcpyAsync(dst, wav, (size_t)ns * 4, cudaMemcpyDeviceToHost, c->stream2));
    QTC_CUDA(cudaEventRecord(c->chunk_ev[so->nchunks], c->stream2));
    so->chunk_end[so->nchunks] = upto;
    so->nchunks++;
    so->decoded = upto;
    return 0;
}
/* chunks whose PCM has landed are handed to the callback (in order, on the caller's thread) */
static void deliver_chunks(qtc_ctx *c, StreamOut *so, bool wait) {
    while (so->delivered < so->nchunks) {
        cudaEvent_t e = c->chunk_ev[so->delivered];
        if (wait) cudaEventSynchronize(e);
        else if (cudaEventQuery(e) != cudaSuccess) break;
        const int t0 = so->delivered ? so->chunk_end[so->delivered - 1] : 0, t1 = so->chunk_end[so->delivered];
        memcpy(so->host + (size_t)t0 * so->spf, c->first_pcm_host + (size_t)t0 * so->spf, (size_t)(t1 - t0) * so->spf * 4);
        if (c->stream_cb) c->stream_cb(so->host + (size_t)t0 * so->spf, (t1 - t0) * so->spf, t0, c->stream_user);
        so->delivered++;
    }
}

/* The loop proper.  c->prefill [plen,H], c->trailing [ntr,H], c->tts_pad [H] are already in HBM.
 * so != NULL: the codec decodes chunks of finished frames on the second stream while the loop goes on (first chunk
 * c->stream_first frames, then c->stream_chunk), PCM lands in so->host. */
int qtc_generate_frames_resident(qtc_ctx *c, int prefill_len, int n_trailing, int *out_codes, int limit, StreamOut *so) {
    const int H = c->H, G = c->G;
    const int fixed = c->params.fixed_codec_tokens > 0 ? c->params.fixed_codec_tokens : 0;
    if (limit <= 0) return 0;
    const double t0 = now_ms();
    if (c->kv.used[0] * QTC_KV_PAGE_POS < prefill_len + limit + 1 && (int)c->kv.free_pages.size() * QTC_KV_PAGE_POS < prefill_len + limit + 1) qtc_kv_release_all(c);   /* pages of an earlier batch */
    if (qtc_ensure_kv(c, &c->tk, &c->talker, prefill_len + limit + 1)) return -1;
    if (limit > c->codes_cap) {
        cudaFree(c->codes); cudaFreeHost(c->codes_host);
        c->codes = NULL; c->codes_host = NULL;
        QTC_CUDA(cudaMalloc(&c->codes, (size_t)(limit + 64) * G * 4));
        QTC_COUNT_MALLOC(c);
        QTC_CUDA(cudaMallocHost(&c->codes_host, (size_t)(limit + 64) * G * 4));
        c->codes_cap = limit + 64;
        c->graph_key = -1;
    }
    GenState s;
    memset(&s, 0, sizeof(s));
    { float f = (float)c->params.sample_seed; memcpy(&s.rng, &f, 4); }
    s.limit = limit; s.fixed = fixed > 0; s.n_trailing = n_trailing; s.stop_step = limit;
    if (push_state(c, s)) return -1;
    QTC_CUDA(cudaMemsetAsync(c->hist_count, 0, (size_t)c->talker.vocab * 4, c->stream));

    if (prefill_dev(c, c->prefill, prefill_len)) return -1;
    GemvSeg h = {c->codec_head, c->tk.logits, c->talker.vocab};             /* step 0: logits from the prefill hidden */
    if (qtc_gemv(c, &h, 1, c->tk.x, H, EPI_STORE, NULL)) return -1;
    QTC_CUDA(cudaStreamSynchronize(c->stream));
    const double t1 = now_ms();

    if (ensure_graphs(c)) return -1;
    long long lf = 0, lt = 0;
    if (count_launches(c, body_front, &lf) || count_launches(c, body_talker, &lt)) return -1;
    while ((int)c->frame_ev.size() < limit) { cudaEvent_t e; QTC_CUDA(cudaEventCreate(&e)); c->frame_ev.push_back(e); }
    if (so) {
        if (qtc_codec_stream_begin(c, limit)) return -1;
        const int big = (c->stream_first > c->stream_chunk ? c->stream_first : c->stream_chunk) + 8;   /* + one EOS poll interval */
        if (qtc_codec_stream_reserve(c, big < limit ? big : limit)) return -1;
        so->chunk_end.assign((size_t)limit + 1, 0);
    }
    QTC_CUDA(cudaEventRecord(c->ev0, c->stream));
    int launched = 0;
    const int pos0 = prefill_len;                    /* KV length after the prefill (the state was pushed with position 0) */
    const int poll = fixed > 0 ? limit : 8;          /* without a fixed budget, look for EOS every 8 frames */
    while (launched < limit) {
        const int upto = launched + poll < limit ? launched + poll : limit;
        for (; launched < upto; launched++) {
            QTC_CUDA(cudaGraphLaunch(c->g_front, c->stream));
            c->launches += lf;
            QTC_CUDA(cudaEventRecord(c->frame_ev[launched], c->stream));       /* frame `launched` has its 16 codes */
            if (so && fixed > 0) {
                /* fixed budget: every frame is real, chunks are enqueued as the loop is (the host runs ahead of the device) */
                const int want = so->decoded == 0 ? c->stream_first : c->stream_chunk, have = launched + 1 - so->decoded;
                if (have >= want || launched + 1 == limit) { if (push_chunk(c, so, launched + 1)) return -1; }
                deliver_chunks(c, so, false);
            }
            /* the step after frame `launched` puts the token at position prefill_len + launched and sees that many + 1 positions */
            if (launched + 1 < limit) { QTC_CUDA(cudaGraphLaunch(talker_graph_for(c, pos0 + launched + 1), c->stream)); c->launches += lt; }
        }
        if (pull_state(c)) return -1;
        if (so && fixed == 0) {
            /* EOS mode: only frames the device has confirmed (state.step) are decoded */
            const int valid = c->state_host->step, want = so->decoded == 0 ? c->stream_first : c->stream_chunk;
            if (valid - so->decoded >= want || (c->state_host->done && valid > so->decoded) || (launched >= limit && valid > so->decoded)) {
                if (push_chunk(c, so, valid)) return -1;
            }
            deliver_chunks(c, so, false);
        }
        if (c->state_host->done) break;
    }
    QTC_CUDA(cudaEventRecord(c->ev1, c->stream));
    if (pull_state(c)) return -1;
    const int n = c->state_host->step;
    if (so && so->decoded < n) { if (push_chunk(c, so, n)) return -1; }
    if (so) {
        deliver_chunks(c, so, true);
        QTC_CUDA(cudaStreamSynchronize(c->stream2));
        so->samples = so->decoded * so->spf;
    }
    /* latency distribution of this utterance (device clock): first frame from the start of the prompt build, then frame to frame */
    if (n > 0) {
        QTC_CUDA(cudaEventSynchronize(c->frame_ev[n - 1]));
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev2, c->frame_ev[0]);
        c->perf.ttff_ms = ms;
        std::vector<float> dt;
        for (int i = 1; i < n; i++) { float d = 0; cudaEventElapsedTime(&d, c->frame_ev[i - 1], c->frame_ev[i]); dt.push_back(d); }
        std::sort(dt.begin(), dt.end());
        c->perf.frame_ms_p50 = dt.empty() ? ms : dt[dt.size() / 2];
        c->perf.frame_ms_p99 = dt.empty() ? ms : dt[(size_t)((dt.size() - 1) * 0.99)];
        c->perf.ttfa_ms = 0.0;
        if (so && so->nchunks > 0) { float a = 0; cudaEventElapsedTime(&a, c->ev2, c->chunk_ev[0]); c->perf.ttfa_ms = a; }
    }
    if (n > 0 && out_codes) {
        QTC_CUDA(cudaMemcpyAsync(c->codes_host, c->codes, (size_t)n * G * 4, cudaMemcpyDeviceToHost, c->stream));
        QTC_CUDA(cudaStreamSynchronize(c->stream));
        memcpy(out_codes, c->codes_host, (size_t)n * G * 4);
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->perf.prefill_ms = t1 - t0;
    c->perf.talker_ms = ms;
    c->perf.codec_tokens = n;
    c->perf.stop_reason = c->state_host->done == 1 ? 1 : 2;
    c->perf.stop_step = c->state_host->done == 1 ? c->state_host->stop_step : limit;
    return n;
}

extern "C" int qwen_tts_cuda_generate_frames(qtc_ctx_t *c, const float *prefill_embeds, int prefill_len,
                                             const float *trailing_text, int n_trailing, const float *tts_pad,
                                             int *out_codes, int max_tokens) {
    if (need_lm(c)) return -1;
    if (!prefill_embeds || prefill_len <= 0 || !trailing_text || n_trailing <= 0 || !tts_pad || !out_codes) {
        qtc_set_error("generate_frames: null or empty argument"); return -1;
    }
    const int H = c->H;
    const int fixed = c->params.fixed_codec_tokens > 0 ? c->params.fixed_codec_tokens : 0;
    int limit = fixed > 0 ? fixed : c->params.max_new_tokens;
    if (limit > max_tokens) limit = max_tokens;
    const float *old_tr = c->trailing;
    if (ensure_cap(c, &c->prefill, &c->prefill_cap, prefill_len, H) || ensure_cap(c, &c->trailing, &c->trailing_cap, n_trailing, H)) return -1;
    if (old_tr != c->trailing) c->graph_key = -1;          /* the graphs hold the buffer's address */
    QTC_CUDA(cudaMemcpyAsync(c->prefill, prefill_embeds, (size_t)prefill_len * H * 4, cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaMemcpyAsync(c->trailing, trailing_text, (size_t)n_trailing * H * 4, cudaMemcpyHostToDevice, c->stream));
    QTC_CUDA(cudaMemcpyAsync(c->tts_pad, tts_pad, (size_t)H * 4, cudaMemcpyHostToDevice, c->stream));
    cudaEventRecord(c->ev2, c->stream);
    return qtc_generate_frames_resident(c, prefill_len, n_trailing, out_codes, limit, NULL);
}

/* ---- whole utterance ----------------------------------------------------------- */
static int lookup(const char *name, int n, const char *const *names, const int *ids) {
    for (int i = 0; i < n; i++) if (names[i] && strcasecmp(names[i], name) == 0) return ids[i];
    return -1;
}

/* c/qwen_tts.c:1147-1243 on device: prefill rows (3 role tokens, codec prefix + tts_pad/tts_bos, first text token + codec_bos),
 * trailing text rows (+ tts_eos) and the tts_pad row.  pre needs plen + 3 rows (two scratch rows follow the prompt).
 * spk / lang: codec ids or -1. */
int qtc_build_prompt_dev(qtc_ctx *c, const int *tok, int ntok, int spk, int lang, float *pre, int *plen_out, float *trailing, int *ntr_out, float *pad) {
    const qtc_model_desc_t *m = &c->host;
    int prefix[8], np = 0;
    if (lang < 0) { prefix[np++] = m->codec_nothink_id; prefix[np++] = m->codec_think_bos_id; prefix[np++] = m->codec_think_eos_id; }
    else { prefix[np++] = m->codec_think_id; prefix[np++] = m->codec_think_bos_id; prefix[np++] = lang; prefix[np++] = m->codec_think_eos_id; }
    if (spk >= 0) prefix[np++] = spk;
    prefix[np++] = m->codec_pad_id; prefix[np++] = m->codec_bos_id;
    const int H = c->H, plen = 3 + np;
    int ntr = (ntok - 4 - 5) + 1;
    if (ntr < 1) ntr = 1;
    float *bos = pre + (size_t)plen * H, *scratch = pre + (size_t)(plen + 1) * H;
    for (int i = 0; i < 3; i++) if (embed_text_dev(c, tok[i], pre + (size_t)i * H)) return -1;
    if (embed_text_dev(c, m->tts_pad_id, pad) || embed_text_dev(c, m->tts_bos_id, bos)) return -1;
    for (int i = 0; i < np - 1; i++)
        if (qtc_add_row(c, pre + (size_t)(3 + i) * H, i < np - 2 ? pad : bos, c->codec_embedding, prefix[i], H)) return -1;
    if (embed_text_dev(c, tok[3], scratch)) return -1;
    if (qtc_add_row(c, pre + (size_t)(3 + np - 1) * H, scratch, c->codec_embedding, m->codec_bos_id, H)) return -1;
    for (int i = 0; i < ntr - 1; i++) if (embed_text_dev(c, tok[4 + i], trailing + (size_t)i * H)) return -1;
    if (embed_text_dev(c, m->tts_eos_id, trailing + (size_t)(ntr - 1) * H)) return -1;
    *plen_out = plen; *ntr_out = ntr;
    return 0;
}

extern "C" float *qwen_tts_cuda_generate(qtc_ctx_t *c, const char *text, const char *speaker, const char *language,
                                         int *out_samples, int *out_codes, int max_codes, int *out_n_frames) {
    if (!c || !out_samples) return NULL;
    *out_samples = 0;
    if (out_n_frames) *out_n_frames = 0;
    if (need_lm(c)) return NULL;
    const double t_start = now_ms();
    /* comma-separated token ids, c/qwen_tts.c:1079-1112 */
    std::vector<int> tok;
    if (text) {
        const char *p = text;
        while (*p) {
            while (*p == ' ' || *p == ',') p++;
            if (!*p) break;
            char *endp = NULL;
            const long v = strtol(p, &endp, 10);
            if (endp == p) { fprintf(stderr, "Error: invalid token ID near '%s'\n", p); return NULL; }
            tok.push_back((int)v);
            p = endp;
        }
    }
    if ((int)tok.size() < 8) { fprintf(stderr, "Error: need at least 8 text tokens (chat template format)\n"); return NULL; }
    const qtc_model_desc_t *m = &c->host;
    int spk = -1, lang = -1;
    if (speaker && strlen(speaker) > 0) {
        spk = lookup(speaker, m->n_speakers, m->speaker_names, m->speaker_ids);
        if (spk < 0) fprintf(stderr, "Warning: speaker '%s' not found, using no speaker embedding\n", speaker);
    }
    if (language && strlen(language) > 0 && strcasecmp(language, "auto") != 0) {
        lang = lookup(language, m->n_languages, m->language_names, m->language_ids);
        if (lang < 0) fprintf(stderr, "Warning: language '%s' not found\n", language);
    }
    cudaEventRecord(c->ev2, c->stream);
    const int plen_max = 3 + 7;
    int ntr = ((int)tok.size() - 4 - 5) + 1;
    if (ntr < 1) ntr = 1;
    /* the prompt is assembled in HBM; nothing crosses the bus except the token ids */
    const float *old_tr = c->trailing;
    if (ensure_cap(c, &c->prefill, &c->prefill_cap, plen_max + 3, c->H) || ensure_cap(c, &c->trailing, &c->trailing_cap, ntr, c->H)) return NULL;
    if (old_tr != c->trailing) c->graph_key = -1;
    int plen = 0;
    if (qtc_build_prompt_dev(c, tok.data(), (int)tok.size(), spk, lang, c->prefill, &plen, c->trailing, &ntr, c->tts_pad)) return NULL;

    const int fixed = c->params.fixed_codec_tokens > 0 ? c->params.fixed_codec_tokens : 0;
    const int limit = fixed > 0 ? fixed : c->params.max_new_tokens;
    std::vector<int> codes((size_t)(limit > 0 ? limit : 1) * c->G);
    if (!c->codec) { qtc_set_error("generate: context has no codec weights"); return NULL; }
    float *pcm = NULL;
    int n = 0;
    double t_codec = 0.0;
    if (c->stream_chunk > 0 && limit > 0) {
        /* codec chunks decode on the second stream while frames are produced; PCM lands in the caller's buffer, page-locked
         * for the duration of the call so that the copies are asynchronous */
        StreamOut so;
        so.spf = samples_per_frame(c);
        const size_t bytes = (size_t)limit * so.spf * 4;
        so.host = (float *)malloc(bytes);
        if (!so.host) { qtc_set_error("generate: PCM buffer of %zu bytes", bytes); return NULL; }
        if (bytes > c->pcm_cap) {                /* pinned landing buffer, grown on demand (page-locking the caller's buffer per call
                                                    cost up to 60 ms in cudaHostUnregister) */
            cudaFreeHost(c->first_pcm_host); c->first_pcm_host = NULL; c->pcm_cap = 0;
            if (cudaMallocHost(&c->first_pcm_host, bytes + (1 << 20)) != cudaSuccess) { cudaGetLastError(); free(so.host); qtc_set_error("generate: pinned PCM buffer"); return NULL; }
            c->pcm_cap = bytes + (1 << 20);
        }
        n = qtc_generate_frames_resident(c, plen, ntr, codes.data(), limit, &so);
        t_codec = now_ms();
        if (n <= 0) { free(so.host); if (n < 0) return NULL; }
        else { pcm = so.host; *out_samples = so.samples; }
    } else {
        n = qtc_generate_frames_resident(c, plen, ntr, codes.data(), limit, NULL);
        if (n < 0) return NULL;
        t_codec = now_ms();
        if (n > 0) pcm = qtc_codec_decode_dev(c, c->codes, n, out_samples, NULL);
    }
    if (out_codes && n > 0) memcpy(out_codes, codes.data(), (size_t)(n < max_codes ? n : max_codes) * c->G * 4);
    if (out_n_frames) *out_n_frames = n > 0 ? n : 0;
    const double t_end = now_ms();
    c->perf.codec_ms = t_end - t_codec;         /* streamed: only what was left after the last frame */
    if (pcm) { float dms = 0; cudaEventElapsedTime(&dms, c->ev2, c->ev3); c->perf.device_ms = dms; }
    c->perf.total_ms = t_end - t_start;
    return pcm;
}

/* chunk schedule of the streamed decode inside qwen_tts_cuda_generate: first chunk / later chunks in frames;
 * chunk_frames <= 0: decode once after the last frame (the reference's order, c/qwen_tts.c:1399-1420) */
extern "C" int qwen_tts_cuda_set_streaming(qtc_ctx_t *c, int first_frames, int chunk_frames) {
    if (!c) return -1;
    c->stream_chunk = chunk_frames;
    c->stream_first = first_frames > 0 ? first_frames : (chunk_frames > 0 ? chunk_frames : 1);
    return 0;
}
/* cb(pcm, n_samples, first_frame, user) is called on the caller's thread, in order, as chunks land in host memory during
 * qwen_tts_cuda_generate (pcm points into the buffer the call will return).  NULL: off. */
extern "C" void qwen_tts_cuda_set_stream_callback(qtc_ctx_t *c, qtc_pcm_callback_t cb, void *user) {
    if (c) { c->stream_cb = cb; c->stream_user = user; }
}
