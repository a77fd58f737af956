/*
 * qtc_api.cu - the C ABI of include/qwen_tts_cuda.h: lifecycle, weight upload,
 * parameters, the stage API (c/qwen_tts.h:483-502) and the device-resident
 * frame loop (c/qwen_tts.c:1258-1373).  Host logic only; kernels live in the
 * .cuh files next to this one.
 */
#include "qtc_internal.h"

#include <chrono>
#include <math.h>
#include <stdarg.h>
#include <strings.h>

static char g_err[512] = "";
static int g_device = -1;
int qwen_tts_verbose_cuda = 0;

void qtc_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    fprintf(stderr, "qwen_tts_cuda: %s\n", g_err);
}
extern "C" const char *qwen_tts_cuda_last_error(void) { return g_err; }

extern "C" int qwen_tts_cuda_init(int device) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { qtc_set_error("no CUDA device"); return -1; }
    if (device < 0 || device >= n) { qtc_set_error("device %d out of range (%d present)", device, n); return -1; }
    cudaDeviceProp p;
    QTC_CUDA(cudaGetDeviceProperties(&p, device));
    if (p.major != 10) { qtc_set_error("device %d is sm_%d%d; this library is built for sm_100a only", device, p.major, p.minor); return -1; }
    QTC_CUDA(cudaSetDevice(device));
    g_device = device;
    return 0;
}
extern "C" int qwen_tts_cuda_is_available(void) { return g_device >= 0; }
extern "C" void qwen_tts_cuda_shutdown(void) { if (g_device >= 0) { cudaDeviceSynchronize(); g_device = -1; } }
extern "C" void qwen_tts_cuda_print_info(void) {
    if (g_device < 0) { fprintf(stderr, "qwen_tts_cuda: not initialised\n"); return; }
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, g_device) != cudaSuccess) return;
    size_t fr = 0, tot = 0;
    cudaMemGetInfo(&fr, &tot);
    fprintf(stderr, "qwen_tts_cuda: device %d %s sm_%d%d, %d SMs, %.1f GB HBM (%.1f free), L2 %.0f MB\n", g_device, p.name,
            p.major, p.minor, p.multiProcessorCount, tot / 1e9, fr / 1e9, p.l2CacheSize / 1048576.0);
}
extern "C" void qwen_tts_cuda_default_params(qtc_gen_params_t *p) {   /* c/qwen_tts.c:871-880 */
    p->temperature = 0.9f; p->subtalker_temperature = 0.9f;
    p->top_k = 50; p->subtalker_top_k = 50;
    p->top_p = 1.0f; p->subtalker_top_p = 1.0f;
    p->repetition_penalty = 1.05f;
    p->max_new_tokens = 4096; p->fixed_codec_tokens = 0; p->sample_seed = 42;
}

void *qtc_dmalloc(qtc_ctx *c, size_t bytes) {
    void *p = NULL;
    if (bytes == 0) bytes = 16;
    if (cudaMalloc(&p, (bytes + 255) & ~(size_t)255) != cudaSuccess) { qtc_set_error("cudaMalloc(%zu) failed", bytes); return NULL; }
    c->allocs.push_back(p);
    return p;
}
/* ---- load path ---------------------------------------------------------------------------------
 * The reference hands over pageable host pointers (mmap'ed safetensors, malloc'ed f32 copies: c/qwen_tts.c:481-489, 604-769).
 * A cudaMemcpy from such memory is synchronous and staged by the driver in small pieces; here tensors go through three 32 MB
 * pinned buffers: up to four host threads copy a piece into a pinned buffer (page faults of an mmap'ed file included) while
 * the previous pieces are on the wire (cudaMemcpyAsync on a blocking stream, so later null-stream kernels - the weight
 * re-layouts - order after them).  Upload time and bytes are reported in qtc_perf_t. */
#include <thread>
struct Uploader {
    static const int NBUF = 3;
    static const size_t BUF = 32u << 20;
    char *pin[NBUF] = {NULL, NULL, NULL};
    cudaEvent_t ev[NBUF];
    cudaStream_t stream = NULL;
    int next = 0;
    bool ok = false;
    double bytes = 0;
    bool init() {
        if (ok) return true;
        if (cudaStreamCreate(&stream) != cudaSuccess) return false;
        for (int i = 0; i < NBUF; i++)
            if (cudaMallocHost(&pin[i], BUF) != cudaSuccess || cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) != cudaSuccess) return false;
        ok = true;
        return true;
    }
    static void host_copy(char *dst, const char *src, size_t n) {
        const int nt = n >= (8u << 20) ? 4 : 1;
        if (nt == 1) { memcpy(dst, src, n); return; }
        std::thread th[4];
        const size_t per = ((n / nt) + 4095) & ~(size_t)4095;
        for (int t = 0; t < nt; t++) {
            const size_t lo = per * t, hi = lo + per < n ? lo + per : n;
            th[t] = std::thread([=]() { if (hi > lo) memcpy(dst + lo, src + lo, hi - lo); });
        }
        for (int t = 0; t < nt; t++) th[t].join();
    }
    bool copy(void *dev, const void *host, size_t n) {
        if (!init()) return false;
        const char *src = (const char *)host;
        char *dst = (char *)dev;
        for (size_t off = 0; off < n; off += BUF) {
            const size_t nb = n - off < BUF ? n - off : BUF;
            const int b = next;
            next = (next + 1) % NBUF;
            if (cudaEventSynchronize(ev[b]) != cudaSuccess) return false;      /* the buffer's previous piece is on the device */
            host_copy(pin[b], src + off, nb);
            if (cudaMemcpyAsync(dst + off, pin[b], nb, cudaMemcpyHostToDevice, stream) != cudaSuccess) return false;
            cudaEventRecord(ev[b], stream);
        }
        bytes += (double)n;
        return true;
    }
    bool finish() { return !ok || cudaStreamSynchronize(stream) == cudaSuccess; }
    void release() {
        if (!ok) return;
        cudaStreamSynchronize(stream);
        for (int i = 0; i < NBUF; i++) { cudaFreeHost(pin[i]); cudaEventDestroy(ev[i]); pin[i] = NULL; }
        cudaStreamDestroy(stream);
        ok = false;
    }
};
static Uploader g_up;
template <class T> T *qtc_upload(qtc_ctx *c, const T *host, size_t n) {
    if (!host || n == 0) return NULL;
    T *d = (T *)qtc_dmalloc(c, n * sizeof(T));
    if (!d) return NULL;
    static int plain = -1;
    if (plain < 0) { const char *e = getenv("QTC_UPLOAD_PLAIN"); plain = e ? atoi(e) : 0; }      /* A/B: the synchronous pageable copy */
    if (plain) {
        if (cudaMemcpy(d, host, n * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) { qtc_set_error("H2D copy of %zu bytes failed", n * sizeof(T)); return NULL; }
        g_up.bytes += (double)(n * sizeof(T));
        return d;
    }
    if (!g_up.copy(d, host, n * sizeof(T))) { qtc_set_error("H2D copy of %zu bytes failed: %s", n * sizeof(T), cudaGetErrorString(cudaGetLastError())); return NULL; }
    return d;
}
template float *qtc_upload<float>(qtc_ctx *, const float *, size_t);
template uint16_t *qtc_upload<uint16_t>(qtc_ctx *, const uint16_t *, size_t);

static int upload_lm(qtc_ctx *c, LmDev *d, const qtc_lm_t *h) {
    d->vocab = h->vocab; d->hidden = h->hidden; d->inter = h->intermediate; d->layers = h->layers;
    d->heads = h->heads; d->kv_heads = h->kv_heads; d->head_dim = h->head_dim; d->eps = h->rms_eps; d->theta = h->rope_theta;
    const size_t H = h->hidden, qd = (size_t)h->heads * h->head_dim, kvd = (size_t)h->kv_heads * h->head_dim, I = h->intermediate;
    for (int l = 0; l < h->layers; l++) {
        const qtc_lm_layer_t *s = &h->layer[l];
        LmLayerDev *o = &d->layer[l];
        if (!s->wq || !s->wk || !s->wv || !s->wo || !s->gate || !s->up || !s->down || !s->q_norm || !s->k_norm || !s->input_norm || !s->post_attn_norm) {
            qtc_set_error("decoder layer %d has a missing tensor", l);
            return -1;
        }
        o->wq = qtc_upload(c, s->wq, qd * H); o->wk = qtc_upload(c, s->wk, kvd * H); o->wv = qtc_upload(c, s->wv, kvd * H);
        o->wo = qtc_upload(c, s->wo, H * qd);
        o->gate = qtc_upload(c, s->gate, I * H); o->up = qtc_upload(c, s->up, I * H); o->down = qtc_upload(c, s->down, H * I);
        o->q_norm = qtc_upload(c, s->q_norm, h->head_dim); o->k_norm = qtc_upload(c, s->k_norm, h->head_dim);
        o->in_norm = qtc_upload(c, s->input_norm, H); o->post_norm = qtc_upload(c, s->post_attn_norm, H);
        if (!o->wq || !o->wk || !o->wv || !o->wo || !o->gate || !o->up || !o->down || !o->post_norm) return -1;
    }
    if (h->layers > 0) {
        if (!h->final_norm) { qtc_set_error("decoder stack without final norm"); return -1; }
        d->final_norm = qtc_upload(c, h->final_norm, H);
    }
    return qtc_cluster_prepare(c, d);
}

static int alloc_lm_buffers(qtc_ctx *c, LmBuffers *b, const LmDev *lm) {
    memset(b, 0, sizeof(*b));
    const size_t H = lm->hidden, qd = (size_t)lm->heads * lm->head_dim, kvd = (size_t)lm->kv_heads * lm->head_dim;
    b->x = (float *)qtc_dmalloc(c, H * 4); b->xn = (float *)qtc_dmalloc(c, H * 4);
    b->q = (float *)qtc_dmalloc(c, qd * 4); b->kk = (float *)qtc_dmalloc(c, kvd * 4); b->vv = (float *)qtc_dmalloc(c, kvd * 4);
    b->att = (float *)qtc_dmalloc(c, qd * 4); b->g = (float *)qtc_dmalloc(c, (size_t)lm->inter * 4);
    b->logits = (float *)qtc_dmalloc(c, (size_t)lm->vocab * 4);
    b->xacc = (long long *)qtc_dmalloc(c, 3 * H * 8); b->cl_phase = (int *)qtc_dmalloc(c, 16);
    if (!b->xacc || !b->cl_phase) return -1;
    cudaMemset(b->xacc, 0, 3 * H * 8); cudaMemset(b->cl_phase, 0, 16);
    return (b->x && b->xn && b->q && b->kk && b->vv && b->att && b->g && b->logits) ? 0 : -1;
}

extern "C" void qwen_tts_cuda_release(qtc_ctx_t *c) {
    if (!c) return;
    cudaStreamSynchronize(c->stream);
    if (c->stream2) cudaStreamSynchronize(c->stream2);
    if (c->g_front) cudaGraphExecDestroy(c->g_front);
    if (c->g_talker) cudaGraphExecDestroy(c->g_talker);
    for (int i = 0; i < 4; i++) if (c->g_talker_b[i]) cudaGraphExecDestroy(c->g_talker_b[i]);
    qtc_batch_release(c);
    qtc_prefill_release(c);
    qtc_codec_release(c->codec);
    for (void *p : c->allocs) { qtc_gemm_tc_forget(p); cudaFree(p); }
    LmBuffers *bs[2] = {&c->tk, &c->st};
    for (LmBuffers *b : bs) { cudaFree(b->k); cudaFree(b->v); cudaFree(b->cos); cudaFree(b->sin); }
    cudaFree(c->kv.tables); cudaFreeHost(c->kv.tables_host);
    cudaFree(c->codes); cudaFree(c->trailing); cudaFree(c->prefill);
    cudaFreeHost(c->state_host); cudaFreeHost(c->codes_host); cudaFreeHost(c->host_stage);
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1); cudaEventDestroy(c->ev2); cudaEventDestroy(c->ev3);
    cudaFree(c->pcm_dev); cudaFreeHost(c->first_pcm_host); cudaFree(c->s16_dev);
    for (cudaEvent_t e : c->frame_ev) cudaEventDestroy(e);
    for (cudaEvent_t e : c->chunk_ev) cudaEventDestroy(e);
    cudaStreamDestroy(c->stream); cudaStreamDestroy(c->stream2);
    delete c;
}

extern "C" qtc_ctx_t *qwen_tts_cuda_upload(const qtc_model_desc_t *desc, int max_batch) {
    if (g_device < 0) { qtc_set_error("upload before qwen_tts_cuda_init"); return NULL; }
    if (!desc) { qtc_set_error("null model descriptor"); return NULL; }
    qtc_ctx *c = new qtc_ctx();
    const auto t_up0 = std::chrono::steady_clock::now();
    g_up.bytes = 0;
    c->host = *desc;
    c->H = desc->talker.hidden; c->G = desc->num_code_groups; c->max_batch = max_batch > 0 ? max_batch : 1;
    c->bb = NULL; c->pf = NULL; c->batch_pieces = 3; c->dev_mallocs = 0;
    c->mode = 2; c->arith = 0;
    { const char *e = getenv("QTC_MODE"); if (e) c->mode = atoi(e) < 0 ? 0 : atoi(e) > 2 ? 2 : atoi(e); } c->graph_key = -1; c->g_front = NULL; c->g_talker = NULL; c->launches = 0;
    for (int i = 0; i < 4; i++) c->g_talker_b[i] = NULL;
    c->tk_reach_hint = 0;
    c->codec = NULL; c->codes = NULL; c->trailing = NULL; c->prefill = NULL; c->codes_cap = c->trailing_cap = c->prefill_cap = 0;
    memset(&c->tk, 0, sizeof(c->tk)); memset(&c->st, 0, sizeof(c->st)); memset(&c->perf, 0, sizeof(c->perf));
    memset(&c->talker, 0, sizeof(c->talker)); memset(&c->sub, 0, sizeof(c->sub));
    qwen_tts_cuda_default_params(&c->params);
    /* two streams: the frame loop with the codec chunks inline (qtc_generate.cu:push_chunk says why), and the copies of
     * finished PCM chunks to the host */
    bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreate(&c->ev0) == cudaSuccess && cudaEventCreate(&c->ev1) == cudaSuccess &&
              cudaEventCreate(&c->ev2) == cudaSuccess && cudaEventCreate(&c->ev3) == cudaSuccess;
    c->host_stage_bytes = 1 << 20;
    ok = ok && cudaMallocHost(&c->state_host, sizeof(GenState)) == cudaSuccess &&
         cudaMallocHost(&c->host_stage, c->host_stage_bytes) == cudaSuccess;
    c->codes_host = NULL;
    c->cstream = c->stream;
    { const char *e = getenv("QTC_CODEC_TC"); c->codec_tc = e ? atoi(e) : 1; }
    c->last_pcm_dev = NULL; c->last_pcm_n = 0; c->s16_dev = NULL; c->s16_cap = 0;
    c->pcm_dev = NULL; c->pcm_cap = 0; c->first_pcm_host = NULL; c->stream_first = 8; c->stream_chunk = 0;      /* decode at the end unless qwen_tts_cuda_set_streaming asks for chunks */ c->stream_cb = NULL; c->stream_user = NULL;
    if (!ok) { qtc_set_error("stream/event/pinned allocation failed"); qwen_tts_cuda_release(c); return NULL; }
    if (c->G < 1 || c->G > QTC_MAX_GROUPS) { qtc_set_error("num_code_groups %d out of range", c->G); qwen_tts_cuda_release(c); return NULL; }

    const size_t H = c->H, TH = desc->text_hidden;
    const bool have_lm = desc->talker.layers > 0;
    if (have_lm) {
        if (upload_lm(c, &c->talker, &desc->talker) || upload_lm(c, &c->sub, &desc->sub)) { qwen_tts_cuda_release(c); return NULL; }
        c->codec_embedding = qtc_upload(c, desc->codec_embedding, (size_t)desc->talker.vocab * H);
        c->text_embedding = qtc_upload(c, desc->text_embedding, (size_t)desc->text_vocab * TH);
        c->text_fc1_w = qtc_upload(c, desc->text_fc1_w, TH * TH); c->text_fc1_b = qtc_upload(c, desc->text_fc1_b, TH);
        c->text_fc2_w = qtc_upload(c, desc->text_fc2_w, H * TH); c->text_fc2_b = qtc_upload(c, desc->text_fc2_b, H);
        c->codec_head = qtc_upload(c, desc->codec_head, (size_t)desc->talker.vocab * H);
        if (!c->codec_embedding || !c->text_embedding || !c->text_fc1_w || !c->text_fc2_w || !c->codec_head) {
            qtc_set_error("a talker embedding / head tensor is missing"); qwen_tts_cuda_release(c); return NULL;
        }
        for (int g = 0; g < c->G - 1; g++) {
            c->sub_emb[g] = qtc_upload(c, desc->sub_codec_embeddings[g], (size_t)desc->sub.vocab * H);
            c->sub_head[g] = qtc_upload(c, desc->sub_lm_heads[g], (size_t)desc->sub.vocab * desc->sub.hidden);
            if (!c->sub_emb[g] || !c->sub_head[g]) { qtc_set_error("sub-talker group %d tensor missing", g); qwen_tts_cuda_release(c); return NULL; }
        }
        c->sub_in_w = qtc_upload(c, desc->sub_input_proj_w, (size_t)desc->sub.hidden * H);
        c->sub_in_b = qtc_upload(c, desc->sub_input_proj_b, desc->sub.hidden);
        if (alloc_lm_buffers(c, &c->tk, &c->talker) || alloc_lm_buffers(c, &c->st, &c->sub)) { qwen_tts_cuda_release(c); return NULL; }
        if (qtc_build_proj_tables(c)) { qwen_tts_cuda_release(c); return NULL; }
        if (qtc_ensure_kv(c, &c->st, &c->sub, c->G + 2) || qtc_kv_pool_init(c, c->max_batch)) { qwen_tts_cuda_release(c); return NULL; }
    } else {
        c->talker.vocab = desc->talker.vocab; c->talker.hidden = desc->talker.hidden;
        c->sub.vocab = desc->sub.vocab; c->sub.hidden = desc->sub.hidden;
    }
    const int V = desc->talker.vocab > desc->sub.vocab ? desc->talker.vocab : desc->sub.vocab;
    c->state = (GenState *)qtc_dmalloc(c, sizeof(GenState));
    c->hist_count = (int *)qtc_dmalloc(c, (size_t)(desc->talker.vocab > 0 ? desc->talker.vocab : 1) * 4);
    c->frame_codes = (int *)qtc_dmalloc(c, QTC_MAX_GROUPS * 4);
    c->next_embed = (float *)qtc_dmalloc(c, H * 4); c->tts_pad = (float *)qtc_dmalloc(c, H * 4);
    c->st_embed = (float *)qtc_dmalloc(c, H * 4); c->st_proj = (float *)qtc_dmalloc(c, H * 4);
    c->text_tmp0 = (float *)qtc_dmalloc(c, (TH > H ? TH : H) * 4); c->text_tmp1 = (float *)qtc_dmalloc(c, (TH > H ? TH : H) * 4);
    c->st_rng = (uint32_t *)qtc_dmalloc(c, 16);
    c->zero_pos = (int *)qtc_dmalloc(c, 16);
    if (c->zero_pos) cudaMemset(c->zero_pos, 0, 16);
    {
        cudaDeviceProp prop;
        cudaGetDeviceProperties(&prop, g_device);
        c->num_sms = prop.multiProcessorCount;
        const int heads = desc->talker.heads > desc->sub.heads ? desc->talker.heads : desc->sub.heads;
        const int hd = desc->talker.head_dim > desc->sub.head_dim ? desc->talker.head_dim : desc->sub.head_dim;
        c->talker_dev = (LmDev *)qtc_dmalloc(c, sizeof(LmDev));
        c->sub_dev = (LmDev *)qtc_dmalloc(c, sizeof(LmDev));
        c->bar = (unsigned long long *)qtc_dmalloc(c, 64);
        c->cl_err = (int *)qtc_dmalloc(c, 16);
        if (c->cl_err) cudaMemset(c->cl_err, 0, 16);
        c->head_cnt = (unsigned *)qtc_dmalloc(c, (size_t)(heads > 0 ? heads : 1) * 4);
        c->attn_part = (float *)qtc_dmalloc(c, (size_t)(heads > 0 ? heads : 1) * 16 * (hd + 2) * 4);   /* FUSED_MAX_SPLIT */
        if (!c->talker_dev || !c->sub_dev || !c->bar || !c->head_cnt || !c->attn_part) { qwen_tts_cuda_release(c); return NULL; }
        cudaMemcpy(c->talker_dev, &c->talker, sizeof(LmDev), cudaMemcpyHostToDevice);
        cudaMemcpy(c->sub_dev, &c->sub, sizeof(LmDev), cudaMemcpyHostToDevice);
        cudaMemset(c->bar, 0, 64);
        cudaMemset(c->head_cnt, 0, (size_t)(heads > 0 ? heads : 1) * 4);
    }
    c->sample_scratch = (int *)qtc_dmalloc(c, (size_t)(V > 0 ? V : 1) * 4);
    c->sample_scratch_f = (float *)qtc_dmalloc(c, (size_t)(V > 0 ? V : 1) * 4);
    if (!c->state || !c->hist_count || !c->frame_codes || !c->next_embed || !c->tts_pad || !c->st_embed || !c->text_tmp0 || !c->text_tmp1 ||
        !c->sample_scratch || !c->sample_scratch_f) { qwen_tts_cuda_release(c); return NULL; }
    cudaMemset(c->state, 0, sizeof(GenState));
    cudaMemset(c->frame_codes, 0, QTC_MAX_GROUPS * 4);
    if (desc->codec.layers > 0 && desc->codec.pre_conv_w) {
        c->codec = qtc_codec_upload(c, &desc->codec);
        if (!c->codec) { qwen_tts_cuda_release(c); return NULL; }
    }
    if (!g_up.finish() || cudaDeviceSynchronize() != cudaSuccess) { qtc_set_error("upload: device error"); qwen_tts_cuda_release(c); return NULL; }
    g_up.release();                                   /* 96 MB of pinned memory are not kept after the model is resident */
    c->perf.upload_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_up0).count();
    c->perf.upload_bytes = g_up.bytes;
    c->dev_mallocs = 0;           /* from here on every cudaMalloc is counted (qtc_perf_t.device_mallocs) */
    return c;
}

extern "C" void qwen_tts_cuda_set_params(qtc_ctx_t *c, const qtc_gen_params_t *p) { if (c && p) { c->params = *p; c->graph_key = -1; } }
extern "C" void qwen_tts_cuda_get_perf(const qtc_ctx_t *c, qtc_perf_t *out) {
    if (!c || !out) return;
    *out = c->perf;
    out->kernel_launches = c->launches;
    out->device_mallocs = c->dev_mallocs;
    out->kv_pool_pages = c->kv.pages; out->kv_page_allocs = c->kv.page_allocs;
    int used = 0;
    for (int u : c->kv.used) used += u;
    out->kv_pages_in_use = used;
}
extern "C" int qwen_tts_cuda_set_mode(qtc_ctx_t *c, int mode) {
    if (!c || mode < 0 || mode > 2) return -1;
    c->mode = mode; c->graph_key = -1;
    return 0;
}
/* 1 (default): the vocoder's convolutions run on the tensor cores (two bf16 pieces per value, three MMAs per product);
 * 0: the fp32 FFMA kernels (round 1's path, kept as the A/B baseline and for decoders whose channel counts do not tile) */
extern "C" int qwen_tts_cuda_set_codec_tc(qtc_ctx_t *c, int on) {
    if (!c) return -1;
    c->codec_tc = on ? 1 : 0;
    return 0;
}
extern "C" int qwen_tts_cuda_set_arith(qtc_ctx_t *c, int strict) {
    if (!c) return -1;
    c->arith = strict ? 1 : 0; c->graph_key = -1;
    return 0;
}
