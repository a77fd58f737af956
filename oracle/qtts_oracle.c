/*
 * qtts_oracle.c - CPU restatement of the reference hot path (LM half).
 * TEST INFRASTRUCTURE ONLY - see qtts_oracle.h.  Codec half: qtts_oracle_codec.c.
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off -fopenmp (oracle/Makefile).  The
 * OpenMP loops only split disjoint output rows, so results do not depend on
 * the thread count (same property the reference has).
 */
#include "qtts_oracle.h"
#include "qtts_oracle_internal.h"

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* bf16 -> f32 is a 16-bit left shift.  c/qwen_tts.h:475-480 */
static inline float bf16f(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

static void bf16_row(float *dst, const uint16_t *src, int n) { /* c/qwen_tts_kernels.c:360-365 */
    for (int i = 0; i < n; i++) dst[i] = bf16f(src[i]);
}

/* ---- kernels ------------------------------------------------------------- */

/* c/qwen_tts_kernels.c:27-39: ss sequential; inv = 1/sqrtf(ss/dim + eps); (x*inv)*w */
void orc_rms_norm(float *out, const float *x, const float *w, int n, float eps) {
    float acc = 0.0f;
    for (int i = 0; i < n; i++) acc += x[i] * x[i];
    const float inv = 1.0f / sqrtf(acc / (float)n + eps);
    for (int i = 0; i < n; i++) out[i] = x[i] * inv * w[i];
}

/* c/qwen_tts_kernels.c:140-147 (the scalar branch this box builds) */
void orc_matvec_bf16(float *out, const uint16_t *A, const float *x, int rows, int cols) {
#pragma omp parallel for schedule(static) if (rows >= 256)
    for (int r = 0; r < rows; r++) {
        const uint16_t *a = A + (size_t)r * cols;
        float acc = 0.0f;
        for (int c = 0; c < cols; c++) acc += bf16f(a[c]) * x[c];
        out[r] = acc;
    }
}

/* c/qwen_tts_kernels.c:155-160 */
void orc_matvec_f32(float *out, const float *A, const float *x, int rows, int cols) {
#pragma omp parallel for schedule(static) if (rows >= 256)
    for (int r = 0; r < rows; r++) {
        const float *a = A + (size_t)r * cols;
        float acc = 0.0f;
        for (int c = 0; c < cols; c++) acc += a[c] * x[c];
        out[r] = acc;
    }
}

/* c/qwen_tts_kernels.c:174-181: C[m,n] = sum_k A[m,k]*B[n,k], k ascending */
void orc_matmul_f32(float *C, const float *A, const float *B, int M, int N, int K) {
#pragma omp parallel for schedule(static)
    for (int m = 0; m < M; m++)
        for (int n = 0; n < N; n++) {
            const float *a = A + (size_t)m * K, *b = B + (size_t)n * K;
            float acc = 0.0f;
            for (int k = 0; k < K; k++) acc += a[k] * b[k];
            C[(size_t)m * N + n] = acc;
        }
}

/* c/qwen_tts_kernels.c:198-205 */
void orc_matmul_bf16(float *C, const float *A, const uint16_t *B, int M, int N, int K) {
#pragma omp parallel for schedule(static)
    for (int m = 0; m < M; m++)
        for (int n = 0; n < N; n++) {
            const float *a = A + (size_t)m * K;
            const uint16_t *b = B + (size_t)n * K;
            float acc = 0.0f;
            for (int k = 0; k < K; k++) acc += a[k] * bf16f(b[k]);
            C[(size_t)m * N + n] = acc;
        }
}

/* c/qwen_tts_kernels.c:213-233: two matvecs over [2I,H], then g/(1+expf(-g))*u */
static void swiglu_split(float *out, float *up_tmp, const uint16_t *gate, const uint16_t *up,
                         const float *x, int inter, int hidden) {
    orc_matvec_bf16(out, gate, x, inter, hidden);
    orc_matvec_bf16(up_tmp, up, x, inter, hidden);
    for (int i = 0; i < inter; i++) {
        const float g = out[i];
        out[i] = (g / (1.0f + expf(-g))) * up_tmp[i];
    }
}

void orc_swiglu_matvec_bf16(float *out, const uint16_t *gate_up, const float *x, int inter, int hidden) {
    float *tmp = (float *)malloc((size_t)inter * sizeof(float));
    swiglu_split(out, tmp, gate_up, gate_up + (size_t)inter * hidden, x, inter, hidden);
    free(tmp);
}

/* c/qwen_tts_kernels.c:371-378 */
void orc_softmax(float *x, int n) {
    float mx = x[0];
    for (int i = 1; i < n; i++) if (x[i] > mx) mx = x[i];
    float tot = 0.0f;
    for (int i = 0; i < n; i++) { x[i] = expf(x[i] - mx); tot += x[i]; }
    const float inv = 1.0f / tot;
    for (int i = 0; i < n; i++) x[i] *= inv;
}

/* One position of the rotary table: c/qwen_tts_talker.c:76-92 (sub-talker) and
 * :97-113 (talker; the three M-RoPE streams are identical so one row suffices). */
void orc_rope_row(float *cos_out, float *sin_out, int pos, int head_dim, float theta) {
    const int half = head_dim / 2;
    for (int i = 0; i < half; i++) {
        const float freq = 1.0f / powf(theta, (float)(2 * i) / (float)head_dim);
        const float ang = (float)pos * freq;
        const float c = cosf(ang), s = sinf(ang);
        cos_out[i] = c; cos_out[i + half] = c;
        sin_out[i] = s; sin_out[i + half] = s;
    }
}

/* c/qwen_tts_kernels.c:395-405: once per OCCURRENCE in the history */
void orc_repetition_penalty(float *logits, const int *ids, int n_ids, int vocab, float penalty) {
    if (penalty == 1.0f) return;
    for (int i = 0; i < n_ids; i++) {
        const int t = ids[i];
        if (t < 0 || t >= vocab) continue;
        if (logits[t] > 0) logits[t] /= penalty; else logits[t] *= penalty;
    }
}

/* c/qwen_tts_kernels.c:384-393: xorshift32 on the storage of a float */
static float rng_next(float *state) {
    uint32_t s;
    memcpy(&s, state, 4);
    s ^= s << 13; s ^= s >> 17; s ^= s << 5;
    memcpy(state, &s, 4);
    return (float)(s & 0x7FFFFFFFu) / (float)0x7FFFFFFF;
}

/* c/qwen_tts_kernels.c:407-558 */
int orc_sample(const float *logits, int vocab, int top_k, float top_p, float temperature, float *rng_state) {
    if (temperature <= 0.0f) temperature = 1e-5f;

    if (top_p >= 1.0f && top_k > 0 && top_k < vocab) {
        /* fast path :419-484: descending list by insertion; ties keep the lower index first */
        const int k = top_k;
        int *idx = (int *)malloc((size_t)k * sizeof(int));
        float *val = (float *)malloc((size_t)k * sizeof(float));
        for (int j = 0; j < k; j++) { idx[j] = -1; val[j] = -FLT_MAX; }
        for (int i = 0; i < vocab; i++) {
            const float v = logits[i] / temperature;
            if (v <= val[k - 1]) continue;
            int p = k - 1;
            for (; p > 0 && v > val[p - 1]; p--) { val[p] = val[p - 1]; idx[p] = idx[p - 1]; }
            val[p] = v; idx[p] = i;
        }
        const float mx = val[0];
        float tot = 0.0f;
        for (int j = 0; j < k; j++) {
            if (idx[j] < 0) { val[j] = 0.0f; continue; }
            const float p = expf(val[j] - mx);
            val[j] = p; tot += p;
        }
        int pick = 0;
        if (tot > 0.0f) {
            const float r = rng_next(rng_state) * tot;
            float run = 0.0f;
            for (int j = 0; j < k; j++) {
                run += val[j];
                if (run >= r) { pick = idx[j] >= 0 ? idx[j] : 0; break; }
            }
        } else {
            pick = idx[0] >= 0 ? idx[0] : 0;
        }
        free(idx); free(val);
        return pick;
    }

    /* slow path :486-557 */
    float *p = (float *)malloc((size_t)vocab * sizeof(float));
    for (int i = 0; i < vocab; i++) p[i] = logits[i] / temperature;
    orc_softmax(p, vocab);
    if (top_k > 0 && top_k < vocab) {
        /* threshold = k-th largest probability (:494-511) */
        float *tmp = (float *)malloc((size_t)vocab * sizeof(float));
        memcpy(tmp, p, (size_t)vocab * sizeof(float));
        for (int i = 0; i < top_k; i++) {
            int best = i;
            for (int j = i + 1; j < vocab; j++) if (tmp[j] > tmp[best]) best = j;
            float t = tmp[i]; tmp[i] = tmp[best]; tmp[best] = t;
        }
        const float thr = tmp[top_k - 1];
        free(tmp);
        for (int i = 0; i < vocab; i++) if (p[i] < thr) p[i] = 0.0f;
    }
    if (top_p < 1.0f) {
        /* stable descending order, cut after the cumulative mass reaches top_p (:514-537) */
        int *ord = (int *)malloc((size_t)vocab * sizeof(int));
        for (int i = 0; i < vocab; i++) ord[i] = i;
        for (int i = 1; i < vocab; i++) {
            const int id = ord[i]; const float v = p[id];
            int j = i - 1;
            while (j >= 0 && p[ord[j]] < v) { ord[j + 1] = ord[j]; j--; }
            ord[j + 1] = id;
        }
        float run = 0.0f; int cut = vocab;
        for (int i = 0; i < vocab; i++) { run += p[ord[i]]; if (run >= top_p) { cut = i + 1; break; } }
        for (int i = cut; i < vocab; i++) p[ord[i]] = 0.0f;
        free(ord);
    }
    float tot = 0.0f;
    for (int i = 0; i < vocab; i++) tot += p[i];
    if (tot > 0.0f) { const float inv = 1.0f / tot; for (int i = 0; i < vocab; i++) p[i] *= inv; }
    const float r = rng_next(rng_state);
    float run = 0.0f; int pick = 0;
    for (int i = 0; i < vocab; i++) { run += p[i]; if (run >= r) { pick = i; break; } }
    free(p);
    return pick;
}

/* ---- decoder stack -------------------------------------------------------- */

void orc_lm_alloc(orc_lm_state_t *st, const qtc_lm_t *lm, int kv_max) {
    memset(st, 0, sizeof(*st));
    st->lm = lm;
    st->kv_max = kv_max;
    const size_t kvd = (size_t)lm->kv_heads * lm->head_dim;
    st->k = (float *)calloc((size_t)lm->layers * kv_max * kvd, sizeof(float));
    st->v = (float *)calloc((size_t)lm->layers * kv_max * kvd, sizeof(float));
    const int qd = lm->heads * lm->head_dim;
    st->x = (float *)calloc(lm->hidden, sizeof(float));
    st->xn = (float *)calloc(lm->hidden, sizeof(float));
    st->q = (float *)calloc(qd, sizeof(float));
    st->kk = (float *)calloc(kvd, sizeof(float));
    st->vv = (float *)calloc(kvd, sizeof(float));
    st->att = (float *)calloc(qd, sizeof(float));
    st->g = (float *)calloc(lm->intermediate, sizeof(float));
    st->u = (float *)calloc(lm->intermediate, sizeof(float));
    st->sc = (float *)calloc(kv_max, sizeof(float));
}

void orc_lm_free(orc_lm_state_t *st) {
    free(st->k); free(st->v); free(st->x); free(st->xn); free(st->q); free(st->kk);
    free(st->vv); free(st->att); free(st->g); free(st->u); free(st->sc);
    memset(st, 0, sizeof(*st));
}

static void lm_grow(orc_lm_state_t *st, int need) {
    if (need <= st->kv_max) return;
    const qtc_lm_t *lm = st->lm;
    const size_t kvd = (size_t)lm->kv_heads * lm->head_dim;
    const int nmax = need + 2048;
    float *nk = (float *)calloc((size_t)lm->layers * nmax * kvd, sizeof(float));
    float *nv = (float *)calloc((size_t)lm->layers * nmax * kvd, sizeof(float));
    for (int l = 0; l < lm->layers; l++) {
        memcpy(nk + (size_t)l * nmax * kvd, st->k + (size_t)l * st->kv_max * kvd, (size_t)st->kv_max * kvd * 4);
        memcpy(nv + (size_t)l * nmax * kvd, st->v + (size_t)l * st->kv_max * kvd, (size_t)st->kv_max * kvd * 4);
    }
    free(st->k); free(st->v); st->k = nk; st->v = nv;
    st->sc = (float *)realloc(st->sc, (size_t)nmax * sizeof(float));
    st->kv_max = nmax;
}

/* One token through every layer + final norm; st->x holds the input on entry and
 * the post-final-norm hidden on exit.
 * Talker: talker_attention_single c/qwen_tts_talker.c:119-248 + :520-527.
 * Sub-talker: ST_FORWARD c/qwen_tts_talker.c:638-691.
 * The prefill (c/qwen_tts_talker.c:320-458) performs, per token, exactly these
 * operations in exactly this order (causal attention, k-ascending sums), so
 * running this step token by token reproduces it bit for bit
 * (checked against oracle/_ref in tests/test_oracle_vs_ref.py). */
void orc_lm_step(orc_lm_state_t *st, int pos) {
    const qtc_lm_t *lm = st->lm;
    const int H = lm->hidden, hd = lm->head_dim, nh = lm->heads, nkv = lm->kv_heads;
    const int qd = nh * hd, kvd = nkv * hd, half = hd / 2, rep = nh / nkv;
    const float eps = lm->rms_eps;
    const float scale = 1.0f / sqrtf((float)hd);
    float cs[512], sn[512];
    lm_grow(st, pos + 1);
    orc_rope_row(cs, sn, pos, hd, lm->rope_theta);

    for (int l = 0; l < lm->layers; l++) {
        const qtc_lm_layer_t *w = &lm->layer[l];
        float *Kc = st->k + (size_t)l * st->kv_max * kvd;
        float *Vc = st->v + (size_t)l * st->kv_max * kvd;

        orc_rms_norm(st->xn, st->x, w->input_norm, H, eps);
        orc_matvec_bf16(st->q, w->wq, st->xn, qd, H);
        orc_matvec_bf16(st->kk, w->wk, st->xn, kvd, H);
        orc_matvec_bf16(st->vv, w->wv, st->xn, kvd, H);
        for (int h = 0; h < nh; h++) orc_rms_norm(st->q + h * hd, st->q + h * hd, w->q_norm, hd, eps);
        for (int h = 0; h < nkv; h++) orc_rms_norm(st->kk + h * hd, st->kk + h * hd, w->k_norm, hd, eps);
        /* rotate-half, c/qwen_tts_talker.c:172-188 / c/qwen_tts_kernels.c:570-577 */
        for (int h = 0; h < nh + nkv; h++) {
            float *p = h < nh ? st->q + h * hd : st->kk + (h - nh) * hd;
            for (int i = 0; i < half; i++) {
                const float a = p[i], b = p[i + half];
                p[i] = a * cs[i] - b * sn[i];
                p[i + half] = b * cs[i + half] + a * sn[i + half];
            }
        }
        memcpy(Kc + (size_t)pos * kvd, st->kk, (size_t)kvd * 4);
        memcpy(Vc + (size_t)pos * kvd, st->vv, (size_t)kvd * 4);

        /* c/qwen_tts_talker.c:199-225 */
        for (int h = 0; h < nh; h++) {
            const int g = h / rep;
            const float *qh = st->q + h * hd;
            for (int t = 0; t <= pos; t++) {
                const float *kt = Kc + (size_t)t * kvd + g * hd;
                float d = 0.0f;
                for (int i = 0; i < hd; i++) d += qh[i] * kt[i];
                st->sc[t] = d * scale;
            }
            orc_softmax(st->sc, pos + 1);
            float *o = st->att + h * hd;
            for (int i = 0; i < hd; i++) o[i] = 0.0f;
            for (int t = 0; t <= pos; t++) {
                const float *vt = Vc + (size_t)t * kvd + g * hd;
                const float p = st->sc[t];
                for (int i = 0; i < hd; i++) o[i] += p * vt[i];
            }
        }
        orc_matvec_bf16(st->xn, w->wo, st->att, H, qd);
        for (int i = 0; i < H; i++) st->x[i] += st->xn[i];

        orc_rms_norm(st->xn, st->x, w->post_attn_norm, H, eps);
        swiglu_split(st->g, st->u, w->gate, w->up, st->xn, lm->intermediate, H);
        orc_matvec_bf16(st->xn, w->down, st->g, H, lm->intermediate);
        for (int i = 0; i < H; i++) st->x[i] += st->xn[i];
    }
    orc_rms_norm(st->x, st->x, lm->final_norm, H, eps);
}

/* ---- state ---------------------------------------------------------------- */

orc_state_t *orc_create(const qtc_model_desc_t *m) {
    orc_state_t *s = (orc_state_t *)calloc(1, sizeof(*s));
    s->m = m;
    orc_lm_alloc(&s->tk, &m->talker, 512);
    orc_lm_alloc(&s->st, &m->sub, m->num_code_groups + 2);
    qwen_tts_cuda_default_params_host(&s->p);
    s->st_logits = (float *)calloc(m->sub.vocab, sizeof(float));
    s->st_embed = (float *)calloc(m->talker.hidden, sizeof(float));
    s->st_proj = (float *)calloc(m->sub.hidden, sizeof(float));
    return s;
}

void orc_destroy(orc_state_t *s) {
    if (!s) return;
    orc_lm_free(&s->tk); orc_lm_free(&s->st);
    free(s->st_logits); free(s->st_embed); free(s->st_proj);
    free(s);
}

/* defaults c/qwen_tts.c:871-880 */
void qwen_tts_cuda_default_params_host(qtc_gen_params_t *p) {
    p->temperature = 0.9f; p->subtalker_temperature = 0.9f;
    p->top_k = 50; p->subtalker_top_k = 50;
    p->top_p = 1.0f; p->subtalker_top_p = 1.0f;
    p->repetition_penalty = 1.05f;
    p->max_new_tokens = 4096; p->fixed_codec_tokens = 0; p->sample_seed = 42;
}

void orc_set_params(orc_state_t *s, const qtc_gen_params_t *p) { s->p = *p; }
void orc_reset(orc_state_t *s) { s->tk_len = 0; }
const float *orc_hidden(const orc_state_t *s) { return s->tk.x; }
int orc_kv_len(const orc_state_t *s) { return s->tk_len; }

/* c/qwen_tts_talker.c:254-472 (see orc_lm_step for why token-by-token is exact) */
void orc_talker_prefill(orc_state_t *s, const float *embeds, int seq_len) {
    const int H = s->m->talker.hidden;
    for (int t = 0; t < seq_len; t++) {
        memcpy(s->tk.x, embeds + (size_t)t * H, (size_t)H * 4);
        orc_lm_step(&s->tk, t);
    }
    s->tk_len = seq_len;
}

/* c/qwen_tts_talker.c:478-533 */
void orc_talker_forward(orc_state_t *s, const float *embed, float *logits) {
    const qtc_lm_t *lm = &s->m->talker;
    memcpy(s->tk.x, embed, (size_t)lm->hidden * 4);
    orc_lm_step(&s->tk, s->tk_len);
    orc_matvec_bf16(logits, s->m->codec_head, s->tk.x, lm->vocab, lm->hidden);
    s->tk_len += 1;
}

void orc_head_logits(orc_state_t *s, float *logits) {
    orc_matvec_bf16(logits, s->m->codec_head, s->tk.x, s->m->talker.vocab, s->m->talker.hidden);
}

/* ST_PROJECT_INPUT c/qwen_tts_talker.c:693-702 */
static void sub_project(orc_state_t *s, float *dst, const float *src) {
    const qtc_model_desc_t *m = s->m;
    const int h = m->sub.hidden, H = m->talker.hidden;
    if (m->sub_input_proj_w) {
        orc_matvec_bf16(dst, m->sub_input_proj_w, src, h, H);
        if (m->sub_input_proj_b) for (int i = 0; i < h; i++) dst[i] += m->sub_input_proj_b[i];
    } else {
        const int n = H < h ? H : h;
        memcpy(dst, src, (size_t)n * 4);
        for (int i = n; i < h; i++) dst[i] = 0.0f;
    }
}

/* c/qwen_tts_talker.c:539-736 */
void orc_subtalker_generate(orc_state_t *s, const float *talker_hidden, int first_code, int *out_codes) {
    const qtc_model_desc_t *m = s->m;
    const int h = m->sub.hidden, H = m->talker.hidden, G = m->num_code_groups, v = m->sub.vocab;
    out_codes[0] = first_code;

    sub_project(s, s->st_proj, talker_hidden);                       /* pos 0: no logits */
    memcpy(s->st.x, s->st_proj, (size_t)h * 4);
    orc_lm_step(&s->st, 0);

    bf16_row(s->st_embed, m->codec_embedding + (size_t)first_code * H, H);   /* pos 1 */
    sub_project(s, s->st_proj, s->st_embed);
    memcpy(s->st.x, s->st_proj, (size_t)h * 4);
    orc_lm_step(&s->st, 1);
    orc_matvec_bf16(s->st_logits, m->sub_lm_heads[0], s->st.x, v, h);
    float rng = (float)s->p.sample_seed;                             /* re-seeded every frame, :718 */
    out_codes[1] = orc_sample(s->st_logits, v, s->p.subtalker_top_k, s->p.subtalker_top_p,
                              s->p.subtalker_temperature, &rng);

    for (int g = 2; g < G; g++) {
        bf16_row(s->st_embed, m->sub_codec_embeddings[g - 2] + (size_t)out_codes[g - 1] * H, H);
        sub_project(s, s->st_proj, s->st_embed);
        memcpy(s->st.x, s->st_proj, (size_t)h * 4);
        orc_lm_step(&s->st, g);
        orc_matvec_bf16(s->st_logits, m->sub_lm_heads[g - 1], s->st.x, v, h);
        out_codes[g] = orc_sample(s->st_logits, v, s->p.subtalker_top_k, s->p.subtalker_top_p,
                                  s->p.subtalker_temperature, &rng);
    }
}

/* c/qwen_tts.c:823-847 */
void orc_embed_text_token(orc_state_t *s, int token_id, float *out) {
    const qtc_model_desc_t *m = s->m;
    const int TH = m->text_hidden, H = m->talker.hidden;
    float *e = (float *)malloc((size_t)TH * 4), *f = (float *)malloc((size_t)TH * 4);
    bf16_row(e, m->text_embedding + (size_t)token_id * TH, TH);
    orc_matvec_bf16(f, m->text_fc1_w, e, TH, TH);
    if (m->text_fc1_b) for (int i = 0; i < TH; i++) f[i] += m->text_fc1_b[i];
    for (int i = 0; i < TH; i++) f[i] = f[i] / (1.0f + expf(-f[i]));
    orc_matvec_bf16(out, m->text_fc2_w, f, H, TH);
    if (m->text_fc2_b) for (int i = 0; i < H; i++) out[i] += m->text_fc2_b[i];
    free(e); free(f);
}

/* c/qwen_tts.c:1258-1373 */
int orc_generate_frames(orc_state_t *s, const float *prefill_embeds, int prefill_len,
                        const float *trailing_text, int n_trailing, const float *tts_pad,
                        int *out_codes, int max_tokens, int *stop_reason, int *stop_step) {
    const qtc_model_desc_t *m = s->m;
    const int H = m->talker.hidden, V = m->talker.vocab, G = m->num_code_groups;
    const int fixed = s->p.fixed_codec_tokens > 0 ? s->p.fixed_codec_tokens : 0;
    int limit = fixed > 0 ? fixed : s->p.max_new_tokens;
    if (limit > max_tokens) limit = max_tokens;

    orc_reset(s);
    orc_talker_prefill(s, prefill_embeds, prefill_len);

    float *logits = (float *)malloc((size_t)V * 4);
    float *next = (float *)calloc(H, 4), *row = (float *)malloc((size_t)H * 4);
    int *hist = (int *)calloc(limit > 0 ? limit : 1, sizeof(int));
    float rng = (float)s->p.sample_seed;
    int n = 0, why = 0, at = limit;

    for (int step = 0; step < limit; step++) {
        if (step == 0) orc_head_logits(s, logits);
        else orc_talker_forward(s, next, logits);

        for (int i = V - 1024; i < V; i++) if (i != m->codec_eos_id) logits[i] = -1e9f;   /* :1272-1305 */
        orc_repetition_penalty(logits, hist, n, V, s->p.repetition_penalty);
        int tok = orc_sample(logits, V, s->p.top_k, s->p.top_p, s->p.temperature, &rng);
        if (fixed > 0 && tok == m->codec_eos_id && n < fixed) {                           /* :1315-1321 */
            const float keep = logits[m->codec_eos_id];
            logits[m->codec_eos_id] = -1e9f;
            tok = orc_sample(logits, V, s->p.top_k, s->p.top_p, s->p.temperature, &rng);
            logits[m->codec_eos_id] = keep;
        }
        if (fixed == 0 && tok == m->codec_eos_id) { why = 1; at = step; break; }
        hist[n] = tok;

        int codes[QTC_MAX_GROUPS];
        orc_subtalker_generate(s, s->tk.x, tok, codes);
        memcpy(out_codes + (size_t)n * G, codes, (size_t)G * sizeof(int));
        n++;

        /* :1345-1363: zero, + talker row, + 15 sub rows (g ascending), + text row */
        for (int i = 0; i < H; i++) next[i] = 0.0f;
        bf16_row(row, m->codec_embedding + (size_t)tok * H, H);
        for (int i = 0; i < H; i++) next[i] += row[i];
        for (int g = 1; g < G; g++) {
            bf16_row(row, m->sub_codec_embeddings[g - 1] + (size_t)codes[g] * H, H);
            for (int i = 0; i < H; i++) next[i] += row[i];
        }
        const float *txt = step < n_trailing ? trailing_text + (size_t)step * H : tts_pad;
        for (int i = 0; i < H; i++) next[i] += txt[i];
    }
    if (why == 0) { why = 2; at = limit; }
    if (stop_reason) *stop_reason = why;
    if (stop_step) *stop_step = at;
    free(logits); free(next); free(row); free(hist);
    return n;
}

/* Prompt layout of qwen_tts_generate, c/qwen_tts.c:1147-1243.  Allocates and
 * returns the three embedding blocks the frame loop needs. */
int orc_build_prompt(orc_state_t *s, const int *tok, int n_tok, int speaker_id, int language_id,
                     float **prefill_out, int *prefill_len_out, float **trailing_out,
                     int *n_trailing_out, float **pad_out) {
    const qtc_model_desc_t *m = s->m;
    const int H = m->talker.hidden;
    if (n_tok < 8) return -1;
    int prefix[8], np = 0;
    if (language_id < 0) {
        prefix[np++] = m->codec_nothink_id; prefix[np++] = m->codec_think_bos_id; prefix[np++] = m->codec_think_eos_id;
    } else {
        prefix[np++] = m->codec_think_id; prefix[np++] = m->codec_think_bos_id;
        prefix[np++] = language_id; prefix[np++] = m->codec_think_eos_id;
    }
    if (speaker_id >= 0) prefix[np++] = speaker_id;
    prefix[np++] = m->codec_pad_id; prefix[np++] = m->codec_bos_id;

    const int plen = 3 + np;
    float *pre = (float *)calloc((size_t)plen * H, 4);
    float *pad = (float *)malloc((size_t)H * 4), *bos = (float *)malloc((size_t)H * 4);
    float *eos = (float *)malloc((size_t)H * 4), *row = (float *)malloc((size_t)H * 4);
    for (int i = 0; i < 3; i++) orc_embed_text_token(s, tok[i], pre + (size_t)i * H);
    orc_embed_text_token(s, m->tts_pad_id, pad);
    orc_embed_text_token(s, m->tts_bos_id, bos);
    orc_embed_text_token(s, m->tts_eos_id, eos);
    for (int i = 0; i < np - 1; i++) {
        float *d = pre + (size_t)(3 + i) * H;
        memcpy(d, i < np - 2 ? pad : bos, (size_t)H * 4);
        bf16_row(row, m->codec_embedding + (size_t)prefix[i] * H, H);
        for (int j = 0; j < H; j++) d[j] += row[j];
    }
    {
        float *d = pre + (size_t)(3 + np - 1) * H;
        orc_embed_text_token(s, tok[3], d);
        bf16_row(row, m->codec_embedding + (size_t)m->codec_bos_id * H, H);
        for (int j = 0; j < H; j++) d[j] += row[j];
    }
    int ntr = (n_tok - 4 - 5) + 1;
    if (ntr < 1) ntr = 1;
    float *tr = (float *)calloc((size_t)ntr * H, 4);
    for (int i = 0; i < ntr - 1; i++) orc_embed_text_token(s, tok[4 + i], tr + (size_t)i * H);
    memcpy(tr + (size_t)(ntr - 1) * H, eos, (size_t)H * 4);
    free(bos); free(eos); free(row);
    *prefill_out = pre; *prefill_len_out = plen; *trailing_out = tr; *n_trailing_out = ntr; *pad_out = pad;
    return 0;
}

/* c/qwen_tts.c:1059-1443 */
float *orc_generate(orc_state_t *s, const int *text_tokens, int n_text_tokens,
                    int speaker_codec_id, int language_codec_id,
                    int *out_samples, int *out_codes, int max_codes, int *out_n_frames) {
    float *pre, *tr, *pad; int plen, ntr;
    *out_samples = 0;
    if (out_n_frames) *out_n_frames = 0;
    if (orc_build_prompt(s, text_tokens, n_text_tokens, speaker_codec_id, language_codec_id,
                         &pre, &plen, &tr, &ntr, &pad) != 0) {
        fprintf(stderr, "Error: need at least 8 text tokens (chat template format)\n");
        return NULL;
    }
    const int G = s->m->num_code_groups;
    const int fixed = s->p.fixed_codec_tokens > 0 ? s->p.fixed_codec_tokens : 0;
    const int limit = fixed > 0 ? fixed : s->p.max_new_tokens;
    int *codes = (int *)calloc((size_t)limit * G, sizeof(int));
    const int n = orc_generate_frames(s, pre, plen, tr, ntr, pad, codes, limit, NULL, NULL);
    free(pre); free(tr); free(pad);
    float *pcm = NULL;
    if (n > 0) pcm = orc_codec_decode(s, codes, n, out_samples);
    if (out_codes) memcpy(out_codes, codes, (size_t)(n < max_codes ? n : max_codes) * G * sizeof(int));
    if (out_n_frames) *out_n_frames = n;
    free(codes);
    return pcm;
}
void orc_free(void *p) { free(p); }
