/*
 * qtts_oracle_codec.c - CPU restatement of the codec decoder / vocoder.
 * TEST INFRASTRUCTURE ONLY - see qtts_oracle.h.
 *
 * Follows c/qwen_tts_codec.c (pipeline) and the non-BLAS branches of
 * c/qwen_tts_kernels.c (conv, transposed conv, SnakeBeta, LayerNorm, GELU).
 * Every accumulation keeps the reference's order: bias first, then input
 * channel ascending, then tap ascending.
 */
#include "qtts_oracle.h"
#include "qtts_oracle_internal.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* c/qwen_tts_kernels.c:45-58 */
void orc_layer_norm(float *out, const float *x, const float *w, const float *b, int n, float eps) {
    float mean = 0.0f;
    for (int i = 0; i < n; i++) mean += x[i];
    mean /= (float)n;
    float var = 0.0f;
    for (int i = 0; i < n; i++) { const float d = x[i] - mean; var += d * d; }
    var /= (float)n;
    const float inv = 1.0f / sqrtf(var + eps);
    for (int i = 0; i < n; i++) {
        float y = (x[i] - mean) * inv;
        if (w) y *= w[i];
        if (b) y += b[i];
        out[i] = y;
    }
}

/* c/qwen_tts_kernels.c:244-249 (tanh form) */
void orc_gelu(float *x, int n) {
    for (int i = 0; i < n; i++) {
        const float v = x[i];
        x[i] = 0.5f * v * (1.0f + tanhf(0.7978845608028654f * (v + 0.044715f * v * v * v)));
    }
}

/* c/qwen_tts_kernels.c:302-310; alpha / inv_beta are the load-time processed values */
void orc_snake_beta(float *out, const float *x, const float *alpha, const float *inv_beta, int C, int L) {
#pragma omp parallel for schedule(static)
    for (int c = 0; c < C; c++) {
        const float a = alpha[c], ib = inv_beta[c];
        const float *xi = x + (size_t)c * L;
        float *yo = out + (size_t)c * L;
        for (int t = 0; t < L; t++) {
            const float s = sinf(xi[t] * a);
            yo[t] = xi[t] + ib * s * s;
        }
    }
}

/* c/qwen_tts_kernels.c:659-871, non-BLAS branches.  All of them evaluate, for
 * every output sample, bias then (input channel ascending, tap ascending)
 * products; taps that would read before t=0 are skipped (left zero padding).
 * The k=1 depthwise branch (:733-748) computes in*w+b, the same sum. */
void orc_causal_conv1d(float *out, const float *in, const float *w, const float *b,
                       int cin, int cout, int k, int L, int dilation, int groups) {
    const int cpg = cin / groups, opg = cout / groups;
    const int pad = (k - 1) * dilation;
#pragma omp parallel for schedule(static)
    for (int oc = 0; oc < cout; oc++) {
        float *y = out + (size_t)oc * L;
        const float bias = b ? b[oc] : 0.0f;
        for (int t = 0; t < L; t++) y[t] = bias;
        const int ic0 = (oc / opg) * cpg;
        for (int j = 0; j < cpg; j++) {
            const float *x = in + (size_t)(ic0 + j) * L;
            const float *wk = w + ((size_t)oc * cpg + j) * k;
            for (int tap = 0; tap < k; tap++) {
                const int shift = pad - tap * dilation;   /* y[t] += wk * x[t - shift] */
                if (shift >= L) continue;
                const float c = wk[tap];
                for (int t = shift; t < L; t++) y[t] += c * x[t - shift];
            }
        }
    }
}

/* c/qwen_tts_kernels.c:948-971 (scalar branch): weight [cin, cout, k];
 * raw length (L-1)*stride+k, trimmed by k-stride on the right. */
void orc_transposed_conv1d(float *out, const float *in, const float *w, const float *b,
                           int cin, int cout, int k, int stride, int L, int *out_len) {
    int n = (L - 1) * stride + k - (k - stride);
    if (n < 0) n = 0;
#pragma omp parallel for schedule(static)
    for (int oc = 0; oc < cout; oc++) {
        float *y = out + (size_t)oc * n;
        const float bias = b ? b[oc] : 0.0f;
        for (int t = 0; t < n; t++) y[t] = bias;
        for (int ic = 0; ic < cin; ic++) {
            const float *x = in + (size_t)ic * L;
            const float *wk = w + ((size_t)ic * cout + oc) * k;
            for (int t = 0; t < L; t++) {
                const float v = x[t];
                float *dst = y + (size_t)t * stride;
                const int lim = n - t * stride < k ? n - t * stride : k;
                for (int tap = 0; tap < lim; tap++) dst[tap] += v * wk[tap];
            }
        }
    }
    if (out_len) *out_len = n;
}

/* c/qwen_tts_codec.c:127-261 -> [latent/2, T] channels-first */
static float *rvq_dequantize(const qtc_codec_t *c, const int *codes, int T) {
    const int nq = c->num_quantizers, vq = c->codebook_dim / 2, half = c->latent / 2;
    float *sem = (float *)calloc((size_t)vq * T, 4), *aco = (float *)calloc((size_t)vq * T, 4);
    for (int q = 0; q < nq; q++) {
        float *dst = q == 0 ? sem : aco;
        const float *E = c->codebooks[q];
        for (int t = 0; t < T; t++) {
            int code = codes[(size_t)t * nq + q];
            if (code < 0 || code >= c->codebook_size) code = 0;
            for (int d = 0; d < vq; d++) dst[(size_t)d * T + t] += E[(size_t)code * vq + d];
        }
    }
    float *out = (float *)malloc((size_t)half * T * 4);
#pragma omp parallel for schedule(static)
    for (int o = 0; o < half; o++)
        for (int t = 0; t < T; t++) {
            float a = 0, b = 0;
            for (int d = 0; d < vq; d++) a += c->semantic_out_proj[(size_t)o * vq + d] * sem[(size_t)d * T + t];
            for (int d = 0; d < vq; d++) b += c->acoustic_out_proj[(size_t)o * vq + d] * aco[(size_t)d * T + t];
            out[(size_t)o * T + t] = a + b;
        }
    free(sem); free(aco);
    return out;
}

static void transpose(float *dst, const float *src, int rows, int cols) { /* dst[c,r] = src[r,c] */
    for (int r = 0; r < rows; r++)
        for (int c = 0; c < cols; c++) dst[(size_t)c * rows + r] = src[(size_t)r * cols + c];
}

/* c/qwen_tts_codec.c:267-461; seq is [T, latent], overwritten with the output */
static void codec_transformer(const qtc_codec_t *c, float *seq, int T) {
    const int Hc = c->hidden, lat = c->latent, nh = c->heads, nkv = c->kv_heads;
    const int hd = Hc / nh, kvd = nkv * hd, qd = nh * hd, I = c->intermediate, W = c->sliding_window;
    const int rep = nh / nkv, half = hd / 2;
    const float eps = c->rms_eps, scale = 1.0f / sqrtf((float)hd);

    float *x = (float *)malloc((size_t)T * Hc * 4), *xn = (float *)malloc((size_t)T * Hc * 4);
    float *q = (float *)malloc((size_t)T * qd * 4), *k = (float *)malloc((size_t)T * kvd * 4);
    float *v = (float *)malloc((size_t)T * kvd * 4), *att = (float *)malloc((size_t)T * qd * 4);
    float *g = (float *)malloc((size_t)T * I * 4), *u = (float *)malloc((size_t)T * I * 4);
    float *cs = (float *)malloc((size_t)T * hd * 4), *sn = (float *)malloc((size_t)T * hd * 4);

    orc_matmul_f32(x, seq, c->in_proj_w, T, Hc, lat);
    if (c->in_proj_b) for (int t = 0; t < T; t++) for (int i = 0; i < Hc; i++) x[(size_t)t * Hc + i] += c->in_proj_b[i];
    for (int t = 0; t < T; t++) orc_rope_row(cs + (size_t)t * hd, sn + (size_t)t * hd, t, hd, 10000.0f); /* :304-320 */

    for (int l = 0; l < c->layers; l++) {
        const qtc_codec_layer_t *w = &c->layer[l];
        for (int t = 0; t < T; t++) orc_rms_norm(xn + (size_t)t * Hc, x + (size_t)t * Hc, w->input_norm, Hc, eps);
        orc_matmul_f32(q, xn, w->wq, T, qd, Hc);
        orc_matmul_f32(k, xn, w->wk, T, kvd, Hc);
        orc_matmul_f32(v, xn, w->wv, T, kvd, Hc);
        for (int t = 0; t < T; t++)
            for (int h = 0; h < nh + nkv; h++) {
                float *p = h < nh ? q + (size_t)t * qd + h * hd : k + (size_t)t * kvd + (h - nh) * hd;
                const float *cc = cs + (size_t)t * hd, *ss = sn + (size_t)t * hd;
                for (int i = 0; i < half; i++) {
                    const float a = p[i], b = p[i + half];
                    p[i] = a * cc[i] - b * ss[i];
                    p[i + half] = b * cc[i] + a * ss[i];
                }
            }
        /* sliding-window causal attention :355-383 */
#pragma omp parallel for schedule(static)
        for (int qi = 0; qi < T; qi++) {
            float sc[512];
            int start = qi - W + 1; if (start < 0) start = 0;
            const int n = qi - start + 1;
            for (int h = 0; h < nh; h++) {
                const int gk = h / rep;
                const float *qh = q + (size_t)qi * qd + h * hd;
                for (int i = 0; i < n; i++) {
                    const float *kh = k + (size_t)(start + i) * kvd + gk * hd;
                    float d = 0.0f;
                    for (int e = 0; e < hd; e++) d += qh[e] * kh[e];
                    sc[i] = d * scale;
                }
                orc_softmax(sc, n);
                float *o = att + (size_t)qi * qd + h * hd;
                for (int e = 0; e < hd; e++) o[e] = 0.0f;
                for (int i = 0; i < n; i++) {
                    const float *vh = v + (size_t)(start + i) * kvd + gk * hd;
                    const float p = sc[i];
                    for (int e = 0; e < hd; e++) o[e] += p * vh[e];
                }
            }
        }
        orc_matmul_f32(xn, att, w->wo, T, Hc, qd);
        for (int t = 0; t < T; t++)
            for (int i = 0; i < Hc; i++) {
                float y = xn[(size_t)t * Hc + i];
                if (w->attn_layer_scale) y *= w->attn_layer_scale[i];
                x[(size_t)t * Hc + i] += y;
            }
        for (int t = 0; t < T; t++) orc_rms_norm(xn + (size_t)t * Hc, x + (size_t)t * Hc, w->post_attn_norm, Hc, eps);
        orc_matmul_f32(g, xn, w->gate, T, I, Hc);
        orc_matmul_f32(u, xn, w->up, T, I, Hc);
        for (size_t i = 0; i < (size_t)T * I; i++) {
            float s = g[i] / (1.0f + expf(-g[i]));
            g[i] = s * u[i];
        }
        orc_matmul_f32(xn, g, w->down, T, Hc, I);
        for (int t = 0; t < T; t++)
            for (int i = 0; i < Hc; i++) {
                float y = xn[(size_t)t * Hc + i];
                if (w->mlp_layer_scale) y *= w->mlp_layer_scale[i];
                x[(size_t)t * Hc + i] += y;
            }
    }
    if (c->final_norm) for (int t = 0; t < T; t++) orc_rms_norm(x + (size_t)t * Hc, x + (size_t)t * Hc, c->final_norm, Hc, eps);
    orc_matmul_f32(seq, x, c->out_proj_w, T, lat, Hc);
    if (c->out_proj_b) for (int t = 0; t < T; t++) for (int i = 0; i < lat; i++) seq[(size_t)t * lat + i] += c->out_proj_b[i];
    free(x); free(xn); free(q); free(k); free(v); free(att); free(g); free(u); free(cs); free(sn);
}

/* c/qwen_tts_codec.c:467-522; h is [dim, L] channels-first, updated in place */
static void convnext(const qtc_convnext_t *b, float *h, int dim, int L) {
    const int d4 = 4 * dim;
    float *cv = (float *)malloc((size_t)dim * L * 4), *ld = (float *)malloc((size_t)L * dim * 4);
    float *pw = (float *)malloc((size_t)L * d4 * 4);
    orc_causal_conv1d(cv, h, b->dwconv_w, b->dwconv_b, dim, dim, 7, L, 1, dim);
    transpose(ld, cv, dim, L);                       /* -> [L, dim] */
#pragma omp parallel for schedule(static)
    for (int t = 0; t < L; t++) {
        float *row = ld + (size_t)t * dim, *mid = pw + (size_t)t * d4;
        orc_layer_norm(row, row, b->norm_w, b->norm_b, dim, 1e-6f);
        for (int r = 0; r < d4; r++) {
            const float *a = b->pw1_w + (size_t)r * dim;
            float acc = 0.0f;
            for (int i = 0; i < dim; i++) acc += a[i] * row[i];
            mid[r] = acc;
        }
        if (b->pw1_b) for (int r = 0; r < d4; r++) mid[r] += b->pw1_b[r];
        orc_gelu(mid, d4);
        for (int r = 0; r < dim; r++) {
            const float *a = b->pw2_w + (size_t)r * d4;
            float acc = 0.0f;
            for (int i = 0; i < d4; i++) acc += a[i] * mid[i];
            row[r] = acc;
        }
        if (b->pw2_b) for (int r = 0; r < dim; r++) row[r] += b->pw2_b[r];
        for (int r = 0; r < dim; r++) row[r] *= b->gamma[r];
    }
    for (int c = 0; c < dim; c++)
        for (int t = 0; t < L; t++) h[(size_t)c * L + t] = ld[(size_t)t * dim + c] + h[(size_t)c * L + t];
    free(cv); free(ld); free(pw);
}

/* c/qwen_tts_codec.c:549-575 */
static void resunit(const qtc_resunit_t *r, float *h, float *tmp1, float *tmp2, int dim, int L, int dil) {
    const size_t n = (size_t)dim * L;
    orc_snake_beta(tmp1, h, r->act1_alpha, r->act1_beta, dim, L);
    orc_causal_conv1d(tmp2, tmp1, r->conv1_w, r->conv1_b, dim, dim, 7, L, dil, 1);
    orc_snake_beta(tmp2, tmp2, r->act2_alpha, r->act2_beta, dim, L);
    orc_causal_conv1d(tmp1, tmp2, r->conv2_w, r->conv2_b, dim, dim, 1, L, 1, 1);
    for (size_t i = 0; i < n; i++) h[i] = tmp1[i] + h[i];
}

/* c/qwen_tts_codec.c:581-749 */
float *orc_codec_decode(orc_state_t *s, const int *codes, int T, int *out_samples) {
    const qtc_codec_t *c = &s->m->codec;
    if (!codes || !out_samples || T <= 0) { if (out_samples) *out_samples = 0; return NULL; }
    const int lat = c->latent;

    float *h = rvq_dequantize(c, codes, T);                                  /* [codebook_dim, T] */
    float *pre = (float *)malloc((size_t)lat * T * 4);
    orc_causal_conv1d(pre, h, c->pre_conv_w, c->pre_conv_b, c->codebook_dim, lat, 3, T, 1, 1);
    free(h);
    float *seq = (float *)malloc((size_t)T * lat * 4);
    transpose(seq, pre, lat, T);                                             /* [T, latent] */
    codec_transformer(c, seq, T);
    transpose(pre, seq, T, lat);                                             /* [latent, T] */
    free(seq);
    h = pre;

    int L = T;
    for (int st = 0; st < 2; st++) {                                         /* :644-663 */
        const int f = c->upsampling_ratios[st];
        float *up = (float *)malloc((size_t)lat * ((size_t)L * f + f) * 4);
        int nl;
        orc_transposed_conv1d(up, h, c->up_tconv_w[st], c->up_tconv_b[st], lat, lat, f, f, L, &nl);
        free(h); h = up; L = nl;
        convnext(&c->convnext[st], h, lat, L);
    }

    int dim = c->decoder_dim;                                                /* :665-719 */
    float *voc = (float *)malloc((size_t)dim * L * 4);
    orc_causal_conv1d(voc, h, c->voc_pre_w, c->voc_pre_b, lat, dim, 7, L, 1, 1);
    free(h);
    static const int dil[3] = {1, 3, 9};
    for (int b = 0; b < 4; b++) {
        const qtc_vocoder_block_t *vb = &c->block[b];
        const int r = c->upsample_rates[b], od = dim / 2;
        orc_snake_beta(voc, voc, vb->act_alpha, vb->act_beta, dim, L);
        float *up = (float *)malloc((size_t)od * ((size_t)L * r + 2 * r) * 4);
        int nl;
        orc_transposed_conv1d(up, voc, vb->tconv_w, vb->tconv_b, dim, od, 2 * r, r, L, &nl);
        free(voc); voc = up; L = nl; dim = od;
        float *t1 = (float *)malloc((size_t)dim * L * 4), *t2 = (float *)malloc((size_t)dim * L * 4);
        for (int u = 0; u < 3; u++) resunit(&vb->res[u], voc, t1, t2, dim, L, dil[u]);
        free(t1); free(t2);
    }
    orc_snake_beta(voc, voc, c->final_alpha, c->final_beta, dim, L);         /* :721-735 */
    float *wav = (float *)malloc((size_t)L * 4);
    orc_causal_conv1d(wav, voc, c->final_conv_w, c->final_conv_b, dim, 1, 7, L, 1, 1);
    free(voc);
    for (int i = 0; i < L; i++) { if (wav[i] < -1.0f) wav[i] = -1.0f; if (wav[i] > 1.0f) wav[i] = 1.0f; }
    *out_samples = L;
    return wav;
}

/* Load-time preprocessing the reference applies to codec tensors, restated so a
 * descriptor built by the test harness carries bit-identical values. */
/* c/qwen_tts.c:596-602 */
void orc_prep_snake(float *alpha, float *beta, int n) {
    for (int i = 0; i < n; i++) {
        alpha[i] = expf(alpha[i]);
        beta[i] = 1.0f / (expf(beta[i]) + 1e-9f);
    }
}
/* c/qwen_tts.c:577-594 */
void orc_prep_codebook(float *dst, const float *embedding_sum, const float *usage, int size, int dim) {
    for (int c = 0; c < size; c++) {
        float u = usage[c];
        if (u < 1e-5f) u = 1e-5f;
        const float inv = 1.0f / u;
        for (int d = 0; d < dim; d++) dst[(size_t)c * dim + d] = embedding_sum[(size_t)c * dim + d] * inv;
    }
}
