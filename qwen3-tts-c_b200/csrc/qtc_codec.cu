/*
 * qtc_codec.cu - codec decoder / vocoder on device: qwen_tts_codec_decode
 * (c/qwen_tts_codec.c:581-749).  Stages: RVQ de-quantise -> pre-conv ->
 * sliding-window transformer -> 2 x (transposed conv + ConvNeXt) -> vocoder
 * (pre-conv, 4 x {SnakeBeta, transposed conv, 3 x ResUnit}, SnakeBeta, final conv, clamp).
 * Activations stay in HBM between kernels as [channels, samples] fp32; SnakeBeta,
 * bias, residual adds and the final clamp are fused into the conv kernels.
 */
#include "qtc_codec_kernels.cuh"
#include "qtc_internal.h"

#include <math.h>

#define CLAUNCH(c)                                                                  \
    do {                                                                            \
        (c)->launches++;                                                            \
        cudaError_t e_ = cudaGetLastError();                                        \
        if (e_ != cudaSuccess) { qtc_set_error("%s:%d launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); return -1; } \
    } while (0)

struct TconvDev { const float *w; const float *b; int cin, cout, s, J; };   /* repacked [cout*s, cin*J] */
struct ResUnitDev { const float *a1, *b1, *w1, *c1, *a2, *b2, *w2, *c2; };
struct CodecLayerDev { const float *in_norm, *post_norm, *attn_scale, *mlp_scale, *wq, *wk, *wv, *wo, *gate, *up, *down; };
struct ConvNextDev { const float *dw_w, *dw_b, *norm_w, *norm_b, *pw1_w, *pw1_b, *pw2_w, *pw2_b, *gamma; };

struct CodecDev {
    qtc_codec_t h;                 /* dims */
    const float *cb[QTC_MAX_QUANTIZERS];
    const float *sem_proj, *aco_proj, *pre_w, *pre_b, *in_w, *in_b, *out_w, *out_b, *final_norm;
    CodecLayerDev layer[QTC_MAX_CODEC_LAYERS];
    TconvDev up_tconv[2];
    ConvNextDev cnx[2];
    const float *voc_pre_w, *voc_pre_b;
    struct { const float *alpha, *beta; TconvDev tc; ResUnitDev ru[3]; } block[4];
    const float *fin_alpha, *fin_beta, *fin_w, *fin_b;
    /* workspace, grown on demand */
    float *ws; size_t ws_bytes;
    int *codes_dev; int codes_cap;
    float *rope_cs, *rope_sn; int rope_T;
};

/* transposed conv weight [cin, cout, k] -> rows (oc*s + p) x (ic*J + j): tap p + (J-1-j)*s (0 if >= k).
 * With it, y[oc, m*s+p] = b + sum_{ic,j} W'[oc*s+p, ic*J+j] * x[ic, m-(J-1-j)] is kernel_transposed_conv1d
 * (c/qwen_tts_kernels.c:873-972) after its right trim, in the same (ic, t) accumulation order. */
static const float *repack_tconv(qtc_ctx *c, const float *w, int cin, int cout, int k, int s, int *J_out) {
    const int J = (k + s - 1) / s;
    std::vector<float> r((size_t)cout * s * cin * J, 0.0f);
    for (int oc = 0; oc < cout; oc++)
        for (int p = 0; p < s; p++)
            for (int ic = 0; ic < cin; ic++)
                for (int j = 0; j < J; j++) {
                    const int tap = p + (J - 1 - j) * s;
                    if (tap < k) r[((size_t)(oc * s + p) * cin + ic) * J + j] = w[((size_t)ic * cout + oc) * k + tap];
                }
    *J_out = J;
    return qtc_upload(c, r.data(), r.size());
}

CodecDev *qtc_codec_upload(qtc_ctx *c, const qtc_codec_t *k) {
    CodecDev *d = new CodecDev();
    memset(d, 0, sizeof(*d));
    d->h = *k;
    const size_t vq = k->codebook_dim / 2, half = k->latent / 2, lat = k->latent, Hc = k->hidden, I = k->intermediate;
    const size_t hd = Hc / k->heads, qd = k->heads * hd, kvd = k->kv_heads * hd;
    bool ok = true;
#define UP(dst, src, n) do { (dst) = qtc_upload(c, (src), (size_t)(n)); if ((src) && !(dst)) ok = false; } while (0)
#define NEED(p) do { if (!(p)) { qtc_set_error("codec tensor missing: %s", #p); ok = false; } } while (0)
    for (int q = 0; q < k->num_quantizers; q++) { NEED(k->codebooks[q]); UP(d->cb[q], k->codebooks[q], (size_t)k->codebook_size * vq); }
    NEED(k->semantic_out_proj); NEED(k->acoustic_out_proj); NEED(k->pre_conv_w); NEED(k->in_proj_w); NEED(k->out_proj_w);
    NEED(k->voc_pre_w); NEED(k->final_conv_w); NEED(k->final_alpha); NEED(k->final_beta);
    if (!ok) { delete d; return NULL; }
    UP(d->sem_proj, k->semantic_out_proj, half * vq); UP(d->aco_proj, k->acoustic_out_proj, half * vq);
    UP(d->pre_w, k->pre_conv_w, lat * k->codebook_dim * 3); UP(d->pre_b, k->pre_conv_b, lat);
    UP(d->in_w, k->in_proj_w, Hc * lat); UP(d->in_b, k->in_proj_b, Hc);
    UP(d->out_w, k->out_proj_w, lat * Hc); UP(d->out_b, k->out_proj_b, lat);
    UP(d->final_norm, k->final_norm, Hc);
    for (int l = 0; l < k->layers; l++) {
        const qtc_codec_layer_t *s = &k->layer[l];
        CodecLayerDev *o = &d->layer[l];
        NEED(s->wq); NEED(s->wk); NEED(s->wv); NEED(s->wo); NEED(s->gate); NEED(s->up); NEED(s->down); NEED(s->input_norm); NEED(s->post_attn_norm);
        if (!ok) break;
        UP(o->in_norm, s->input_norm, Hc); UP(o->post_norm, s->post_attn_norm, Hc);
        UP(o->attn_scale, s->attn_layer_scale, Hc); UP(o->mlp_scale, s->mlp_layer_scale, Hc);
        UP(o->wq, s->wq, qd * Hc); UP(o->wk, s->wk, kvd * Hc); UP(o->wv, s->wv, kvd * Hc); UP(o->wo, s->wo, Hc * qd);
        UP(o->gate, s->gate, I * Hc); UP(o->up, s->up, I * Hc); UP(o->down, s->down, Hc * I);
    }
    for (int s = 0; s < 2 && ok; s++) {
        const int f = k->upsampling_ratios[s];
        NEED(k->up_tconv_w[s]);
        if (!ok) break;
        d->up_tconv[s].w = repack_tconv(c, k->up_tconv_w[s], (int)lat, (int)lat, f, f, &d->up_tconv[s].J);
        UP(d->up_tconv[s].b, k->up_tconv_b[s], lat);
        d->up_tconv[s].cin = d->up_tconv[s].cout = (int)lat; d->up_tconv[s].s = f;
        const qtc_convnext_t *b = &k->convnext[s];
        ConvNextDev *o = &d->cnx[s];
        NEED(b->dwconv_w); NEED(b->pw1_w); NEED(b->pw2_w); NEED(b->gamma);
        UP(o->dw_w, b->dwconv_w, lat * 7); UP(o->dw_b, b->dwconv_b, lat); UP(o->norm_w, b->norm_w, lat); UP(o->norm_b, b->norm_b, lat);
        UP(o->pw1_w, b->pw1_w, 4 * lat * lat); UP(o->pw1_b, b->pw1_b, 4 * lat);
        UP(o->pw2_w, b->pw2_w, 4 * lat * lat); UP(o->pw2_b, b->pw2_b, lat); UP(o->gamma, b->gamma, lat);
    }
    UP(d->voc_pre_w, k->voc_pre_w, (size_t)k->decoder_dim * lat * 7); UP(d->voc_pre_b, k->voc_pre_b, k->decoder_dim);
    int dim = k->decoder_dim;
    for (int b = 0; b < 4 && ok; b++) {
        const qtc_vocoder_block_t *s = &k->block[b];
        const int r = k->upsample_rates[b], od = dim / 2;
        NEED(s->act_alpha); NEED(s->act_beta); NEED(s->tconv_w);
        if (!ok) break;
        UP(d->block[b].alpha, s->act_alpha, dim); UP(d->block[b].beta, s->act_beta, dim);
        d->block[b].tc.w = repack_tconv(c, s->tconv_w, dim, od, 2 * r, r, &d->block[b].tc.J);
        UP(d->block[b].tc.b, s->tconv_b, od);
        d->block[b].tc.cin = dim; d->block[b].tc.cout = od; d->block[b].tc.s = r;
        for (int u = 0; u < 3; u++) {
            const qtc_resunit_t *ru = &s->res[u];
            ResUnitDev *o = &d->block[b].ru[u];
            NEED(ru->conv1_w); NEED(ru->conv2_w); NEED(ru->act1_alpha); NEED(ru->act2_alpha);
            UP(o->a1, ru->act1_alpha, od); UP(o->b1, ru->act1_beta, od); UP(o->w1, ru->conv1_w, (size_t)od * od * 7); UP(o->c1, ru->conv1_b, od);
            UP(o->a2, ru->act2_alpha, od); UP(o->b2, ru->act2_beta, od); UP(o->w2, ru->conv2_w, (size_t)od * od); UP(o->c2, ru->conv2_b, od);
        }
        dim = od;
    }
    UP(d->fin_alpha, k->final_alpha, dim); UP(d->fin_beta, k->final_beta, dim);
    UP(d->fin_w, k->final_conv_w, (size_t)dim * 7); UP(d->fin_b, k->final_conv_b, 1);
#undef UP
#undef NEED
    if (!ok) { delete d; return NULL; }
    return d;
}

void qtc_codec_release(CodecDev *d) {
    if (!d) return;
    cudaFree(d->ws); cudaFree(d->codes_dev); cudaFree(d->rope_cs); cudaFree(d->rope_sn);
    delete d;
}

/* ---- launch helpers ------------------------------------------------------------ */
static int conv_smem_bytes(int BM, int K, int dil) {
    int CI = CV_MAXCK / K; if (CI > 32) CI = 32; if (CI < 1) CI = 1;
    return (CI * K * (BM + 4) + CI * (CV_BN + (K - 1) * dil + 1)) * 4;
}
static int g_conv_attr = 0;

int qtc_conv(qtc_ctx *c, float *out, const float *in, const float *w, const float *bias, const float *res, const float *alpha,
             const float *inv_beta, int cin, int rows, int K, int dil, int Lin, int up) {
    if (!g_conv_attr) {
        QTC_CUDA(cudaFuncSetAttribute(k_conv_tile<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        QTC_CUDA(cudaFuncSetAttribute(k_conv_tile<96>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        QTC_CUDA(cudaFuncSetAttribute(k_conv_tile<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        g_conv_attr = 1;
    }
    ConvArgs a;
    a.in = in; a.w = w; a.bias = bias; a.res = res; a.alpha = alpha; a.inv_beta = inv_beta; a.out = out;
    a.Cin = cin; a.rows = rows; a.K = K; a.dil = dil; a.Lin = Lin; a.up = up;
    const int gx = (Lin + CV_BN - 1) / CV_BN;
    /* tile height: 96 where it divides the rows better than 128 (96-, 192-, 288-channel layers) */
    const int w128 = ((rows + 127) / 128) * 128, w96 = ((rows + 95) / 96) * 96, w64 = ((rows + 63) / 64) * 64;
    if (gx <= 0 || rows <= 0) return 0;
    if (w64 < w96 && w64 < w128) {
        const int sm = conv_smem_bytes(64, K, dil);
        if (sm > 200 * 1024) { qtc_set_error("conv: dilation too large for the tile"); return -1; }
        k_conv_tile<64><<<dim3(gx, w64 / 64), 128, sm, c->stream>>>(a);
    } else if (w96 < w128) {
        const int sm = conv_smem_bytes(96, K, dil);
        if (sm > 200 * 1024) { qtc_set_error("conv: dilation too large for the tile"); return -1; }
        k_conv_tile<96><<<dim3(gx, w96 / 96), 192, sm, c->stream>>>(a);
    } else {
        const int sm = conv_smem_bytes(128, K, dil);
        if (sm > 200 * 1024) { qtc_set_error("conv: dilation too large for the tile"); return -1; }
        k_conv_tile<128><<<dim3(gx, w128 / 128), 256, sm, c->stream>>>(a);
    }
    CLAUNCH(c);
    return 0;
}

int qtc_sgemm(qtc_ctx *c, float *C, const float *A, const float *B, int M, int N, int K, const float *bias, const float *scale,
              const float *res, int gelu, int transposed, int ldc) {
    if (K % 4 != 0) { qtc_set_error("sgemm: K %% 4 != 0 (K=%d)", K); return -1; }
    if (M <= 0 || N <= 0) return 0;
    GemmEpi e;
    e.bias = bias; e.scale = scale; e.res = res; e.gelu = gelu; e.transposed = transposed; e.ldc = ldc;
    k_sgemm_nt<<<dim3((N + SG_BN - 1) / SG_BN, (M + SG_BM - 1) / SG_BM), 256, 0, c->stream>>>(C, A, B, M, N, K, e);
    CLAUNCH(c);
    return 0;
}

int qtc_conv_grouped(qtc_ctx *c, float *out, const float *in, const float *w, const float *bias, int cin, int cout, int K, int L, int dil, int groups) {
    k_conv_grouped<<<dim3((L + 127) / 128, cout), 128, 0, c->stream>>>(out, in, w, bias, cin, cout, K, L, dil, groups);
    CLAUNCH(c);
    return 0;
}
int qtc_snake(qtc_ctx *c, float *out, const float *x, const float *alpha, const float *inv_beta, int C, int L) {
    const size_t n = (size_t)C * L;
    k_snake<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(out, x, alpha, inv_beta, C, (size_t)L);
    CLAUNCH(c);
    return 0;
}

static int transpose(qtc_ctx *c, float *dst, const float *src, int rows, int cols) {
    k_transpose<<<dim3((cols + 31) / 32, (rows + 31) / 32), dim3(32, 8), 0, c->stream>>>(dst, src, rows, cols);
    CLAUNCH(c);
    return 0;
}
static int row_norm(qtc_ctx *c, float *out, const float *x, const float *w, const float *b, int rows, int n, float eps, int mode) {
    k_row_norm<<<rows, 256, 0, c->stream>>>(out, x, w, b, n, eps, mode);
    CLAUNCH(c);
    return 0;
}

static int tconv(qtc_ctx *c, float *out, const float *in, const TconvDev &t, const float *alpha, const float *beta, int Lin) {
    return qtc_conv(c, out, in, t.w, t.b, NULL, alpha, beta, t.cin, t.cout * t.s, t.J, 1, Lin, t.s);
}

/* workspace carve-up: a bump allocator over one grown-on-demand arena */
struct Arena { char *base; size_t off, cap; float *take(size_t n) { off = (off + 255) & ~(size_t)255; float *p = (float *)(base + off); off += n * 4; return p; } };

/* The workspace is two halves used as a ping-pong: every stage reads its input from one half and puts
 * its temporaries and output in the other.  Returns the bytes ONE half must hold. */
static size_t codec_half_bytes(const qtc_codec_t *k, int T) {
    const size_t lat = k->latent, Hc = k->hidden, I = k->intermediate, vq = k->codebook_dim / 2;
    const size_t hd = Hc / k->heads, qd = k->heads * hd, kvd = k->kv_heads * hd;
    size_t L = T;
    size_t need = (size_t)T * (2 * vq + lat / 2 + 2 * lat + 2 * Hc + 2 * qd + 2 * kvd + 2 * I);      /* transformer stage */
    for (int s = 0; s < 2; s++) { L *= k->upsampling_ratios[s]; const size_t b = L * lat * 7; if (b > need) need = b; }
    size_t dim = k->decoder_dim;
    if (L * dim > need) need = L * dim;
    for (int b = 0; b < 4; b++) {
        const size_t Lo = L * k->upsample_rates[b], od = dim / 2;
        if (2 * Lo * od > need) need = 2 * Lo * od;
        L = Lo; dim = od;
    }
    return need * 4 + 64 * 256;       /* + alignment slack for the bump allocator */
}

/* codes_dev [T, nq] on device.  Returns malloc'd host PCM (caller frees); *dev_out (optional) receives the
 * device pointer of the PCM inside the workspace (valid until the next decode). */
float *qtc_codec_decode_dev(qtc_ctx *c, const int *codes_dev, int T, int *out_samples, float **dev_out) {
    CodecDev *d = c->codec;
    if (out_samples) *out_samples = 0;
    if (!d || !codes_dev || T <= 0 || !out_samples) return NULL;
    const qtc_codec_t *k = &d->h;
    const int lat = k->latent, Hc = k->hidden, I = k->intermediate, nh = k->heads, nkv = k->kv_heads, hd = Hc / nh;
    const int qd = nh * hd, kvd = nkv * hd, vq = k->codebook_dim / 2, half = lat / 2;
    if (k->sliding_window > 128 || hd > 256 || hd % 4) { qtc_set_error("codec: window %d / head_dim %d unsupported", k->sliding_window, hd); return NULL; }
    if (k->codebook_dim != half * 1 && k->codebook_dim != lat / 2) { qtc_set_error("codec: codebook_dim must equal latent/2 (SURVEY 7.0)"); return NULL; }
    const size_t half_bytes = (codec_half_bytes(k, T) + 255) & ~(size_t)255;
    const size_t need = 2 * half_bytes;
    if (need > d->ws_bytes) {
        cudaStreamSynchronize(c->stream);
        cudaFree(d->ws); d->ws = NULL; d->ws_bytes = 0;
        if (cudaMalloc(&d->ws, need) != cudaSuccess) { qtc_set_error("codec: workspace of %.1f MB failed", need / 1e6); return NULL; }
        d->ws_bytes = need;
    }
    if (T > d->rope_T) {          /* codec RoPE: theta 10000 hard-coded (c/qwen_tts_codec.c:309), host libm */
        cudaStreamSynchronize(c->stream);
        cudaFree(d->rope_cs); cudaFree(d->rope_sn);
        const int n = T + 256;
        std::vector<float> cs((size_t)n * hd), sn((size_t)n * hd);
        for (int t = 0; t < n; t++)
            for (int i = 0; i < hd / 2; i++) {
                const float freq = 1.0f / powf(10000.0f, (float)(2 * i) / (float)hd), ang = (float)t * freq;
                cs[(size_t)t * hd + i] = cs[(size_t)t * hd + i + hd / 2] = cosf(ang);
                sn[(size_t)t * hd + i] = sn[(size_t)t * hd + i + hd / 2] = sinf(ang);
            }
        if (cudaMalloc(&d->rope_cs, cs.size() * 4) != cudaSuccess || cudaMalloc(&d->rope_sn, sn.size() * 4) != cudaSuccess) return NULL;
        cudaMemcpy(d->rope_cs, cs.data(), cs.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d->rope_sn, sn.data(), sn.size() * 4, cudaMemcpyHostToDevice);
        d->rope_T = n;
    }
#define CK(x) do { if ((x) != 0) return NULL; } while (0)
    Arena ar = {(char *)d->ws, 0, half_bytes};
    /* ---- RVQ + pre-conv + transformer ---- */
    float *sem = ar.take((size_t)T * vq), *aco = ar.take((size_t)T * vq);
    float *h0 = ar.take((size_t)half * T), *pre = ar.take((size_t)lat * T), *seq = ar.take((size_t)T * lat);
    float *x = ar.take((size_t)T * Hc), *xn = ar.take((size_t)T * Hc), *q = ar.take((size_t)T * qd), *kk = ar.take((size_t)T * kvd);
    float *v = ar.take((size_t)T * kvd), *att = ar.take((size_t)T * qd), *g = ar.take((size_t)T * I), *u = ar.take((size_t)T * I);
    {
        RvqArgs ra;
        for (int qi = 0; qi < k->num_quantizers; qi++) ra.cb[qi] = d->cb[qi];
        ra.codes = codes_dev; ra.sem = sem; ra.aco = aco; ra.T = T; ra.nq = k->num_quantizers; ra.vq = vq; ra.cb_size = k->codebook_size;
        k_rvq_gather<<<T, 128, 0, c->stream>>>(ra);
        c->launches++;
    }
    CK(qtc_sgemm(c, h0, sem, d->sem_proj, T, half, vq, NULL, NULL, NULL, 0, 1, T));           /* -> [half, T] */
    CK(qtc_sgemm(c, h0, aco, d->aco_proj, T, half, vq, NULL, NULL, h0, 0, 1, T));
    CK(qtc_conv(c, pre, h0, d->pre_w, d->pre_b, NULL, NULL, NULL, k->codebook_dim, lat, 3, 1, T, 1));
    CK(transpose(c, seq, pre, lat, T));                                                         /* [T, lat] */
    CK(qtc_sgemm(c, x, seq, d->in_w, T, Hc, lat, d->in_b, NULL, NULL, 0, 0, Hc));
    const float scale = 1.0f / sqrtf((float)hd);
    for (int l = 0; l < k->layers; l++) {
        const CodecLayerDev *w = &d->layer[l];
        CK(row_norm(c, xn, x, w->in_norm, NULL, T, Hc, k->rms_eps, 0));
        CK(qtc_sgemm(c, q, xn, w->wq, T, qd, Hc, NULL, NULL, NULL, 0, 0, qd));
        CK(qtc_sgemm(c, kk, xn, w->wk, T, kvd, Hc, NULL, NULL, NULL, 0, 0, kvd));
        CK(qtc_sgemm(c, v, xn, w->wv, T, kvd, Hc, NULL, NULL, NULL, 0, 0, kvd));
        k_rope_rows<<<T, 256, 0, c->stream>>>(q, kk, d->rope_cs, d->rope_sn, T, nh, nkv, hd);
        c->launches++;
        k_window_attn<<<(T * nh + 3) / 4, 128, 0, c->stream>>>(att, q, kk, v, T, nh, nkv, hd, k->sliding_window, scale);
        c->launches++;
        CK(qtc_sgemm(c, x, att, w->wo, T, Hc, qd, NULL, w->attn_scale, x, 0, 0, Hc));
        CK(row_norm(c, xn, x, w->post_norm, NULL, T, Hc, k->rms_eps, 0));
        CK(qtc_sgemm(c, g, xn, w->gate, T, I, Hc, NULL, NULL, NULL, 0, 0, I));
        CK(qtc_sgemm(c, u, xn, w->up, T, I, Hc, NULL, NULL, NULL, 0, 0, I));
        k_swiglu_mul<<<(unsigned)(((size_t)T * I + 255) / 256), 256, 0, c->stream>>>(g, u, (size_t)T * I);
        c->launches++;
        CK(qtc_sgemm(c, x, g, w->down, T, Hc, I, NULL, w->mlp_scale, x, 0, 0, Hc));
    }
    if (d->final_norm) CK(row_norm(c, x, x, d->final_norm, NULL, T, Hc, k->rms_eps, 0));
    CK(qtc_sgemm(c, pre, x, d->out_w, T, lat, Hc, d->out_b, NULL, NULL, 0, 1, T));            /* -> [lat, T] */

    /* ---- 2 x (transposed conv + ConvNeXt) ---- */
    int L = T, side = 1;                 /* the transformer stage used half 0; `side` = half the next stage writes */
    float *cur = pre;
    auto half_arena = [&](int sd) { Arena a = {(char *)d->ws + (size_t)sd * half_bytes, 0, half_bytes}; return a; };
    for (int s = 0; s < 2; s++) {
        const int f = k->upsampling_ratios[s], Lo = L * f;
        Arena a2 = half_arena(side);
        float *up = a2.take((size_t)lat * Lo), *cv = a2.take((size_t)lat * Lo), *ld = a2.take((size_t)Lo * lat), *pw = a2.take((size_t)Lo * 4 * lat);
        if (a2.off > half_bytes) { qtc_set_error("codec: workspace underestimated (upsample)"); return NULL; }
        CK(tconv(c, up, cur, d->up_tconv[s], NULL, NULL, L));
        const ConvNextDev *b = &d->cnx[s];
        k_dwconv<<<dim3((Lo + 255) / 256, lat), 256, 0, c->stream>>>(cv, up, b->dw_w, b->dw_b, lat, 7, Lo, 1);
        c->launches++;
        CK(transpose(c, ld, cv, lat, Lo));                                                      /* [Lo, lat] */
        CK(row_norm(c, ld, ld, b->norm_w, b->norm_b, Lo, lat, 1e-6f, 1));
        CK(qtc_sgemm(c, pw, ld, b->pw1_w, Lo, 4 * lat, lat, b->pw1_b, NULL, NULL, 1, 0, 4 * lat));
        CK(qtc_sgemm(c, up, pw, b->pw2_w, Lo, lat, 4 * lat, b->pw2_b, b->gamma, up, 0, 1, Lo)); /* h[c,t] += gamma*(..) */
        cur = up; L = Lo; side ^= 1;
    }
    /* ---- vocoder ---- */
    int dim = k->decoder_dim;
    Arena a3 = half_arena(side);
    float *voc = a3.take((size_t)dim * L);
    if (a3.off > half_bytes) { qtc_set_error("codec: workspace underestimated (vocoder pre-conv)"); return NULL; }
    CK(qtc_conv(c, voc, cur, d->voc_pre_w, d->voc_pre_b, NULL, NULL, NULL, lat, dim, 7, 1, L, 1));
    side ^= 1;
    static const int dil[3] = {1, 3, 9};
    for (int b = 0; b < 4; b++) {
        const int r = k->upsample_rates[b], od = dim / 2, Lo = L * r;
        Arena a4 = half_arena(side);
        float *hh = a4.take((size_t)od * Lo), *t1 = a4.take((size_t)od * Lo);
        if (a4.off > half_bytes) { qtc_set_error("codec: workspace underestimated (vocoder)"); return NULL; }
        CK(tconv(c, hh, voc, d->block[b].tc, d->block[b].alpha, d->block[b].beta, L));
        for (int uu = 0; uu < 3; uu++) {
            const ResUnitDev *ru = &d->block[b].ru[uu];
            CK(qtc_conv(c, t1, hh, ru->w1, ru->c1, NULL, ru->a1, ru->b1, od, od, 7, dil[uu], Lo, 1));
            CK(qtc_conv(c, hh, t1, ru->w2, ru->c2, hh, ru->a2, ru->b2, od, od, 1, 1, Lo, 1));
        }
        voc = hh; L = Lo; dim = od; side ^= 1;
    }
    Arena a5 = half_arena(side);
    float *wav = a5.take((size_t)L);
    k_conv_to_mono<<<(L + 255) / 256, 256, 0, c->stream>>>(wav, voc, d->fin_w, d->fin_b, d->fin_alpha, d->fin_beta, dim, 7, L, 1);
    c->launches++;
#undef CK
    cudaEventRecord(c->ev3, c->stream);              /* end of device work; what follows is the D2H of the PCM */
    float *host = (float *)malloc((size_t)L * 4);
    if (!host) return NULL;
    if (cudaMemcpyAsync(host, wav, (size_t)L * 4, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
        qtc_set_error("codec: device error %s", cudaGetErrorString(cudaGetLastError()));
        free(host);
        return NULL;
    }
    *out_samples = L;
    if (dev_out) *dev_out = wav;
    return host;
}

extern "C" float *qwen_tts_cuda_codec_decode(qtc_ctx_t *c, const int *codes, int T, int *out_samples) {
    if (out_samples) *out_samples = 0;
    if (!c || !c->codec) { qtc_set_error("codec_decode: context has no codec"); return NULL; }
    if (!codes || T <= 0 || !out_samples) return NULL;          /* c/qwen_tts_codec.c:583-587 */
    CodecDev *d = c->codec;
    const int nq = d->h.num_quantizers;
    if (T > d->codes_cap) {
        cudaStreamSynchronize(c->stream);
        cudaFree(d->codes_dev);
        if (cudaMalloc(&d->codes_dev, (size_t)(T + 64) * nq * 4) != cudaSuccess) { qtc_set_error("codec: codes buffer"); return NULL; }
        d->codes_cap = T + 64;
    }
    cudaEventRecord(c->ev0, c->stream);
    if (cudaMemcpyAsync(d->codes_dev, codes, (size_t)T * nq * 4, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) return NULL;
    float *pcm = qtc_codec_decode_dev(c, d->codes_dev, T, out_samples, NULL);
    cudaEventRecord(c->ev1, c->stream);
    cudaEventSynchronize(c->ev1);
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->perf.codec_ms = ms;
    if (pcm) { float dms = 0; cudaEventElapsedTime(&dms, c->ev0, c->ev3); c->perf.device_ms = dms; }   /* codes H2D .. last codec kernel */
    return pcm;
}
