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
    /* the same vocoder weights as bf16 hi/lo planes for the tensor-core path (qtc_conv_tc.cu); tc_ok: every layer is tileable */
    ConvTcW tc_pre, tc_t[4], tc_c1[4][3], tc_c2[4][3];
    bool tc_ok;
    /* ... and the transformer / upsample projections (k = 1 "convolutions" over time-major rows); tcx_ok: they tile too */
    struct { ConvTcW q, k, v, o, gate, up, down; } tcl[QTC_MAX_CODEC_LAYERS];
    ConvTcW tc_in, tc_out, tc_up[2], tc_pw1[2], tc_pw2[2];
    bool tcx_ok;
    /* workspace, grown on demand */
    float *ws; size_t ws_bytes;
    int *codes_dev; int codes_cap;
    float *rope_cs, *rope_sn; int rope_T;
    size_t ws_half;
    /* streaming session (qtc_codec_stream_*): frames decoded so far, capacity, conv-stack context in frames */
    int s_T, s_cap, s_ctx, s_max;
    float *s_tf;                                    /* [cap, latent] transformer output rows, time-major */
    float *s_k[QTC_MAX_CODEC_LAYERS], *s_v[QTC_MAX_CODEC_LAYERS];   /* [cap, kv_heads*hd] per layer, by absolute position */
    int *s_codes;                                   /* [cap, nq] codes pushed so far */
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
    /* tensor-core path: all vocoder channel counts must tile (multiples of 16 out, 8 in); the decoder of the 12 Hz tokenizer
     * (1536 / 768 / 384 / 192 / 96) does, toy decoders fall back to the FFMA kernels */
    d->tc_ok = ok && qtc_conv_tc_ok((int)lat, k->decoder_dim) && dim % 2 == 0;
    for (int b = 0, dd = k->decoder_dim; b < 4 && d->tc_ok; b++, dd /= 2) d->tc_ok = qtc_conv_tc_ok(dd, dd / 2) && qtc_conv_tc_ok(dd / 2, dd / 2);
    if (d->tc_ok) {
        bool g = qtc_conv_tc_prep(c, &d->tc_pre, k->voc_pre_w, k->decoder_dim, (int)lat, 7) == 0;
        for (int b = 0, dd = k->decoder_dim; b < 4 && g; b++, dd /= 2) {
            const int r = k->upsample_rates[b], od = dd / 2;
            g = qtc_tconv_tc_prep(c, &d->tc_t[b], k->block[b].tconv_w, dd, od, 2 * r, r) == 0;
            for (int u = 0; u < 3 && g; u++)
                g = qtc_conv_tc_prep(c, &d->tc_c1[b][u], k->block[b].res[u].conv1_w, od, od, 7) == 0 &&
                    qtc_conv_tc_prep(c, &d->tc_c2[b][u], k->block[b].res[u].conv2_w, od, od, 1) == 0;
        }
        if (!g) ok = false;
    }
    d->tcx_ok = d->tc_ok && ok && qtc_conv_tc_ok((int)lat, (int)Hc) && qtc_conv_tc_ok((int)Hc, (int)qd) && qtc_conv_tc_ok((int)Hc, (int)kvd) &&
                qtc_conv_tc_ok((int)qd, (int)Hc) && qtc_conv_tc_ok((int)Hc, (int)I) && qtc_conv_tc_ok((int)I, (int)Hc) && qtc_conv_tc_ok((int)Hc, (int)lat) &&
                lat % 32 == 0;
    if (d->tcx_ok) {
        bool g = qtc_conv_tc_prep(c, &d->tc_in, k->in_proj_w, (int)Hc, (int)lat, 1) == 0 && qtc_conv_tc_prep(c, &d->tc_out, k->out_proj_w, (int)lat, (int)Hc, 1) == 0;
        for (int l = 0; l < k->layers && g; l++) {
            const qtc_codec_layer_t *s = &k->layer[l];
            g = qtc_conv_tc_prep(c, &d->tcl[l].q, s->wq, (int)qd, (int)Hc, 1) == 0 && qtc_conv_tc_prep(c, &d->tcl[l].k, s->wk, (int)kvd, (int)Hc, 1) == 0 &&
                qtc_conv_tc_prep(c, &d->tcl[l].v, s->wv, (int)kvd, (int)Hc, 1) == 0 && qtc_conv_tc_prep(c, &d->tcl[l].o, s->wo, (int)Hc, (int)qd, 1) == 0 &&
                qtc_conv_tc_prep(c, &d->tcl[l].gate, s->gate, (int)I, (int)Hc, 1) == 0 && qtc_conv_tc_prep(c, &d->tcl[l].up, s->up, (int)I, (int)Hc, 1) == 0 &&
                qtc_conv_tc_prep(c, &d->tcl[l].down, s->down, (int)Hc, (int)I, 1) == 0;
        }
        for (int s = 0; s < 2 && g; s++) {
            const int f = k->upsampling_ratios[s];
            g = qtc_tconv_tc_prep(c, &d->tc_up[s], k->up_tconv_w[s], (int)lat, (int)lat, f, f) == 0 &&
                qtc_conv_tc_prep(c, &d->tc_pw1[s], k->convnext[s].pw1_w, 4 * (int)lat, (int)lat, 1) == 0 &&
                qtc_conv_tc_prep(c, &d->tc_pw2[s], k->convnext[s].pw2_w, (int)lat, 4 * (int)lat, 1) == 0;
        }
        if (!g) ok = false;
    }
#undef UP
#undef NEED
    if (!ok) { delete d; return NULL; }
    return d;
}

static void stream_free(CodecDev *d);
void qtc_codec_release(CodecDev *d) {
    if (!d) return;
    cudaFree(d->ws); cudaFree(d->codes_dev); cudaFree(d->rope_cs); cudaFree(d->rope_sn);
    stream_free(d);
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
        k_conv_tile<64><<<dim3(gx, w64 / 64), 128, sm, c->cstream>>>(a);
    } else if (w96 < w128) {
        const int sm = conv_smem_bytes(96, K, dil);
        if (sm > 200 * 1024) { qtc_set_error("conv: dilation too large for the tile"); return -1; }
        k_conv_tile<96><<<dim3(gx, w96 / 96), 192, sm, c->cstream>>>(a);
    } else {
        const int sm = conv_smem_bytes(128, K, dil);
        if (sm > 200 * 1024) { qtc_set_error("conv: dilation too large for the tile"); return -1; }
        k_conv_tile<128><<<dim3(gx, w128 / 128), 256, sm, c->cstream>>>(a);
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
    k_sgemm_nt<<<dim3((N + SG_BN - 1) / SG_BN, (M + SG_BM - 1) / SG_BM), 256, 0, c->cstream>>>(C, A, B, M, N, K, e);
    CLAUNCH(c);
    return 0;
}

int qtc_conv_grouped(qtc_ctx *c, float *out, const float *in, const float *w, const float *bias, int cin, int cout, int K, int L, int dil, int groups) {
    k_conv_grouped<<<dim3((L + 127) / 128, cout), 128, 0, c->cstream>>>(out, in, w, bias, cin, cout, K, L, dil, groups);
    CLAUNCH(c);
    return 0;
}
int qtc_snake(qtc_ctx *c, float *out, const float *x, const float *alpha, const float *inv_beta, int C, int L) {
    const size_t n = (size_t)C * L;
    k_snake<<<(unsigned)((n + 255) / 256), 256, 0, c->cstream>>>(out, x, alpha, inv_beta, C, (size_t)L);
    CLAUNCH(c);
    return 0;
}

static int transpose(qtc_ctx *c, float *dst, const float *src, int rows, int cols) {
    k_transpose<<<dim3((cols + 31) / 32, (rows + 31) / 32), dim3(32, 8), 0, c->cstream>>>(dst, src, rows, cols);
    CLAUNCH(c);
    return 0;
}
static int row_norm(qtc_ctx *c, float *out, const float *x, const float *w, const float *b, int rows, int n, float eps, int mode) {
    k_row_norm<<<rows, 256, 0, c->cstream>>>(out, x, w, b, n, eps, mode);
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
    size_t need = (size_t)T * (2 * vq + lat / 2 + 3 * lat + 4 * Hc + 2 * qd + 2 * kvd + 3 * I);      /* transformer stage */
    for (int s = 0; s < 2; s++) { L *= k->upsampling_ratios[s]; const size_t b = L * lat * 9; if (b > need) need = b; }
    size_t dim = k->decoder_dim;
    if (L * dim > need) need = L * dim;
    for (int b = 0; b < 4; b++) {
        const size_t Lo = L * k->upsample_rates[b], od = dim / 2;
        if (3 * Lo * od > need) need = 3 * Lo * od;          /* raw rows + two pairs of bf16 planes (tensor-core path); 2 x fp32 otherwise */
        L = Lo; dim = od;
    }
    return need * 4 + 64 * 256;       /* + alignment slack for the bump allocator */
}

/* ---- streaming state ------------------------------------------------------------------------------------------
 * Every layer of the decoder is causal, so frames can be decoded in chunks with EXACTLY the one-shot result:
 *   - the transformer keeps its per-layer K (after RoPE) and V rows by absolute position and attends over the window of
 *     72 frames across chunk borders; its output rows are kept time-major in `tf`;
 *   - everything after it is convolutions whose total receptive field is < 10 frames (ctx_frames below): a chunk re-runs
 *     the conv stack on [T - ctx, T + n) and drops the first ctx frames of PCM.  Products and summation order of every kept
 *     sample are those of the one-shot decode (accumulation order in k_conv_tile / k_sgemm_nt does not depend on the tile
 *     position), so streamed PCM == one-shot PCM bit for bit (tests/test_gpu_stream.py).
 * The Python reference streams 300-frame chunks with 25 frames of left context and accepts the seam
 * (qwen_tts/core/tokenizer_12hz/modeling_qwen3_tts_tokenizer_v2.py:886-896); the C reference decodes once at the end. */
static int ctx_frames(const qtc_codec_t *k) {
    double f = 0.0, rate = 1.0;
    for (int s = 0; s < 2; s++) { rate *= k->upsampling_ratios[s]; f += 6.0 / rate; }          /* ConvNeXt dwconv k=7 (the k=s transposed conv has no halo) */
    f += 6.0 / rate;                                                                             /* vocoder pre-conv k=7 */
    for (int b = 0; b < 4; b++) { f += 1.0 / rate; rate *= k->upsample_rates[b]; f += (6.0 + 18.0 + 54.0) / rate; }   /* tconv k=2r: one input back; ResUnits dil 1,3,9 */
    f += 6.0 / rate;                                                                             /* final conv k=7 */
    return (int)ceil(f) + 1;
}

static void stream_free(CodecDev *d) {
    cudaFree(d->s_tf); cudaFree(d->s_codes);
    for (int l = 0; l < QTC_MAX_CODEC_LAYERS; l++) { cudaFree(d->s_k[l]); cudaFree(d->s_v[l]); d->s_k[l] = d->s_v[l] = NULL; }
    d->s_tf = NULL; d->s_codes = NULL; d->s_cap = 0; d->s_T = 0;
}

/* start a decode of up to max_frames frames (one-shot decodes are a session of one chunk) */
int qtc_codec_stream_begin(qtc_ctx *c, int max_frames) {
    c->last_pcm_dev = NULL; c->last_pcm_n = 0;        /* the workspace is about to be reused */
    CodecDev *d = c->codec;
    if (!d || max_frames <= 0) { qtc_set_error("codec stream: no codec / bad frame count"); return -1; }
    const qtc_codec_t *k = &d->h;
    const int Hc = k->hidden, hd = Hc / k->heads, kvd = k->kv_heads * hd;
    if (k->sliding_window > 128 || hd > 256 || hd % 4) { qtc_set_error("codec: window %d / head_dim %d unsupported", k->sliding_window, hd); return -1; }
    if (k->codebook_dim != k->latent / 2) { qtc_set_error("codec: codebook_dim must equal latent/2 (SURVEY 7.0)"); return -1; }
    if (max_frames > d->s_cap) {
        cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->cstream);
        stream_free(d);
        const int cap = max_frames + 64;
        c->dev_mallocs += 2 + 2 * k->layers;
        bool ok = cudaMalloc(&d->s_tf, (size_t)cap * k->latent * 4) == cudaSuccess && cudaMalloc(&d->s_codes, (size_t)cap * k->num_quantizers * 4) == cudaSuccess;
        for (int l = 0; l < k->layers && ok; l++)
            ok = cudaMalloc(&d->s_k[l], (size_t)cap * kvd * 4) == cudaSuccess && cudaMalloc(&d->s_v[l], (size_t)cap * kvd * 4) == cudaSuccess;
        if (!ok) { stream_free(d); qtc_set_error("codec stream: state for %d frames failed", cap); return -1; }
        d->s_cap = cap;
    }
    if (max_frames > d->rope_T) {          /* codec RoPE: theta 10000 hard-coded (c/qwen_tts_codec.c:309), host libm */
        cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->cstream);
        cudaFree(d->rope_cs); cudaFree(d->rope_sn);
        d->rope_cs = d->rope_sn = NULL; d->rope_T = 0;
        const int n = max_frames + 256;
        std::vector<float> cs((size_t)n * hd), sn((size_t)n * hd);
        for (int t = 0; t < n; t++)
            for (int i = 0; i < hd / 2; i++) {
                const float freq = 1.0f / powf(10000.0f, (float)(2 * i) / (float)hd), ang = (float)t * freq;
                cs[(size_t)t * hd + i] = cs[(size_t)t * hd + i + hd / 2] = cosf(ang);
                sn[(size_t)t * hd + i] = sn[(size_t)t * hd + i + hd / 2] = sinf(ang);
            }
        c->dev_mallocs += 2;
        if (cudaMalloc(&d->rope_cs, cs.size() * 4) != cudaSuccess || cudaMalloc(&d->rope_sn, sn.size() * 4) != cudaSuccess) { qtc_set_error("codec: rope tables"); return -1; }
        cudaMemcpy(d->rope_cs, cs.data(), cs.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d->rope_sn, sn.data(), sn.size() * 4, cudaMemcpyHostToDevice);
        d->rope_T = n;
    }
    d->s_T = 0; d->s_max = max_frames;
    d->s_ctx = ctx_frames(k);
    return 0;
}

/* workspace for chunks of at most `chunk` new frames; call before enqueueing chunks that overlap other work */
int qtc_codec_stream_reserve(qtc_ctx *c, int chunk) {
    CodecDev *d = c->codec;
    const size_t half_bytes = (codec_half_bytes(&d->h, chunk + d->s_ctx + 2) + 255) & ~(size_t)255;
    if (2 * half_bytes > d->ws_bytes) {
        cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->cstream);
        cudaFree(d->ws); d->ws = NULL; d->ws_bytes = 0;
        QTC_COUNT_MALLOC(c);
        if (cudaMalloc(&d->ws, 2 * half_bytes) != cudaSuccess) { qtc_set_error("codec: workspace of %.1f MB failed", 2 * half_bytes / 1e6); return -1; }
        d->ws_bytes = 2 * half_bytes;
    }
    d->ws_half = d->ws_bytes / 2;
    return 0;
}

/* decode the next n frames.  codes_dev [n, nq] (device) are the codes of frames [T, T+n).  All work is enqueued on c->cstream.
 * *pcm_dev receives the device pointer of the n*1920 new samples (inside the workspace: valid until the next push). */
int qtc_codec_stream_push(qtc_ctx *c, const int *codes_dev, int n, float **pcm_dev) {
    CodecDev *d = c->codec;
    if (!d || !codes_dev || n <= 0) return -1;
    const qtc_codec_t *k = &d->h;
    const int T0 = d->s_T, T1 = T0 + n;
    if (T1 > d->s_max || T1 > d->s_cap) { qtc_set_error("codec stream: %d frames pushed, session opened for %d", T1, d->s_max); return -1; }
    const int lat = k->latent, Hc = k->hidden, I = k->intermediate, nh = k->heads, nkv = k->kv_heads, hd = Hc / nh;
    const int qd = nh * hd, kvd = nkv * hd, vq = k->codebook_dim / 2, half = lat / 2, nq = k->num_quantizers;
    if (qtc_codec_stream_reserve(c, n)) return -1;
    const size_t half_bytes = d->ws_half;
    cudaStream_t st = c->cstream;
#define CK(x) do { if ((x) != 0) return -1; } while (0)
    if (cudaMemcpyAsync(d->s_codes + (size_t)T0 * nq, codes_dev, (size_t)n * nq * 4, cudaMemcpyDeviceToDevice, st) != cudaSuccess) { qtc_set_error("codec stream: codes copy"); return -1; }
    Arena ar = {(char *)d->ws, 0, half_bytes};
    /* ---- RVQ + pre-conv (k = 3: two frames of left context, recomputed from the codes) + transformer on the n new frames ---- */
    const int r0 = T0 >= 2 ? T0 - 2 : 0, m = T1 - r0, skip = T0 - r0;
    float *sem = ar.take((size_t)m * vq), *aco = ar.take((size_t)m * vq);
    float *h0 = ar.take((size_t)half * m), *pre = ar.take((size_t)lat * m), *seq = ar.take((size_t)m * lat);
    float *x = ar.take((size_t)n * Hc), *xn = ar.take((size_t)n * Hc), *q = ar.take((size_t)n * qd);
    float *att = ar.take((size_t)n * qd), *g = ar.take((size_t)n * I), *u = ar.take((size_t)n * I);
    if (ar.off > half_bytes) { qtc_set_error("codec: workspace underestimated (transformer)"); return -1; }
    {
        RvqArgs ra;
        for (int qi = 0; qi < nq; qi++) ra.cb[qi] = d->cb[qi];
        ra.codes = d->s_codes + (size_t)r0 * nq; ra.sem = sem; ra.aco = aco; ra.T = m; ra.nq = nq; ra.vq = vq; ra.cb_size = k->codebook_size;
        k_rvq_gather<<<m, 128, 0, st>>>(ra);
        c->launches++;
    }
    CK(qtc_sgemm(c, h0, sem, d->sem_proj, m, half, vq, NULL, NULL, NULL, 0, 1, m));           /* -> [half, m] */
    CK(qtc_sgemm(c, h0, aco, d->aco_proj, m, half, vq, NULL, NULL, h0, 0, 1, m));
    CK(qtc_conv(c, pre, h0, d->pre_w, d->pre_b, NULL, NULL, NULL, k->codebook_dim, lat, 3, 1, m, 1));
    CK(transpose(c, seq, pre, lat, m));                                                         /* [m, lat] */
    const float scale = 1.0f / sqrtf((float)hd);
    const bool tcx = d->tcx_ok && c->codec_tc;
    const int w0 = T0 > d->s_ctx ? T0 - d->s_ctx : 0, W = T1 - w0;
    int L = W, side = 1;                 /* the transformer stage uses half 0; `side` = half the next stage writes */
    auto half_arena = [&](int sd) { Arena a = {(char *)d->ws + (size_t)sd * half_bytes, 0, half_bytes}; return a; };
    uint16_t *uph = NULL, *upl = NULL;   /* tensor-core path: planes [4W, lat] the upsample stages leave for the vocoder pre-conv */
    float *cur = NULL;
    if (tcx) {
        /* ---- transformer + upsample on the tensor cores: every projection is a k = 1 convolution over time-major rows; RMSNorm,
         * attention and SwiGLU emit the bf16 planes the next GEMM reads; LayerScale and residual ride in the GEMM's epilogue ---- */
        const size_t pmax = (size_t)n * (I > Hc ? I : Hc) / 2 + 8;
        uint16_t *Ph = (uint16_t *)ar.take(pmax), *Pl = (uint16_t *)ar.take(pmax);
        if (ar.off > half_bytes) { qtc_set_error("codec: workspace underestimated (transformer, tc)"); return -1; }
        CK(qtc_rows_to_planes(c, Ph, Pl, seq + (size_t)skip * lat, (size_t)n * lat));
        CK(qtc_conv_tc(c, &d->tc_in, Ph, Pl, n, 1, d->in_b, NULL, x, NULL, NULL, NULL, NULL));
        for (int l = 0; l < k->layers; l++) {
            const CodecLayerDev *w = &d->layer[l];
            float *kk = d->s_k[l] + (size_t)T0 * kvd, *v = d->s_v[l] + (size_t)T0 * kvd;
            k_row_norm<<<n, 256, 0, st>>>(NULL, x, w->in_norm, NULL, Hc, k->rms_eps, 0, Ph, Pl);
            c->launches++;
            CK(qtc_conv_tc(c, &d->tcl[l].q, Ph, Pl, n, 1, NULL, NULL, q, NULL, NULL, NULL, NULL));
            CK(qtc_conv_tc(c, &d->tcl[l].k, Ph, Pl, n, 1, NULL, NULL, kk, NULL, NULL, NULL, NULL));
            CK(qtc_conv_tc(c, &d->tcl[l].v, Ph, Pl, n, 1, NULL, NULL, v, NULL, NULL, NULL, NULL));
            k_rope_rows<<<n, 256, 0, st>>>(q, kk, d->rope_cs + (size_t)T0 * hd, d->rope_sn + (size_t)T0 * hd, n, nh, nkv, hd);
            c->launches++;
            k_window_attn<<<(n * nh + 3) / 4, 128, 0, st>>>(NULL, q, d->s_k[l], d->s_v[l], n, nh, nkv, hd, k->sliding_window, scale, T0, Ph, Pl);
            c->launches++;
            CK(qtc_conv_tc(c, &d->tcl[l].o, Ph, Pl, n, 1, NULL, x, x, NULL, NULL, NULL, NULL, w->attn_scale, 0));
            k_row_norm<<<n, 256, 0, st>>>(NULL, x, w->post_norm, NULL, Hc, k->rms_eps, 0, Ph, Pl);
            c->launches++;
            CK(qtc_conv_tc(c, &d->tcl[l].gate, Ph, Pl, n, 1, NULL, NULL, g, NULL, NULL, NULL, NULL));
            CK(qtc_conv_tc(c, &d->tcl[l].up, Ph, Pl, n, 1, NULL, NULL, u, NULL, NULL, NULL, NULL));
            k_swiglu_mul<<<(unsigned)(((size_t)n * I + 255) / 256), 256, 0, st>>>(g, u, (size_t)n * I, Ph, Pl);
            c->launches++;
            CK(qtc_conv_tc(c, &d->tcl[l].down, Ph, Pl, n, 1, NULL, x, x, NULL, NULL, NULL, NULL, w->mlp_scale, 0));
        }
        if (d->final_norm) { k_row_norm<<<n, 256, 0, st>>>(NULL, x, d->final_norm, NULL, Hc, k->rms_eps, 0, Ph, Pl); c->launches++; }
        else CK(qtc_rows_to_planes(c, Ph, Pl, x, (size_t)n * Hc));
        CK(qtc_conv_tc(c, &d->tc_out, Ph, Pl, n, 1, d->out_b, NULL, d->s_tf + (size_t)T0 * lat, NULL, NULL, NULL, NULL));   /* time-major rows [T0, T1) */
        /* upsample: transposed conv (k = s: one tap per phase) + ConvNeXt, all time-major */
        Arena a1 = half_arena(side);
        uint16_t *ih = (uint16_t *)a1.take((size_t)W * lat / 2 + 8), *il = (uint16_t *)a1.take((size_t)W * lat / 2 + 8);
        CK(qtc_rows_to_planes(c, ih, il, d->s_tf + (size_t)w0 * lat, (size_t)W * lat));
        for (int s = 0; s < 2; s++) {
            const int f = k->upsampling_ratios[s], Lo = L * f;
            const ConvNextDev *b = &d->cnx[s];
            Arena a2 = s == 0 ? a1 : half_arena(side);
            float *U = a2.take((size_t)Lo * lat), *cv = a2.take((size_t)Lo * lat);
            uint16_t *p1h = (uint16_t *)a2.take((size_t)Lo * lat / 2 + 8), *p1l = (uint16_t *)a2.take((size_t)Lo * lat / 2 + 8);
            uint16_t *p2h = (uint16_t *)a2.take((size_t)Lo * 2 * lat + 8), *p2l = (uint16_t *)a2.take((size_t)Lo * 2 * lat + 8);
            uint16_t *oh = (uint16_t *)a2.take((size_t)Lo * lat / 2 + 8), *ol = (uint16_t *)a2.take((size_t)Lo * lat / 2 + 8);
            if (a2.off > half_bytes) { qtc_set_error("codec: workspace underestimated (upsample, tc)"); return -1; }
            CK(qtc_conv_tc(c, &d->tc_up[s], ih, il, L, 1, d->up_tconv[s].b, NULL, U, NULL, NULL, NULL, NULL));
            CK(qtc_dwconv_tm(c, cv, U, b->dw_w, b->dw_b, lat, 7, Lo));
            k_row_norm<<<Lo, 256, 0, st>>>(NULL, cv, b->norm_w, b->norm_b, lat, 1e-6f, 1, p1h, p1l);
            c->launches++;
            CK(qtc_conv_tc(c, &d->tc_pw1[s], p1h, p1l, Lo, 1, b->pw1_b, NULL, NULL, p2h, p2l, NULL, NULL, NULL, 1));
            CK(qtc_conv_tc(c, &d->tc_pw2[s], p2h, p2l, Lo, 1, b->pw2_b, U, NULL, oh, ol, NULL, NULL, b->gamma, 0));
            ih = oh; il = ol; L = Lo; side ^= 1;
        }
        uph = ih; upl = il;
    } else {
    CK(qtc_sgemm(c, x, seq + (size_t)skip * lat, d->in_w, n, Hc, lat, d->in_b, NULL, NULL, 0, 0, Hc));
    for (int l = 0; l < k->layers; l++) {
        const CodecLayerDev *w = &d->layer[l];
        float *kk = d->s_k[l] + (size_t)T0 * kvd, *v = d->s_v[l] + (size_t)T0 * kvd;          /* rows of this chunk in the persistent caches */
        CK(row_norm(c, xn, x, w->in_norm, NULL, n, Hc, k->rms_eps, 0));
        CK(qtc_sgemm(c, q, xn, w->wq, n, qd, Hc, NULL, NULL, NULL, 0, 0, qd));
        CK(qtc_sgemm(c, kk, xn, w->wk, n, kvd, Hc, NULL, NULL, NULL, 0, 0, kvd));
        CK(qtc_sgemm(c, v, xn, w->wv, n, kvd, Hc, NULL, NULL, NULL, 0, 0, kvd));
        k_rope_rows<<<n, 256, 0, st>>>(q, kk, d->rope_cs + (size_t)T0 * hd, d->rope_sn + (size_t)T0 * hd, n, nh, nkv, hd);
        c->launches++;
        k_window_attn<<<(n * nh + 3) / 4, 128, 0, st>>>(att, q, d->s_k[l], d->s_v[l], n, nh, nkv, hd, k->sliding_window, scale, T0);
        c->launches++;
        CK(qtc_sgemm(c, x, att, w->wo, n, Hc, qd, NULL, w->attn_scale, x, 0, 0, Hc));
        CK(row_norm(c, xn, x, w->post_norm, NULL, n, Hc, k->rms_eps, 0));
        CK(qtc_sgemm(c, g, xn, w->gate, n, I, Hc, NULL, NULL, NULL, 0, 0, I));
        CK(qtc_sgemm(c, u, xn, w->up, n, I, Hc, NULL, NULL, NULL, 0, 0, I));
        k_swiglu_mul<<<(unsigned)(((size_t)n * I + 255) / 256), 256, 0, st>>>(g, u, (size_t)n * I);
        c->launches++;
        CK(qtc_sgemm(c, x, g, w->down, n, Hc, I, NULL, w->mlp_scale, x, 0, 0, Hc));
    }
    if (d->final_norm) CK(row_norm(c, x, x, d->final_norm, NULL, n, Hc, k->rms_eps, 0));
    CK(qtc_sgemm(c, d->s_tf + (size_t)T0 * lat, x, d->out_w, n, lat, Hc, d->out_b, NULL, NULL, 0, 0, lat));   /* time-major rows [T0, T1) */

    /* ---- conv stack on the window [w0, T1): 2 x (transposed conv + ConvNeXt), vocoder ---- */
    cur = pre;                                                                                  /* half 0, free again: [lat, W] */
    CK(transpose(c, cur, d->s_tf + (size_t)w0 * lat, W, lat));
    for (int s = 0; s < 2; s++) {
        const int f = k->upsampling_ratios[s], Lo = L * f;
        Arena a2 = half_arena(side);
        float *up = a2.take((size_t)lat * Lo), *cv = a2.take((size_t)lat * Lo), *ld = a2.take((size_t)Lo * lat), *pw = a2.take((size_t)Lo * 4 * lat);
        if (a2.off > half_bytes) { qtc_set_error("codec: workspace underestimated (upsample)"); return -1; }
        CK(tconv(c, up, cur, d->up_tconv[s], NULL, NULL, L));
        const ConvNextDev *b = &d->cnx[s];
        k_dwconv<<<dim3((Lo + 255) / 256, lat), 256, 0, st>>>(cv, up, b->dw_w, b->dw_b, lat, 7, Lo, 1);
        c->launches++;
        CK(transpose(c, ld, cv, lat, Lo));                                                      /* [Lo, lat] */
        CK(row_norm(c, ld, ld, b->norm_w, b->norm_b, Lo, lat, 1e-6f, 1));
        CK(qtc_sgemm(c, pw, ld, b->pw1_w, Lo, 4 * lat, lat, b->pw1_b, NULL, NULL, 1, 0, 4 * lat));
        CK(qtc_sgemm(c, up, pw, b->pw2_w, Lo, lat, 4 * lat, b->pw2_b, b->gamma, up, 0, 1, Lo)); /* h[c,t] += gamma*(..) */
        cur = up; L = Lo; side ^= 1;
    }
    }
    int dim = k->decoder_dim;
    static const int dil[3] = {1, 3, 9};
    if (d->tc_ok && c->codec_tc) {
        /* ---- vocoder on the tensor cores: time-major, two bf16 planes per activation (qtc_conv_tc.cu) ---- */
        Arena a3 = half_arena(side);
        uint16_t *p0h = uph, *p0l = upl;
        if (!uph) { p0h = (uint16_t *)a3.take((size_t)L * lat / 2 + 8); p0l = (uint16_t *)a3.take((size_t)L * lat / 2 + 8); }
        uint16_t *sh = (uint16_t *)a3.take((size_t)L * dim / 2 + 8), *sl = (uint16_t *)a3.take((size_t)L * dim / 2 + 8);
        if (a3.off > half_bytes) { qtc_set_error("codec: workspace underestimated (vocoder pre-conv, tc)"); return -1; }
        if (!uph) CK(qtc_cf_to_planes(c, p0h, p0l, cur, lat, L, NULL, NULL));
        CK(qtc_conv_tc(c, &d->tc_pre, p0h, p0l, L, 1, d->voc_pre_b, NULL, NULL, sh, sl, d->block[0].alpha, d->block[0].beta));
        side ^= 1;
        for (int b = 0; b < 4; b++) {
            const int r = k->upsample_rates[b], od = dim / 2, Lo = L * r;
            Arena a4 = half_arena(side);
            float *X = a4.take((size_t)Lo * od);
            uint16_t *s2h = (uint16_t *)a4.take((size_t)Lo * od / 2 + 8), *s2l = (uint16_t *)a4.take((size_t)Lo * od / 2 + 8);
            uint16_t *th = (uint16_t *)a4.take((size_t)Lo * od / 2 + 8), *tl = (uint16_t *)a4.take((size_t)Lo * od / 2 + 8);
            if (a4.off > half_bytes) { qtc_set_error("codec: workspace underestimated (vocoder, tc)"); return -1; }
            const ResUnitDev *r0 = &d->block[b].ru[0];
            CK(qtc_conv_tc(c, &d->tc_t[b], sh, sl, L, 1, d->block[b].tc.b, NULL, X, s2h, s2l, r0->a1, r0->b1));
            for (int uu = 0; uu < 3; uu++) {
                const ResUnitDev *ru = &d->block[b].ru[uu];
                CK(qtc_conv_tc(c, &d->tc_c1[b][uu], s2h, s2l, Lo, dil[uu], ru->c1, NULL, NULL, th, tl, ru->a2, ru->b2));
                /* conv1 + residual; the planes it leaves carry the NEXT consumer's SnakeBeta: next ResUnit, next block, or the final conv */
                const float *na = uu < 2 ? d->block[b].ru[uu + 1].a1 : (b < 3 ? d->block[b + 1].alpha : d->fin_alpha);
                const float *nb = uu < 2 ? d->block[b].ru[uu + 1].b1 : (b < 3 ? d->block[b + 1].beta : d->fin_beta);
                CK(qtc_conv_tc(c, &d->tc_c2[b][uu], th, tl, Lo, 1, ru->c2, X, uu < 2 ? X : NULL, s2h, s2l, na, nb));
            }
            sh = s2h; sl = s2l; L = Lo; dim = od; side ^= 1;
        }
        Arena a5 = half_arena(side);
        float *wav = a5.take((size_t)L);
        CK(qtc_mono_from_planes(c, wav, sh, sl, d->fin_w, d->fin_b, dim, 7, L));
        d->s_T = T1;
        if (pcm_dev) *pcm_dev = wav + (size_t)(T0 - w0) * (L / W);
        return n * (L / W);
    }
    Arena a3 = half_arena(side);
    float *voc = a3.take((size_t)dim * L);
    if (a3.off > half_bytes) { qtc_set_error("codec: workspace underestimated (vocoder pre-conv)"); return -1; }
    CK(qtc_conv(c, voc, cur, d->voc_pre_w, d->voc_pre_b, NULL, NULL, NULL, lat, dim, 7, 1, L, 1));
    side ^= 1;
    for (int b = 0; b < 4; b++) {
        const int r = k->upsample_rates[b], od = dim / 2, Lo = L * r;
        Arena a4 = half_arena(side);
        float *hh = a4.take((size_t)od * Lo), *t1 = a4.take((size_t)od * Lo);
        if (a4.off > half_bytes) { qtc_set_error("codec: workspace underestimated (vocoder)"); return -1; }
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
    k_conv_to_mono<<<(L + 255) / 256, 256, 0, st>>>(wav, voc, d->fin_w, d->fin_b, d->fin_alpha, d->fin_beta, dim, 7, L, 1);
    c->launches++;
#undef CK
    d->s_T = T1;
    if (pcm_dev) *pcm_dev = wav + (size_t)(T0 - w0) * (L / W);
    return n * (L / W);
}

/* codes_dev [T, nq] on device.  Returns malloc'd host PCM (caller frees); *dev_out (optional) receives the
 * device pointer of the PCM inside the workspace (valid until the next decode).  A session of one chunk. */
float *qtc_codec_decode_dev(qtc_ctx *c, const int *codes_dev, int T, int *out_samples, float **dev_out) {
    CodecDev *d = c->codec;
    if (out_samples) *out_samples = 0;
    if (!d || !codes_dev || T <= 0 || !out_samples) return NULL;
    float *wav = NULL;
    if (qtc_codec_stream_begin(c, T)) return NULL;
    const int L = qtc_codec_stream_push(c, codes_dev, T, &wav);
    if (L <= 0) return NULL;
    cudaEventRecord(c->ev3, c->cstream);             /* end of device work; what follows is the D2H of the PCM */
    float *host = (float *)malloc((size_t)L * 4);
    if (!host) return NULL;
    if (cudaMemcpyAsync(host, wav, (size_t)L * 4, cudaMemcpyDeviceToHost, c->cstream) != cudaSuccess ||
        cudaStreamSynchronize(c->cstream) != cudaSuccess) {
        qtc_set_error("codec: device error %s", cudaGetErrorString(cudaGetLastError()));
        free(host);
        return NULL;
    }
    *out_samples = L;
    if (dev_out) *dev_out = wav;
    c->last_pcm_dev = wav; c->last_pcm_n = L;
    return host;
}

/* c/qwen_tts_audio.c:70-82 on the device: clamp to [-1, 1], s * 32767.0f (one rounding), truncate toward zero */
__global__ void k_pcm_to_s16(int16_t *out, const float *in, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = in[i];
    if (s > 1.0f) s = 1.0f;
    if (s < -1.0f) s = -1.0f;
    out[i] = (int16_t)__float2int_rz(__fmul_rn(s, 32767.0f));
}
extern "C" int qwen_tts_cuda_last_pcm_s16(qtc_ctx_t *c, int16_t *out, int max_samples) {
    if (!c || !out || max_samples <= 0) { qtc_set_error("last_pcm_s16: null or empty argument"); return -1; }
    if (!c->last_pcm_dev || c->last_pcm_n <= 0) { qtc_set_error("last_pcm_s16: no one-shot decode to convert (a streamed decode keeps no device copy of the utterance)"); return -1; }
    const int n = c->last_pcm_n < max_samples ? c->last_pcm_n : max_samples;
    if ((size_t)n > c->s16_cap) {
        cudaFree(c->s16_dev); c->s16_dev = NULL; c->s16_cap = 0;
        QTC_COUNT_MALLOC(c);
        if (cudaMalloc(&c->s16_dev, (size_t)n * 2 + 4096) != cudaSuccess) { qtc_set_error("last_pcm_s16: device buffer"); return -1; }
        c->s16_cap = (size_t)n + 2048;
    }
    k_pcm_to_s16<<<(n + 255) / 256, 256, 0, c->cstream>>>(c->s16_dev, c->last_pcm_dev, n);
    c->launches++;
    QTC_CUDA(cudaMemcpyAsync(out, c->s16_dev, (size_t)n * 2, cudaMemcpyDeviceToHost, c->cstream));
    QTC_CUDA(cudaStreamSynchronize(c->cstream));
    return n;
}

/* ---- streaming through the C ABI (host buffers): SURVEY.md 8f rank 3 ------------------------------------------------ */
extern "C" int qwen_tts_cuda_codec_stream_begin(qtc_ctx_t *c, int max_frames) {
    if (!c || !c->codec) { qtc_set_error("codec_stream_begin: context has no codec"); return -1; }
    return qtc_codec_stream_begin(c, max_frames);
}
/* codes [n, nq] (host): the next n frames; pcm_out (host, n * 1920 floats) receives exactly their samples.  Returns the
 * sample count or -1. */
extern "C" int qwen_tts_cuda_codec_stream_push(qtc_ctx_t *c, const int *codes, int n, float *pcm_out) {
    if (!c || !c->codec) { qtc_set_error("codec_stream_push: context has no codec"); return -1; }
    if (!codes || n <= 0 || !pcm_out) { qtc_set_error("codec_stream_push: null or empty argument"); return -1; }
    CodecDev *d = c->codec;
    const int nq = d->h.num_quantizers;
    if (n > d->codes_cap) {
        cudaStreamSynchronize(c->cstream);
        cudaFree(d->codes_dev);
        d->codes_dev = NULL; d->codes_cap = 0;
        QTC_COUNT_MALLOC(c);
        if (cudaMalloc(&d->codes_dev, (size_t)(n + 64) * nq * 4) != cudaSuccess) { qtc_set_error("codec: codes buffer"); return -1; }
        d->codes_cap = n + 64;
    }
    if (cudaMemcpyAsync(d->codes_dev, codes, (size_t)n * nq * 4, cudaMemcpyHostToDevice, c->cstream) != cudaSuccess) return -1;
    float *wav = NULL;
    const int L = qtc_codec_stream_push(c, d->codes_dev, n, &wav);
    if (L <= 0) return -1;
    if (cudaMemcpyAsync(pcm_out, wav, (size_t)L * 4, cudaMemcpyDeviceToHost, c->cstream) != cudaSuccess || cudaStreamSynchronize(c->cstream) != cudaSuccess) {
        qtc_set_error("codec stream: device error %s", cudaGetErrorString(cudaGetLastError()));
        return -1;
    }
    return L;
}

extern "C" float *qwen_tts_cuda_codec_decode(qtc_ctx_t *c, const int *codes, int T, int *out_samples) {
    if (out_samples) *out_samples = 0;
    if (!c || !c->codec) { qtc_set_error("codec_decode: context has no codec"); return NULL; }
    if (!codes || T <= 0 || !out_samples) return NULL;          /* c/qwen_tts_codec.c:583-587 */
    CodecDev *d = c->codec;
    const int nq = d->h.num_quantizers;
    if (T > d->codes_cap) {
        cudaStreamSynchronize(c->cstream);
        cudaFree(d->codes_dev);
        d->codes_dev = NULL; d->codes_cap = 0;
        QTC_COUNT_MALLOC(c);
        if (cudaMalloc(&d->codes_dev, (size_t)(T + 64) * nq * 4) != cudaSuccess) { qtc_set_error("codec: codes buffer"); return NULL; }
        d->codes_cap = T + 64;
    }
    cudaEventRecord(c->ev0, c->cstream);
    if (cudaMemcpyAsync(d->codes_dev, codes, (size_t)T * nq * 4, cudaMemcpyHostToDevice, c->cstream) != cudaSuccess) return NULL;
    float *pcm = qtc_codec_decode_dev(c, d->codes_dev, T, out_samples, NULL);
    cudaEventRecord(c->ev1, c->cstream);
    cudaEventSynchronize(c->ev1);
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->perf.codec_ms = ms;
    if (pcm) { float dms = 0; cudaEventElapsedTime(&dms, c->ev0, c->ev3); c->perf.device_ms = dms; }   /* codes H2D .. last codec kernel */
    return pcm;
}
