/*
 * qwen_tts_cuda_adapter.c - reference-side glue: fills qtc_model_desc_t from a
 * loaded qwen_tts_ctx_t.
 *
 * This is the file a maintainer of muonium-ai/Qwen3-TTS-C adds next to
 * c/qwen_tts_cuda.h.  It is compiled against the
 * reference's own c/qwen_tts.h (include path given on the command line, the
 * header is NOT copied into this repository), so it only builds where the
 * reference sources are present; oracle/Makefile builds it into oracle/_ref/.
 *
 * Field map: qwen_tts_ctx_t (c/qwen_tts.h:362-441) -> qtc_model_desc_t.
 */
#include "qwen_tts.h"          /* the reference's header, -I<reference>/c */
#include "qwen_tts_cuda.h"
#include <string.h>

static void fill_lm_layer(qtc_lm_layer_t *d, const uint16_t *wq, const uint16_t *wk, const uint16_t *wv,
                          const uint16_t *wo, const float *qn, const float *kn, const float *in_norm,
                          const float *post_norm, const uint16_t *gate, const uint16_t *up,
                          const uint16_t *down) {
    d->wq = wq; d->wk = wk; d->wv = wv; d->wo = wo;
    d->q_norm = qn; d->k_norm = kn; d->input_norm = in_norm; d->post_attn_norm = post_norm;
    d->gate = gate; d->up = up; d->down = down;
}

void qwen_tts_cuda_fill_desc(const qwen_tts_ctx_t *ctx, qtc_model_desc_t *d) {
    const qwen_tts_config_t *c = &ctx->config;
    memset(d, 0, sizeof(*d));

    d->talker.vocab = c->talker_vocab_size;   d->talker.hidden = c->talker_hidden;
    d->talker.intermediate = c->talker_intermediate; d->talker.layers = c->talker_layers;
    d->talker.heads = c->talker_heads;        d->talker.kv_heads = c->talker_kv_heads;
    d->talker.head_dim = c->talker_head_dim;
    d->talker.rms_eps = c->talker_rms_norm_eps; d->talker.rope_theta = c->talker_rope_theta;
    for (int i = 0; i < c->talker_layers; i++) {
        const qwen_tts_talker_layer_t *l = &ctx->talker.layers[i];
        fill_lm_layer(&d->talker.layer[i], l->wq_bf16, l->wk_bf16, l->wv_bf16, l->wo_bf16,
                      l->q_norm_weight, l->k_norm_weight, l->input_norm, l->post_attn_norm,
                      l->gate_bf16, l->up_bf16, l->down_bf16);
    }
    d->talker.final_norm = ctx->talker.norm;
    d->text_hidden = c->talker_text_hidden; d->text_vocab = c->talker_text_vocab;
    d->num_code_groups = c->num_code_groups;
    memcpy(d->mrope_section, c->mrope_section, sizeof(d->mrope_section));
    d->codec_embedding = ctx->talker.codec_embedding_bf16;
    d->text_embedding = ctx->talker.text_embedding_bf16;
    d->text_fc1_w = ctx->talker.text_proj_fc1_bf16; d->text_fc1_b = ctx->talker.text_proj_fc1_bias;
    d->text_fc2_w = ctx->talker.text_proj_fc2_bf16; d->text_fc2_b = ctx->talker.text_proj_fc2_bias;
    d->codec_head = ctx->talker.codec_head_bf16;

    /* the sub-talker reuses the talker's eps and theta: c/qwen_tts_talker.c:555,619 */
    d->sub.vocab = c->subtalker_vocab_size;   d->sub.hidden = c->subtalker_hidden;
    d->sub.intermediate = c->subtalker_intermediate; d->sub.layers = c->subtalker_layers;
    d->sub.heads = c->subtalker_heads;        d->sub.kv_heads = c->subtalker_kv_heads;
    d->sub.head_dim = c->subtalker_head_dim;
    d->sub.rms_eps = c->talker_rms_norm_eps;  d->sub.rope_theta = c->talker_rope_theta;
    for (int i = 0; i < c->subtalker_layers; i++) {
        const qwen_tts_subtalker_layer_t *l = &ctx->subtalker.layers[i];
        fill_lm_layer(&d->sub.layer[i], l->wq_bf16, l->wk_bf16, l->wv_bf16, l->wo_bf16,
                      l->q_norm_weight, l->k_norm_weight, l->input_norm, l->post_attn_norm,
                      l->gate_bf16, l->up_bf16, l->down_bf16);
    }
    d->sub.final_norm = ctx->subtalker.norm;
    for (int g = 0; g < c->num_code_groups - 1 && g < QTC_MAX_GROUPS - 1; g++) {
        d->sub_codec_embeddings[g] = ctx->subtalker.codec_embeddings_bf16[g];
        d->sub_lm_heads[g] = ctx->subtalker.lm_heads_bf16[g];
    }
    d->sub_input_proj_w = ctx->subtalker.input_proj_bf16;
    d->sub_input_proj_b = ctx->subtalker.input_proj_bias;

    qtc_codec_t *k = &d->codec;
    const qwen_tts_codec_decoder_t *s = &ctx->codec;
    k->num_quantizers = c->codec_num_quantizers; k->codebook_size = c->codec_codebook_size;
    k->codebook_dim = c->codec_codebook_dim;     k->hidden = c->codec_hidden;
    k->latent = c->codec_latent;                 k->layers = c->codec_layers;
    k->heads = c->codec_heads;                   k->kv_heads = c->codec_kv_heads;
    k->intermediate = c->codec_intermediate;     k->sliding_window = c->codec_sliding_window;
    k->decoder_dim = c->codec_decoder_dim;       k->rms_eps = c->codec_rms_norm_eps;
    memcpy(k->upsample_rates, c->codec_upsample_rates, sizeof(k->upsample_rates));
    memcpy(k->upsampling_ratios, c->codec_upsampling_ratios, sizeof(k->upsampling_ratios));
    k->codebooks[0] = s->rvq.semantic_codebooks[0].embeddings;
    for (int q = 1; q < c->codec_num_quantizers && q < QTC_MAX_QUANTIZERS; q++)
        k->codebooks[q] = s->rvq.acoustic_codebooks[q - 1].embeddings;
    k->semantic_out_proj = s->rvq.semantic_output_proj;
    k->acoustic_out_proj = s->rvq.acoustic_output_proj;
    k->pre_conv_w = s->pre_conv_weight;          k->pre_conv_b = s->pre_conv_bias;
    k->in_proj_w = s->transformer_input_proj_weight;   k->in_proj_b = s->transformer_input_proj_bias;
    k->out_proj_w = s->transformer_output_proj_weight; k->out_proj_b = s->transformer_output_proj_bias;
    for (int i = 0; i < c->codec_layers; i++) {
        const qwen_tts_codec_transformer_layer_t *l = &s->transformer_layers[i];
        qtc_codec_layer_t *o = &k->layer[i];
        o->input_norm = l->input_norm; o->post_attn_norm = l->post_attn_norm;
        o->attn_layer_scale = l->attn_layer_scale; o->mlp_layer_scale = l->mlp_layer_scale;
        o->wq = l->wq; o->wk = l->wk; o->wv = l->wv; o->wo = l->wo;
        o->gate = l->gate; o->up = l->up; o->down = l->down;
    }
    k->final_norm = s->transformer_norm;
    for (int i = 0; i < 2; i++) {
        k->up_tconv_w[i] = s->upsample_transconv_weight[i]; k->up_tconv_b[i] = s->upsample_transconv_bias[i];
        const qwen_tts_convnext_block_t *b = &s->upsample_convnext[i];
        qtc_convnext_t *o = &k->convnext[i];
        o->dwconv_w = b->dwconv_weight; o->dwconv_b = b->dwconv_bias;
        o->norm_w = b->norm_weight; o->norm_b = b->norm_bias;
        o->pw1_w = b->pwconv1_weight; o->pw1_b = b->pwconv1_bias;
        o->pw2_w = b->pwconv2_weight; o->pw2_b = b->pwconv2_bias; o->gamma = b->gamma;
    }
    k->voc_pre_w = s->vocoder_pre_conv_weight; k->voc_pre_b = s->vocoder_pre_conv_bias;
    for (int i = 0; i < 4; i++) {
        const qwen_tts_vocoder_block_t *b = &s->vocoder_blocks[i];
        qtc_vocoder_block_t *o = &k->block[i];
        o->act_alpha = b->act_alpha; o->act_beta = b->act_beta;
        o->tconv_w = b->transconv_weight; o->tconv_b = b->transconv_bias;
        for (int r = 0; r < 3; r++) {
            const qwen_tts_vocoder_resunit_t *u = &b->resunits[r];
            qtc_resunit_t *v = &o->res[r];
            v->act1_alpha = u->act1_alpha; v->act1_beta = u->act1_beta;
            v->conv1_w = u->conv1_weight;  v->conv1_b = u->conv1_bias;
            v->act2_alpha = u->act2_alpha; v->act2_beta = u->act2_beta;
            v->conv2_w = u->conv2_weight;  v->conv2_b = u->conv2_bias;
        }
    }
    k->final_alpha = s->vocoder_final_act_alpha; k->final_beta = s->vocoder_final_act_beta;
    k->final_conv_w = s->vocoder_final_conv_weight; k->final_conv_b = s->vocoder_final_conv_bias;

    d->codec_pad_id = c->codec_pad_id; d->codec_bos_id = c->codec_bos_id; d->codec_eos_id = c->codec_eos_id;
    d->codec_nothink_id = c->codec_nothink_id; d->codec_think_id = c->codec_think_id;
    d->codec_think_bos_id = c->codec_think_bos_id; d->codec_think_eos_id = c->codec_think_eos_id;
    d->tts_pad_id = QWEN_TTS_TOKEN_TTS_PAD; d->tts_bos_id = QWEN_TTS_TOKEN_TTS_BOS;
    d->tts_eos_id = QWEN_TTS_TOKEN_TTS_EOS;
    d->n_speakers = c->n_speakers < QTC_MAX_NAMES ? c->n_speakers : QTC_MAX_NAMES;
    for (int i = 0; i < d->n_speakers; i++) { d->speaker_names[i] = c->speaker_names[i]; d->speaker_ids[i] = c->speaker_ids[i]; }
    d->n_languages = c->n_languages < QTC_MAX_NAMES ? c->n_languages : QTC_MAX_NAMES;
    for (int i = 0; i < d->n_languages; i++) { d->language_names[i] = c->language_names[i]; d->language_ids[i] = c->language_ids[i]; }
}
