/* qtts_oracle_internal.h - shared structs of the oracle (TEST INFRASTRUCTURE ONLY). */
#ifndef QTTS_ORACLE_INTERNAL_H
#define QTTS_ORACLE_INTERNAL_H
#include "qtts_oracle.h"

/* scratch + KV of one decoder stack: tk_* / st_* buffers of qwen_tts_ctx_t, c/qwen_tts.h:373-413 */
typedef struct {
    const qtc_lm_t *lm;
    float *k, *v;        /* [layers, kv_max, kv_heads*head_dim] */
    int kv_max;
    float *x, *xn, *q, *kk, *vv, *att, *g, *u, *sc;
} orc_lm_state_t;

struct orc_state {
    const qtc_model_desc_t *m;
    qtc_gen_params_t p;
    orc_lm_state_t tk, st;
    int tk_len;
    float *st_logits, *st_embed, *st_proj;
};

void orc_lm_alloc(orc_lm_state_t *st, const qtc_lm_t *lm, int kv_max);
void orc_lm_free(orc_lm_state_t *st);
void orc_lm_step(orc_lm_state_t *st, int pos);
void qwen_tts_cuda_default_params_host(qtc_gen_params_t *p);
#endif
