/*
 * ref_shim.c — batch driver over the UNMODIFIED reference (symbols ref_*).  TEST INFRASTRUCTURE ONLY.
 *
 * This file is ours; it is compiled together with the reference's own src/*.c (taken where they
 * lie under $(REF)/src, never copied into the repo) into oracle/_ref/libpoporon_ref.so by
 * oracle/Makefile.  It calls the reference's public API (poporon_create / poporon_encode /
 * poporon_decode) once per codeword, with one independent poporon_t per worker thread — the
 * "one poporon_t per host core" CPU baseline of BASELINE.md §3.
 *
 * It reaches into the reference's struct _poporon_erasure_t (src/internal/common.h:39-44) for one
 * purpose: to define EVERY slot of erasure_positions[] before a decode, because the reference
 * reads slots beyond erasure_count and malloc leaves them uninitialised.
 */
#include <poporon.h>

#include "internal/common.h"

#include "oracle.h"
#include "par.h"

#include <string.h>

const char *ref_kind(void)
{
    return "reference";
}

uint8_t ref_gf_mod(uint8_t m, uint16_t poly, uint16_t value)
{
    poporon_gf_t *gf = poporon_gf_create(m, poly);
    uint8_t r;
    if (!gf)
        return 0;
    r = poporon_gf_mod(gf, value);
    poporon_gf_destroy(gf);
    return r;
}

static poporon_t *make_rs(const orc_rs_params_t *p, poporon_erasure_t *er, uint16_t *syn)
{
    poporon_config_t *cfg = poporon_rs_config_create(p->symbol_size, p->gfpoly, p->fcr, p->prim, p->nroots, er, syn);
    poporon_t *h;
    if (!cfg)
        return NULL;
    h = poporon_create(cfg);
    poporon_config_destroy(cfg);
    return h;
}

int ref_rs_info(const orc_rs_params_t *p, uint16_t *genpoly_index, uint16_t *alpha_to, uint16_t *index_of,
                uint16_t *iprim)
{
    poporon_t *h = make_rs(p, NULL, NULL);
    poporon_rs_t *rs;
    if (!h)
        return -1;
    rs = h->ctx.rs.rs;
    if (genpoly_index)
        memcpy(genpoly_index, rs->generator_polynomial, (rs->num_roots + 1u) * sizeof(uint16_t));
    if (alpha_to)
        memcpy(alpha_to, rs->gf->log2exp, (rs->gf->field_size + 1u) * sizeof(uint16_t));
    if (index_of)
        memcpy(index_of, rs->gf->exp2log, (rs->gf->field_size + 1u) * sizeof(uint16_t));
    if (iprim)
        *iprim = h->ctx.rs.buffer->primitive_inverse;
    poporon_destroy(h);
    return 0;
}

/* struct _poporon_bch_t is private to src/bch.c; use the accessors. gen_poly is not exposed. */
int ref_bch_info(uint8_t m, uint16_t poly, uint8_t t, uint32_t out[4])
{
    poporon_bch_t *b = poporon_bch_create(m, poly, t);
    if (!b)
        return -1;
    out[0] = poporon_bch_get_codeword_length(b);
    out[1] = poporon_bch_get_data_length(b);
    out[2] = (uint32_t)(uint16_t)(out[0] - out[1]);
    out[3] = 0;
    poporon_bch_destroy(b);
    return 0;
}

/* ---------------------------------------------------------------- RS encode */
typedef struct {
    const orc_rs_params_t *p;
    const uint8_t *data;
    size_t data_stride, size;
    uint8_t *parity;
    size_t parity_stride;
    uint8_t *ok;
} enc_job_t;

static void enc_range(void *arg, size_t lo, size_t hi)
{
    enc_job_t *j = (enc_job_t *)arg;
    poporon_t *h = make_rs(j->p, NULL, NULL);
    for (size_t i = lo; i < hi; i++) {
        bool r = poporon_encode(h, (uint8_t *)(j->data + i * j->data_stride), j->size, j->parity + i * j->parity_stride);
        if (j->ok)
            j->ok[i] = r ? 1 : 0;
    }
    poporon_destroy(h);
}

int ref_rs_encode_batch(const orc_rs_params_t *p, const uint8_t *data, size_t data_stride, size_t size,
                        uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, int threads)
{
    enc_job_t job = {p, data, data_stride, size, parity, parity_stride, ok};
    poporon_t *probe = make_rs(p, NULL, NULL);
    if (!probe)
        return -1;
    poporon_destroy(probe);
    par_for(n, threads, enc_range, &job);
    return 0;
}

/* ---------------------------------------------------------------- RS decode */
typedef struct {
    const orc_rs_params_t *p;
    int mode;
    uint8_t *data;
    size_t data_stride, size;
    uint8_t *parity;
    size_t parity_stride;
    const uint32_t *pos;
    size_t pos_stride;
    const uint32_t *cnt;
    const uint16_t *syn;
    uint8_t *ok;
    uint32_t *corrected;
} dec_job_t;

static void dec_range(void *arg, size_t lo, size_t hi)
{
    dec_job_t *j = (dec_job_t *)arg;
    poporon_erasure_t *er = NULL;
    uint16_t synbuf[256];
    poporon_t *h;

    if (j->mode == ORC_MODE_ERASURE) {
        /* capacity large enough that all pos_stride slots exist; they are overwritten per codeword */
        er = poporon_erasure_create(j->p->nroots, (uint32_t)(j->pos_stride > 0 ? j->pos_stride : 1));
    }
    h = make_rs(j->p, er, j->mode == ORC_MODE_EXTSYN ? synbuf : NULL);
    for (size_t i = lo; i < hi; i++) {
        size_t nc = 0;
        bool r;
        if (er) {
            memcpy(er->erasure_positions, j->pos + i * j->pos_stride, j->pos_stride * sizeof(uint32_t));
            er->erasure_count = j->cnt[i];
        }
        if (j->mode == ORC_MODE_EXTSYN)
            memcpy(synbuf, j->syn + i * j->p->nroots, j->p->nroots * sizeof(uint16_t));
        r = poporon_decode(h, j->data + i * j->data_stride, j->size, j->parity + i * j->parity_stride, &nc);
        if (j->ok)
            j->ok[i] = r ? 1 : 0;
        if (j->corrected)
            j->corrected[i] = (uint32_t)nc;
    }
    poporon_destroy(h);
    poporon_erasure_destroy(er);
}

int ref_rs_decode_batch(const orc_rs_params_t *p, int mode, uint8_t *data, size_t data_stride, size_t size,
                        uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *pos, size_t pos_stride,
                        const uint32_t *cnt, const uint16_t *syn, uint8_t *ok, uint32_t *corrected, int threads)
{
    dec_job_t job = {p, mode, data, data_stride, size, parity, parity_stride, pos, pos_stride, cnt, syn, ok, corrected};
    poporon_t *probe = make_rs(p, NULL, NULL);
    if (!probe)
        return -1;
    poporon_destroy(probe);
    par_for(n, threads, dec_range, &job);
    return 0;
}

/* ---------------------------------------------------------------- BCH */
typedef struct {
    uint8_t m;
    uint16_t poly;
    uint8_t t;
    uint8_t *data;
    size_t data_stride, size;
    uint8_t *parity;
    size_t parity_stride;
    uint8_t *ok;
    uint32_t *corrected;
    int decode;
} bch_job_t;

static poporon_t *make_bch(uint8_t m, uint16_t poly, uint8_t t)
{
    poporon_config_t *cfg = poporon_bch_config_create(m, poly, t);
    poporon_t *h;
    if (!cfg)
        return NULL;
    h = poporon_create(cfg);
    poporon_config_destroy(cfg);
    return h;
}

static void bch_range(void *arg, size_t lo, size_t hi)
{
    bch_job_t *j = (bch_job_t *)arg;
    poporon_t *h = make_bch(j->m, j->poly, j->t);
    for (size_t i = lo; i < hi; i++) {
        bool r;
        if (j->decode) {
            size_t nc = 0;
            r = poporon_decode(h, j->data + i * j->data_stride, j->size, j->parity + i * j->parity_stride, &nc);
            if (j->corrected)
                j->corrected[i] = (uint32_t)nc;
        } else {
            r = poporon_encode(h, j->data + i * j->data_stride, j->size, j->parity + i * j->parity_stride);
        }
        if (j->ok)
            j->ok[i] = r ? 1 : 0;
    }
    poporon_destroy(h);
}

static int bch_defined(uint8_t m, uint16_t poly, uint8_t t)
{
    /* the reference's word arithmetic is undefined for n >= 32 (src/bch.c:39,396); never run it there */
    poporon_t *h;
    if (m > 5)
        return 0;
    h = make_bch(m, poly, t);
    if (!h)
        return 0;
    poporon_destroy(h);
    return 1;
}

int ref_bch_encode_batch(uint8_t m, uint16_t poly, uint8_t t, const uint8_t *data, size_t data_stride, size_t size,
                         uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, int threads)
{
    bch_job_t job = {m, poly, t, (uint8_t *)data, data_stride, size, parity, parity_stride, ok, NULL, 0};
    if (!bch_defined(m, poly, t))
        return -1;
    par_for(n, threads, bch_range, &job);
    return 0;
}

int ref_bch_decode_batch(uint8_t m, uint16_t poly, uint8_t t, uint8_t *data, size_t data_stride, size_t size,
                         const uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, uint32_t *corrected,
                         int threads)
{
    bch_job_t job = {m, poly, t, data, data_stride, size, (uint8_t *)parity, parity_stride, ok, corrected, 1};
    if (!bch_defined(m, poly, t))
        return -1;
    par_for(n, threads, bch_range, &job);
    return 0;
}
