/*
 * oracle_runner.c — batch runners over the plain-C restatement (symbols orc_*).
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 */
#include "oracle_internal.h"
#include "par.h"

#include <stdlib.h>
#include <string.h>

const char *orc_kind(void)
{
    return "port";
}

uint8_t orc_gf_mod(uint8_t m, uint16_t poly, uint16_t value)
{
    orc_field_t f;
    if (orc_field_init(&f, m, poly) != 0)
        return 0;
    return (uint8_t)(value % f.nn);
}

int orc_rs_info(const orc_rs_params_t *p, uint16_t *genpoly_index, uint16_t *alpha_to, uint16_t *index_of,
                uint16_t *iprim)
{
    orc_rs_t rs;
    if (orc_rs_init(&rs, p) != 0)
        return -1;
    if (genpoly_index)
        memcpy(genpoly_index, rs.genpoly, (rs.nroots + 1u) * sizeof(uint16_t));
    if (alpha_to)
        memcpy(alpha_to, rs.f.alpha_to, (rs.nn + 1u) * sizeof(uint16_t));
    if (index_of)
        memcpy(index_of, rs.f.index_of, (rs.nn + 1u) * sizeof(uint16_t));
    if (iprim)
        *iprim = rs.iprim;
    return 0;
}

int orc_bch_info(uint8_t m, uint16_t poly, uint8_t t, uint32_t out[4])
{
    orc_bch_t b;
    if (orc_bch_init(&b, m, poly, t) != 0)
        return -1;
    out[0] = b.n;
    out[1] = b.k;
    out[2] = b.parity_bits;
    out[3] = b.gen;
    return 0;
}

/* ---------------------------------------------------------------- RS encode */
typedef struct {
    const orc_rs_t *rs;
    const uint8_t *data;
    size_t data_stride, size;
    uint8_t *parity;
    size_t parity_stride;
    uint8_t *ok;
} enc_job_t;

static void enc_range(void *arg, size_t lo, size_t hi)
{
    enc_job_t *j = (enc_job_t *)arg;
    for (size_t i = lo; i < hi; i++) {
        orc_rs_encode(j->rs, j->data + i * j->data_stride, j->size, j->parity + i * j->parity_stride);
        if (j->ok)
            j->ok[i] = 1;
    }
}

int orc_rs_encode_batch(const orc_rs_params_t *p, const uint8_t *data, size_t data_stride, size_t size,
                        uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, int threads)
{
    orc_rs_t rs;
    enc_job_t job = {&rs, data, data_stride, size, parity, parity_stride, ok};
    if (orc_rs_init(&rs, p) != 0)
        return -1;
    par_for(n, threads, enc_range, &job);
    return 0;
}

/* ---------------------------------------------------------------- RS decode */
typedef struct {
    const orc_rs_t *rs;
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
    for (size_t i = lo; i < hi; i++) {
        size_t nc = 0;
        const uint32_t *pos = (j->mode == ORC_MODE_ERASURE) ? j->pos + i * j->pos_stride : NULL;
        uint32_t cnt = (j->mode == ORC_MODE_ERASURE) ? j->cnt[i] : 0;
        const uint16_t *syn = (j->mode == ORC_MODE_EXTSYN) ? j->syn + i * j->rs->nroots : NULL;
        int ok = orc_rs_decode(j->rs, j->mode, j->data + i * j->data_stride, j->size,
                               j->parity + i * j->parity_stride, cnt, pos, syn, &nc);
        if (j->ok)
            j->ok[i] = (uint8_t)ok;
        if (j->corrected)
            j->corrected[i] = (uint32_t)nc;
    }
}

int orc_rs_decode_batch(const orc_rs_params_t *p, int mode, uint8_t *data, size_t data_stride, size_t size,
                        uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *pos, size_t pos_stride,
                        const uint32_t *cnt, const uint16_t *syn, uint8_t *ok, uint32_t *corrected, int threads)
{
    orc_rs_t rs;
    dec_job_t job = {&rs, mode, data, data_stride, size, parity, parity_stride, pos, pos_stride, cnt, syn, ok, corrected};
    if (orc_rs_init(&rs, p) != 0)
        return -1;
    if (size == 0) { /* poporon_decode: false, *corrected_num untouched (we report 0) */
        for (size_t i = 0; i < n; i++) {
            if (ok)
                ok[i] = 0;
            if (corrected)
                corrected[i] = 0;
        }
        return 0;
    }
    par_for(n, threads, dec_range, &job);
    return 0;
}

/* ---------------------------------------------------------------- BCH */
typedef struct {
    const orc_bch_t *b;
    uint8_t *data;
    size_t data_stride, size;
    uint8_t *parity;
    size_t parity_stride;
    uint8_t *ok;
    uint32_t *corrected;
    int decode;
} bch_job_t;

static void bch_range(void *arg, size_t lo, size_t hi)
{
    bch_job_t *j = (bch_job_t *)arg;
    for (size_t i = lo; i < hi; i++) {
        int ok;
        if (j->decode) {
            size_t nc = 0;
            ok = orc_bch_decode_bytes(j->b, j->data + i * j->data_stride, j->size, j->parity + i * j->parity_stride,
                                      &nc);
            if (j->corrected)
                j->corrected[i] = (uint32_t)nc;
        } else {
            ok = orc_bch_encode_bytes(j->b, j->data + i * j->data_stride, j->size, j->parity + i * j->parity_stride);
        }
        if (j->ok)
            j->ok[i] = (uint8_t)ok;
    }
}

int orc_bch_encode_batch(uint8_t m, uint16_t poly, uint8_t t, const uint8_t *data, size_t data_stride, size_t size,
                         uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, int threads)
{
    orc_bch_t b;
    bch_job_t job = {&b, (uint8_t *)data, data_stride, size, parity, parity_stride, ok, NULL, 0};
    if (orc_bch_init(&b, m, poly, t) != 0)
        return -1;
    par_for(n, threads, bch_range, &job);
    return 0;
}

int orc_bch_decode_batch(uint8_t m, uint16_t poly, uint8_t t, uint8_t *data, size_t data_stride, size_t size,
                         const uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, uint32_t *corrected,
                         int threads)
{
    orc_bch_t b;
    bch_job_t job = {&b, data, data_stride, size, (uint8_t *)parity, parity_stride, ok, corrected, 1};
    if (orc_bch_init(&b, m, poly, t) != 0)
        return -1;
    par_for(n, threads, bch_range, &job);
    return 0;
}
