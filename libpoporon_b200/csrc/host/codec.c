/* codec.c — poporon_encode / poporon_decode and the batched entry points.
 *
 * Argument validation and result reporting follow reference src/encode.c:236-252 and
 * src/decode.c:431-487,542-612 (outcome table: SURVEY Appendix A.5); the codeword arithmetic is
 * executed by libpoporon_cuda.so through pipeline.c.  There is no CPU path. */
#include "pp_internal.h"

static bool rs_handle(const poporon_t *h)
{
    return h->fec_type == PPLN_FEC_RS && h->rs != NULL;
}

static bool bch_handle(const poporon_t *h)
{
    return h->fec_type == PPLN_FEC_BCH && h->bch != NULL;
}

/* 0 host, 2 device (ordinal in *device), -1 mixed */
static int classify(const void *const *ptrs, size_t count, int *device)
{
    int kind = -2, dev = -1;
    size_t i;
    for (i = 0; i < count; i++) {
        int d = -1, k;
        if (!ptrs[i])
            continue;
        k = poporon_cuda_pointer_kind(ptrs[i], &d);
        k = (k == 2) ? 2 : 0;
        if (kind == -2) {
            kind = k;
            dev = d;
        } else if (kind != k || (k == 2 && d != dev)) {
            return -1;
        }
    }
    if (device)
        *device = dev;
    return kind == -2 ? 0 : kind;
}

static bool run_job(poporon_t *h, const pp_job_t *job, void *stream, bool async)
{
    const void *ptrs[8] = {job->data, job->parity, job->pos, job->cnt, job->syn, job->ok, job->corrected, job->syn_out};
    int device = -1;
    const int kind = classify(ptrs, 8, &device);
    if (kind < 0)
        return false;
    if (kind == 2)
        return pp_run_device_job(h, job, device, stream, !async);
    if (async)
        return false; /* the async forms are device-pointer only */
    return pp_run_host_job(h, job);
}

static bool job_args_ok(const poporon_t *h, const pp_job_t *j)
{
    if (!h || !j->data || !j->parity)
        return false;
    if (h->fec_type == PPLN_FEC_RS) {
        if (!rs_handle(h))
            return false;
        if (j->op == 0 && j->size >= 65536)
            return false;
        if (j->mode == POPORON_CUDA_MODE_ERASURE &&
            (!j->pos || !j->cnt || j->pos_stride < (size_t)h->rs->nroots))
            return false;
        if (j->mode == POPORON_CUDA_MODE_EXTSYN && !j->syn)
            return false;
        return true;
    }
    if (h->fec_type == PPLN_FEC_BCH)
        return bch_handle(h) && h->bch->n <= 31 && j->mode == POPORON_CUDA_MODE_PLAIN;
    return false;
}

/* ---------------------------------------------------------------------------------- batch API */

bool poporon_encode_batch(poporon_t *h, const uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                          size_t parity_stride, size_t n)
{
    pp_job_t j;
    memset(&j, 0, sizeof j);
    j.op = 0;
    j.data = (uint8_t *)data;
    j.data_stride = data_stride;
    j.size = size;
    j.parity = parity;
    j.parity_stride = parity_stride;
    j.n = n;
    if (!job_args_ok(h, &j))
        return false;
    if (bch_handle(h) && size < (size_t)((h->bch->k + 7u) / 8u))
        return false; /* every codeword would report false (reference src/encode.c:211-213) */
    return run_job(h, &j, NULL, false);
}

bool poporon_encode_batch_async(poporon_t *h, const uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                                size_t parity_stride, size_t n, void *stream)
{
    pp_job_t j;
    memset(&j, 0, sizeof j);
    j.op = 0;
    j.data = (uint8_t *)data;
    j.data_stride = data_stride;
    j.size = size;
    j.parity = parity;
    j.parity_stride = parity_stride;
    j.n = n;
    if (!job_args_ok(h, &j))
        return false;
    return run_job(h, &j, stream, true);
}

bool poporon_syndromes_batch(poporon_t *h, const uint8_t *data, size_t data_stride, size_t size,
                             const uint8_t *parity, size_t parity_stride, size_t n, uint16_t *syndromes)
{
    pp_job_t j;
    memset(&j, 0, sizeof j);
    j.op = 2;
    j.data = (uint8_t *)data;
    j.data_stride = data_stride;
    j.size = size;
    j.parity = (uint8_t *)parity;
    j.parity_stride = parity_stride;
    j.n = n;
    j.syn_out = syndromes;
    if (!h || !rs_handle(h) || !data || !parity || !syndromes)
        return false;
    /* the reference reaches calculate_syndrome_u8 only for sizes its padding rule accepts (src/decode.c:418-429) */
    if (size < 1 || size > (size_t)(h->rs->gf->nn - h->rs->nroots))
        return false;
    return run_job(h, &j, NULL, false);
}

static bool decode_common(poporon_t *h, int mode, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                          size_t parity_stride, size_t n, const uint32_t *pos, size_t pos_stride,
                          const uint32_t *cnt, const uint16_t *syn, uint8_t *ok, uint32_t *corrected, void *stream,
                          bool async)
{
    pp_job_t j;
    memset(&j, 0, sizeof j);
    j.op = 1;
    j.mode = mode;
    j.data = data;
    j.data_stride = data_stride;
    j.size = size;
    j.parity = parity;
    j.parity_stride = parity_stride;
    j.n = n;
    j.pos = pos;
    j.pos_stride = pos_stride;
    j.cnt = cnt;
    j.syn = syn;
    j.ok = ok;
    j.corrected = corrected;
    if (!job_args_ok(h, &j))
        return false;
    return run_job(h, &j, stream, async);
}

bool poporon_decode_batch(poporon_t *h, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                          size_t parity_stride, size_t n, uint8_t *ok, uint32_t *corrected_num)
{
    return decode_common(h, POPORON_CUDA_MODE_PLAIN, data, data_stride, size, parity, parity_stride, n, NULL, 0, NULL,
                         NULL, ok, corrected_num, NULL, false);
}

bool poporon_decode_batch_erasure(poporon_t *h, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                                  size_t parity_stride, size_t n, const uint32_t *positions, size_t pos_stride,
                                  const uint32_t *counts, uint8_t *ok, uint32_t *corrected_num)
{
    if (!h || !rs_handle(h))
        return false;
    return decode_common(h, POPORON_CUDA_MODE_ERASURE, data, data_stride, size, parity, parity_stride, n, positions,
                         pos_stride, counts, NULL, ok, corrected_num, NULL, false);
}

bool poporon_decode_batch_syndrome(poporon_t *h, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                                   size_t parity_stride, size_t n, const uint16_t *syndromes, uint8_t *ok,
                                   uint32_t *corrected_num)
{
    if (!h || !rs_handle(h))
        return false;
    return decode_common(h, POPORON_CUDA_MODE_EXTSYN, data, data_stride, size, parity, parity_stride, n, NULL, 0, NULL,
                         syndromes, ok, corrected_num, NULL, false);
}

bool poporon_decode_batch_async(poporon_t *h, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                                size_t parity_stride, size_t n, const uint32_t *positions, size_t pos_stride,
                                const uint32_t *counts, uint8_t *ok, uint32_t *corrected_num, void *stream)
{
    const int mode = positions ? POPORON_CUDA_MODE_ERASURE : POPORON_CUDA_MODE_PLAIN;
    return decode_common(h, mode, data, data_stride, size, parity, parity_stride, n, positions, pos_stride, counts,
                         NULL, ok, corrected_num, stream, true);
}

/* ---------------------------------------------------------------------------- single codeword */

bool poporon_encode(poporon_t *h, uint8_t *data, size_t size, uint8_t *parity)
{
    if (!h || !data || !parity)
        return false;
    if (h->fec_type != PPLN_FEC_RS && h->fec_type != PPLN_FEC_BCH)
        return false;
    return poporon_encode_batch(h, data, size ? size : 1, size, parity, h->parity_bytes ? h->parity_bytes : 1, 1);
}

bool poporon_decode(poporon_t *h, uint8_t *data, size_t size, uint8_t *parity, size_t *corrected_num)
{
    uint8_t ok = 0;
    uint32_t nc = 0;
    bool ran;

    if (!h || !data || !parity || !size)
        return false; /* *corrected_num untouched, reference src/decode.c:598-600 */

    if (h->fec_type == PPLN_FEC_BCH) {
        if (!bch_handle(h))
            return false;
        ran = poporon_decode_batch(h, data, size, size, parity, h->parity_bytes ? h->parity_bytes : 1, 1, &ok, &nc);
        if (!ran || !ok) {
            h->last_num_errors = -1;
            return false; /* data and *corrected_num untouched, reference src/decode.c:576-579 */
        }
        h->last_num_errors = (int32_t)nc;
        if (corrected_num)
            *corrected_num = nc;
        return true;
    }
    if (!rs_handle(h))
        return false;

    if (h->ext_syndrome) { /* takes precedence over an erasure object, reference src/decode.c:446 */
        ran = poporon_decode_batch_syndrome(h, data, size, size, parity, h->parity_bytes, 1, h->ext_syndrome, &ok, &nc);
    } else if (h->erasure) {
        /* Present the borrowed list as one fully defined row of max(nroots, count) slots: the
         * reference reads slots past erasure_count when it meets more roots than erasures
         * (SURVEY Appendix A.6); our list storage is zero-initialised so those reads are 0. */
        const poporon_erasure_t *e = h->erasure;
        const uint32_t nroots = h->rs->nroots;
        uint32_t row[256];
        uint32_t cnt = e->count, i;
        /* count > nroots is passed on unchanged: the reference looks at the syndromes first (src/decode.c:466-470) and
         * returns true / 0 for a clean codeword whatever the erasure count; only a dirty one fails, in the kernel */
        for (i = 0; i < nroots; i++)
            row[i] = (i < e->capacity) ? e->positions[i] : 0u;
        ran = poporon_decode_batch_erasure(h, data, size, size, parity, h->parity_bytes, 1, row, nroots, &cnt, &ok, &nc);
    } else {
        ran = poporon_decode_batch(h, data, size, size, parity, h->parity_bytes, 1, &ok, &nc);
    }
    if (!ran) {
        ok = 0;
        nc = 0;
    }
    h->last_corrected = nc;
    if (corrected_num)
        *corrected_num = nc; /* written on success AND failure, reference src/decode.c:479-486 */
    return ok != 0;
}
