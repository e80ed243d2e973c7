/* pipeline.c — moves codewords between the caller's buffers and the device kernels.
 *
 *   device-pointer job : one launch sequence on the caller's stream, no copies.
 *   host-pointer job   : the batch is split into contiguous slices, one per lane (GPU); each lane is
 *                        driven by its own host thread and streams its slice through PP_SLOTS
 *                        staging slots (stream + device buffers) in chunks, so the H2D copy of
 *                        chunk c+1, the kernels of chunk c and the D2H copy of chunk c-1 overlap
 *                        when the caller's memory is page-locked.  No inter-GPU traffic, no
 *                        collective: codewords are independent.
 *
 * Host buffers are mirrored on the device with the caller's own strides: the data span
 * [data, data + (n-1)*data_stride + size) and the parity span are copied as contiguous byte
 * ranges (merged into one range when they overlap, e.g. packed 255-byte records), which keeps every
 * transfer a single large DMA instead of one descriptor per codeword.
 */
#include "pp_internal.h"

#include <pthread.h>

/* ------------------------------------------------------------------------------------------------
 * lanes
 * ---------------------------------------------------------------------------------------------- */

bool pp_lane_prepare(poporon_t *h, pp_lane_t *lane)
{
    int rc, s;
    if (lane->ready)
        return true;
    if (h->fec_type == PPLN_FEC_RS) {
        poporon_cuda_rs_desc_t d;
        memset(&d, 0, sizeof d);
        d.symbol_size = h->rs->gf->symbol_size;
        d.nroots = h->rs->nroots;
        d.nn = h->rs->gf->nn;
        d.gfpoly = h->rs->gf->poly;
        d.fcr = h->rs->fcr;
        d.prim = h->rs->prim;
        d.iprim = h->rs->iprim;
        d.alpha_to = h->rs->gf->alpha_to;
        d.index_of = h->rs->gf->index_of;
        d.genpoly_index = h->rs->genpoly;
        rc = poporon_cuda_rs_create(&d, lane->device, &lane->rs);
    } else {
        poporon_cuda_bch_desc_t d;
        memset(&d, 0, sizeof d);
        d.symbol_size = h->bch->gf->symbol_size;
        d.t = h->bch->t;
        d.n = h->bch->n;
        d.k = h->bch->k;
        d.parity_bits = h->bch->parity_bits;
        d.gen_poly = h->bch->gen;
        d.nn = h->bch->gf->nn;
        d.alpha_to = h->bch->gf->alpha_to;
        d.index_of = h->bch->gf->index_of;
        rc = poporon_cuda_bch_create(&d, lane->device, &lane->bch);
    }
    if (rc != 0)
        return false;
    for (s = 0; s < PP_SLOTS; s++) {
        if (poporon_cuda_stream_create(lane->device, &lane->slot[s].stream) != 0 ||
            poporon_cuda_event_create(lane->device, &lane->slot[s].kdone) != 0)
            return false;
        lane->slot[s].kdone_set = false;
    }
    lane->ready = true;
    return true;
}

void pp_handle_release_lanes(poporon_t *h)
{
    size_t l;
    int s;
    for (l = 0; l < PP_MAX_DEVICES; l++) {
        pp_lane_t *lane = &h->lane[l];
        for (s = 0; s < PP_SLOTS; s++) {
            pp_slot_t *sl = &lane->slot[s];
            if (sl->stream) {
                poporon_cuda_stream_sync(lane->device, sl->stream);
                poporon_cuda_stream_destroy(lane->device, sl->stream);
            }
            if (sl->kdone)
                poporon_cuda_event_destroy(lane->device, sl->kdone);
            poporon_cuda_free(lane->device, sl->d_span);
            poporon_cuda_free(lane->device, sl->d_pspan);
            poporon_cuda_free(lane->device, sl->d_aux);
        }
        poporon_cuda_rs_destroy(lane->rs);
        poporon_cuda_bch_destroy(lane->bch);
        memset(lane, 0, sizeof *lane);
    }
}

bool poporon_use_devices(poporon_t *h, const int *devices, size_t count)
{
    int avail = poporon_cuda_device_count();
    size_t i;
    if (!h || avail <= 0)
        return false;
    if (count == 0) {
        count = (size_t)avail;
        devices = NULL;
    }
    if (count > PP_MAX_DEVICES || (devices == NULL && count > (size_t)avail))
        return false;
    for (i = 0; i < count; i++) {
        int d = devices ? devices[i] : (int)i;
        if (d < 0 || d >= avail)
            return false;
    }
    pp_handle_release_lanes(h);
    h->nlanes = count;
    for (i = 0; i < count; i++) {
        h->lane[i].device = devices ? devices[i] : (int)i;
        if (!pp_lane_prepare(h, &h->lane[i]))
            return false;
    }
    return true;
}

size_t poporon_device_count(void)
{
    int n = poporon_cuda_device_count();
    return n > 0 ? (size_t)n : 0;
}

void *poporon_host_alloc(size_t bytes)
{
    void *p = NULL;
    if (poporon_cuda_host_alloc(bytes, &p) != 0)
        return NULL;
    return p;
}

void poporon_host_free(void *ptr)
{
    poporon_cuda_host_free(ptr);
}

const char *poporon_last_error(void)
{
    return poporon_cuda_last_error();
}

uint64_t poporon_kernel_launches(const poporon_t *h)
{
    uint64_t total = 0;
    size_t l;
    if (!h)
        return 0;
    for (l = 0; l < PP_MAX_DEVICES; l++) {
        total += poporon_cuda_rs_launch_count(h->lane[l].rs);
        total += poporon_cuda_bch_launch_count(h->lane[l].bch);
    }
    return total;
}

bool poporon_kernel_profile(poporon_t *h, bool on)
{
    size_t l;
    bool any = false;
    if (!h)
        return false;
    for (l = 0; l < PP_MAX_DEVICES; l++)
        if (h->lane[l].rs && poporon_cuda_rs_profile(h->lane[l].rs, on ? 1 : 0) == 0)
            any = true;
    return any;
}

const char *poporon_kernel_profile_read(poporon_t *h, int kernel, uint64_t *launches, double *total_ms)
{
    size_t l;
    uint64_t n = 0, ln;
    double ms = 0.0, lms;
    if (!h || !poporon_cuda_rs_profile_name(kernel))
        return NULL;
    for (l = 0; l < PP_MAX_DEVICES; l++)
        if (h->lane[l].rs && poporon_cuda_rs_profile_read(h->lane[l].rs, kernel, &ln, &lms) == 0) {
            n += ln;
            ms += lms;
        }
    if (launches)
        *launches = n;
    if (total_ms)
        *total_ms = ms;
    return poporon_cuda_rs_profile_name(kernel);
}

/* ------------------------------------------------------------------------------------------------
 * kernel dispatch on device-resident buffers
 * ---------------------------------------------------------------------------------------------- */

static int launch(poporon_t *h, pp_lane_t *lane, const pp_job_t *j, void *stream)
{
    if (h->fec_type == PPLN_FEC_RS) {
        if (j->op == 2)
            return poporon_cuda_rs_syndromes_dev(lane->rs, j->data, j->data_stride, j->size, j->parity,
                                                 j->parity_stride, j->n, j->syn_out, stream);
        if (j->op == 0)
            return poporon_cuda_rs_encode_dev(lane->rs, j->data, j->data_stride, j->size, j->parity, j->parity_stride,
                                              j->n, j->ok, stream);
        return poporon_cuda_rs_decode_dev(lane->rs, j->mode, j->data, j->data_stride, j->size, j->parity,
                                          j->parity_stride, j->n, j->pos, j->pos_stride, j->cnt, j->syn, j->ok,
                                          j->corrected, stream);
    }
    if (j->op == 0)
        return poporon_cuda_bch_encode_dev(lane->bch, j->data, j->data_stride, j->size, j->parity, j->parity_stride,
                                           j->n, j->ok, stream);
    return poporon_cuda_bch_decode_dev(lane->bch, j->data, j->data_stride, j->size, j->parity, j->parity_stride, j->n,
                                       j->ok, j->corrected, stream);
}

static pp_lane_t *lane_for_device(poporon_t *h, int device)
{
    size_t l, free_slot = PP_MAX_DEVICES;
    for (l = 0; l < PP_MAX_DEVICES; l++) {
        if (h->lane[l].ready && h->lane[l].device == device)
            return &h->lane[l];
        if (!h->lane[l].ready && l >= h->nlanes && free_slot == PP_MAX_DEVICES)
            free_slot = l;
    }
    if (free_slot == PP_MAX_DEVICES)
        return NULL;
    h->lane[free_slot].device = device;
    if (!pp_lane_prepare(h, &h->lane[free_slot]))
        return NULL;
    return &h->lane[free_slot];
}

bool pp_run_device_job(poporon_t *h, const pp_job_t *job, int device, void *stream, bool sync)
{
    pp_lane_t *lane = lane_for_device(h, device);
    if (!lane)
        return false;
    if (launch(h, lane, job, stream) != 0)
        return false;
    if (sync && poporon_cuda_stream_sync(device, stream) != 0)
        return false;
    return true;
}

/* ------------------------------------------------------------------------------------------------
 * host-pointer jobs
 * ---------------------------------------------------------------------------------------------- */

static bool ensure(int device, uint8_t **buf, size_t *cap, size_t need)
{
    void *p = NULL;
    if (need < 16)
        need = 16; /* never hand a NULL buffer to the device layer */
    if (need <= *cap)
        return true;
    poporon_cuda_free(device, *buf);
    *buf = NULL;
    *cap = 0;
    if (poporon_cuda_malloc(device, need, &p) != 0)
        return false;
    *buf = (uint8_t *)p;
    *cap = need;
    return true;
}

static size_t align_up(size_t v, size_t a)
{
    return (v + a - 1) / a * a;
}

typedef struct {
    poporon_t *h;
    pp_lane_t *lane;
    const pp_job_t *job;
    size_t lo, hi; /* codeword range of this lane */
    bool ok;
} lane_task_t;

/* Streams the codewords [t->lo, t->hi) of the job through the lane's staging slots.  Returns false at the first
 * failure, possibly with copies and kernels of earlier chunks still in flight (run_lane drains them). */
static bool run_lane_body(lane_task_t *t)
{
    poporon_t *h = t->h;
    pp_lane_t *lane = t->lane;
    const pp_job_t *j = t->job;
    const int dev = lane->device;
    const size_t P = h->parity_bytes;
    const size_t nroots = (h->fec_type == PPLN_FEC_RS) ? h->rs->nroots : 0;
    const size_t chunk = h->chunk_codewords ? h->chunk_codewords : 1;
    size_t c0;
    unsigned turn = 0;
    int s;

    for (c0 = t->lo; c0 < t->hi; c0 += chunk, turn++) {
        pp_slot_t *sl = &lane->slot[turn % PP_SLOTS];
        const size_t m = (t->hi - c0 < chunk) ? t->hi - c0 : chunk;
        uint8_t *hd = j->data + c0 * j->data_stride;
        uint8_t *hp = j->parity + c0 * j->parity_stride;
        const size_t dspan = (m - 1) * j->data_stride + j->size;
        const size_t pspan = (m - 1) * j->parity_stride + P;
        /* merged when the two spans overlap (e.g. packed records) */
        const bool merged = !(hp >= hd + dspan || hd >= hp + pspan);
        uint8_t *lo_ptr = hd, *hi_ptr = hd + dspan;
        uint8_t *dd, *dp, *d_ok, *d_cor, *d_pos = NULL, *d_cnt = NULL, *d_syn = NULL, *d_sout = NULL;
        size_t aux_need, off;
        pp_job_t dj;

        /* the slot's previous chunk (two turns ago) must have drained before its buffers are reused */
        if (poporon_cuda_stream_sync(dev, sl->stream) != 0)
            return false;

        if (merged) {
            if (hp < lo_ptr)
                lo_ptr = hp;
            if (hp + pspan > hi_ptr)
                hi_ptr = hp + pspan;
            if (!ensure(dev, &sl->d_span, &sl->span_cap, (size_t)(hi_ptr - lo_ptr)))
                return false;
            dd = sl->d_span + (hd - lo_ptr);
            dp = sl->d_span + (hp - lo_ptr);
        } else {
            if (!ensure(dev, &sl->d_span, &sl->span_cap, dspan) || !ensure(dev, &sl->d_pspan, &sl->pspan_cap, pspan))
                return false;
            dd = sl->d_span;
            dp = sl->d_pspan;
        }
        aux_need = align_up(m, 16) + m * 4;
        if (j->mode == POPORON_CUDA_MODE_ERASURE)
            aux_need += m * j->pos_stride * 4 + m * 4;
        if (j->mode == POPORON_CUDA_MODE_EXTSYN)
            aux_need = align_up(aux_need, 16) + m * nroots * 2;
        if (j->op == 2)
            aux_need = align_up(aux_need, 16) + m * nroots * 2;
        if (!ensure(dev, &sl->d_aux, &sl->aux_cap, aux_need))
            return false;
        d_ok = sl->d_aux;
        off = align_up(m, 16);
        d_cor = sl->d_aux + off;
        off += m * 4;

        /* ---- host -> device ---- */
        if (merged) {
            if (poporon_cuda_memcpy_h2d(dev, sl->d_span, lo_ptr, (size_t)(hi_ptr - lo_ptr), sl->stream) != 0)
                return false;
        } else {
            if (poporon_cuda_memcpy_h2d(dev, dd, hd, dspan, sl->stream) != 0)
                return false;
            /* encode only writes parity, but gaps between parity records must survive the round trip */
            if ((j->op != 0 || j->parity_stride != P) && poporon_cuda_memcpy_h2d(dev, dp, hp, pspan, sl->stream) != 0)
                return false;
        }
        if (j->mode == POPORON_CUDA_MODE_ERASURE && j->op == 1) {
            d_pos = sl->d_aux + off;
            off += m * j->pos_stride * 4;
            d_cnt = sl->d_aux + off;
            off += m * 4;
            if (poporon_cuda_memcpy_h2d(dev, d_pos, j->pos + c0 * j->pos_stride, m * j->pos_stride * 4, sl->stream) ||
                poporon_cuda_memcpy_h2d(dev, d_cnt, j->cnt + c0, m * 4, sl->stream))
                return false;
        }
        if (j->mode == POPORON_CUDA_MODE_EXTSYN && j->op == 1) {
            off = align_up(off, 16); /* 16-byte aligned rows: the bitsliced path reads them as vectors */
            d_syn = sl->d_aux + off;
            off += m * nroots * 2;
            if (poporon_cuda_memcpy_h2d(dev, d_syn, j->syn + c0 * nroots, m * nroots * 2, sl->stream))
                return false;
        }

        if (j->op == 2) {
            off = align_up(off, 16);
            d_sout = sl->d_aux + off;
            off += m * nroots * 2;
        }

        /* ---- kernels ---- */
        dj = *j;
        dj.data = dd;
        dj.parity = dp;
        dj.n = m;
        dj.pos = (const uint32_t *)d_pos;
        dj.cnt = (const uint32_t *)d_cnt;
        dj.syn = (const uint16_t *)d_syn;
        dj.ok = d_ok;
        dj.corrected = (uint32_t *)d_cor;
        dj.syn_out = (uint16_t *)d_sout;
        /* The device context owns ONE set of decode scratch (dirty queue, syndrome rows, counters), so the kernel
         * sections of consecutive chunks must not overlap even though they sit on different streams: wait for the
         * other slot's kernels; copies on either side still overlap them. */
        {
            int o;
            for (o = 0; o < PP_SLOTS; o++) {
                pp_slot_t *other = &lane->slot[o];
                if (other != sl && other->kdone_set &&
                    poporon_cuda_stream_wait_event(dev, sl->stream, other->kdone) != 0)
                    return false;
            }
        }
        if (launch(h, lane, &dj, sl->stream) != 0)
            return false;
        if (poporon_cuda_event_record(dev, sl->kdone, sl->stream) != 0)
            return false;
        sl->kdone_set = true;

        /* ---- device -> host ---- */
        if (j->op == 2) {
            if (poporon_cuda_memcpy_d2h(dev, j->syn_out + c0 * nroots, d_sout, m * nroots * 2, sl->stream) != 0)
                return false;
            continue;
        }
        if (j->op == 0) {
            if (poporon_cuda_memcpy_d2h(dev, hp, dp, pspan, sl->stream) != 0)
                return false;
        } else if (merged) {
            if (poporon_cuda_memcpy_d2h(dev, lo_ptr, sl->d_span, (size_t)(hi_ptr - lo_ptr), sl->stream) != 0)
                return false;
        } else {
            if (poporon_cuda_memcpy_d2h(dev, hd, dd, dspan, sl->stream) != 0)
                return false;
            /* RS corrects parity in place as well; BCH leaves it untouched */
            if (h->fec_type == PPLN_FEC_RS && poporon_cuda_memcpy_d2h(dev, hp, dp, pspan, sl->stream) != 0)
                return false;
        }
        if (j->ok && poporon_cuda_memcpy_d2h(dev, j->ok + c0, d_ok, m, sl->stream) != 0)
            return false;
        if (j->op == 1 && j->corrected &&
            poporon_cuda_memcpy_d2h(dev, j->corrected + c0, d_cor, m * 4, sl->stream) != 0)
            return false;
    }
    for (s = 0; s < PP_SLOTS; s++)
        if (poporon_cuda_stream_sync(dev, lane->slot[s].stream) != 0)
            return false;
    return true;
}


static void run_lane(lane_task_t *t)
{
    t->ok = run_lane_body(t);
    if (!t->ok) {
        /* nothing may still be writing into the caller's buffers when the job reports failure */
        int s;
        for (s = 0; s < PP_SLOTS; s++)
            (void)poporon_cuda_stream_sync(t->lane->device, t->lane->slot[s].stream);
    }
}

static void *lane_thread(void *arg)
{
    run_lane((lane_task_t *)arg);
    return NULL;
}

bool pp_run_host_job(poporon_t *h, const pp_job_t *job)
{
    lane_task_t task[PP_MAX_DEVICES];
    pthread_t tid[PP_MAX_DEVICES];
    bool spawned[PP_MAX_DEVICES];
    size_t lanes = h->nlanes ? h->nlanes : 1;
    size_t per, l, used = 0;
    bool all_ok = true;

    if (job->n == 0)
        return true;
    /* tiny batches are not worth a second GPU: keep at least 32 Ki codewords per lane */
    while (lanes > 1 && job->n / lanes < ((size_t)1 << 15))
        lanes--;
    per = align_up((job->n + lanes - 1) / lanes, 32);
    for (l = 0; l < lanes; l++) {
        size_t lo = l * per, hi = lo + per;
        if (lo >= job->n)
            break;
        if (hi > job->n)
            hi = job->n;
        if (!pp_lane_prepare(h, &h->lane[l]))
            return false;
        task[l].h = h;
        task[l].lane = &h->lane[l];
        task[l].job = job;
        task[l].lo = lo;
        task[l].hi = hi;
        task[l].ok = false;
        used = l + 1;
    }
    if (used == 1) {
        run_lane(&task[0]);
        return task[0].ok;
    }
    for (l = 0; l < used; l++) {
        spawned[l] = pthread_create(&tid[l], NULL, lane_thread, &task[l]) == 0;
        if (!spawned[l])
            run_lane(&task[l]);
    }
    for (l = 0; l < used; l++) {
        if (spawned[l])
            pthread_join(tid[l], NULL);
        all_ok = all_ok && task[l].ok;
    }
    return all_ok;
}
