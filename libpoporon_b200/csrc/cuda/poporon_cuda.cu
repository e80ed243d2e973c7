// poporon_cuda.cu — the extern "C" shim of libpoporon_cuda.so (see include/poporon_cuda.h).
// Device/stream/memory helpers, context lifecycle and dispatch to the kernel families.
#include "common.cuh"
#include "bch.cuh"
#include "rs_generic.cuh"
#include "rs255.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

namespace ppcu {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int check(cudaError_t e, const char *what)
{
    if (e == cudaSuccess)
        return POPORON_CUDA_OK;
    set_error("%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    if (e == cudaErrorMemoryAllocation)
        return POPORON_CUDA_ERR_NOMEM;
    return POPORON_CUDA_ERR_DEVICE;
}

ProfScope::ProfScope(poporon_cuda_rs *c, int id, cudaStream_t s) : ctx(c), stream(s), slot(-1)
{
    if (!c || !c->prof_on || !c->prof_ev || c->prof_used >= kProfCap)
        return;
    slot = c->prof_used++;
    c->prof_id[slot] = (uint8_t)id;
    cudaEventRecord(c->prof_ev[2 * slot], s);
}

ProfScope::~ProfScope()
{
    if (slot >= 0)
        cudaEventRecord(ctx->prof_ev[2 * slot + 1], stream);
}

static void prof_drain(poporon_cuda_rs *ctx)
{
    for (int i = 0; i < ctx->prof_used; i++) {
        float ms = 0.f;
        if (cudaEventSynchronize(ctx->prof_ev[2 * i + 1]) == cudaSuccess &&
            cudaEventElapsedTime(&ms, ctx->prof_ev[2 * i], ctx->prof_ev[2 * i + 1]) == cudaSuccess) {
            ctx->prof_ms[ctx->prof_id[i]] += ms;
            ctx->prof_n[ctx->prof_id[i]]++;
        }
    }
    ctx->prof_used = 0;
}

static void prof_release(poporon_cuda_rs *ctx)
{
    if (ctx->prof_ev)
        for (int i = 0; i < 2 * kProfCap; i++)
            if (ctx->prof_ev[i])
                cudaEventDestroy(ctx->prof_ev[i]);
    free(ctx->prof_ev);
    free(ctx->prof_id);
    ctx->prof_ev = nullptr;
    ctx->prof_id = nullptr;
    ctx->prof_used = 0;
}

} // namespace ppcu

using namespace ppcu;

#define GUARD(dev)                                                   \
    DeviceGuard _guard(dev);                                         \
    if (!_guard.ok)                                                  \
    return check(cudaErrorInvalidDevice, "cudaSetDevice")

extern "C" {

const char *poporon_cuda_last_error(void)
{
    return g_err;
}

void poporon_cuda_note_error(const char *text)
{
    set_error("%s", text ? text : "");
}

int poporon_cuda_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        check(e, "cudaGetDeviceCount");
        (void)cudaGetLastError();
        return 0;
    }
    return n;
}

int poporon_cuda_device_sm_count(int device)
{
    int n = 0;
    if (check(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device), "sm count") != 0)
        return POPORON_CUDA_ERR_DEVICE;
    return n;
}

int poporon_cuda_malloc(int device, size_t bytes, void **dptr)
{
    GUARD(device);
    if (!dptr)
        return POPORON_CUDA_ERR_ARG;
    *dptr = nullptr;
    if (bytes == 0)
        return 0;
    return check(cudaMalloc(dptr, bytes), "cudaMalloc");
}

int poporon_cuda_free(int device, void *dptr)
{
    GUARD(device);
    return dptr ? check(cudaFree(dptr), "cudaFree") : 0;
}

int poporon_cuda_host_alloc(size_t bytes, void **hptr)
{
    if (!hptr)
        return POPORON_CUDA_ERR_ARG;
    *hptr = nullptr;
    return check(cudaHostAlloc(hptr, bytes ? bytes : 1, cudaHostAllocPortable), "cudaHostAlloc");
}

int poporon_cuda_host_free(void *hptr)
{
    return hptr ? check(cudaFreeHost(hptr), "cudaFreeHost") : 0;
}

int poporon_cuda_stream_create(int device, void **stream)
{
    GUARD(device);
    cudaStream_t s;
    PP_CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = (void *)s;
    return 0;
}

int poporon_cuda_stream_destroy(int device, void *stream)
{
    GUARD(device);
    return check(cudaStreamDestroy((cudaStream_t)stream), "cudaStreamDestroy");
}

int poporon_cuda_stream_sync(int device, void *stream)
{
    GUARD(device);
    return check(cudaStreamSynchronize((cudaStream_t)stream), "cudaStreamSynchronize");
}

int poporon_cuda_memcpy_h2d(int device, void *dst, const void *src, size_t bytes, void *stream)
{
    GUARD(device);
    if (bytes == 0)
        return 0;
    return check(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream), "memcpy h2d");
}

int poporon_cuda_memcpy_d2h(int device, void *dst, const void *src, size_t bytes, void *stream)
{
    GUARD(device);
    if (bytes == 0)
        return 0;
    return check(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream), "memcpy d2h");
}

int poporon_cuda_memcpy2d_h2d(int device, void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                              size_t height, void *stream)
{
    GUARD(device);
    if (width == 0 || height == 0)
        return 0;
    return check(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, cudaMemcpyHostToDevice,
                                   (cudaStream_t)stream),
                 "memcpy2d h2d");
}

int poporon_cuda_memcpy2d_d2h(int device, void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                              size_t height, void *stream)
{
    GUARD(device);
    if (width == 0 || height == 0)
        return 0;
    return check(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, cudaMemcpyDeviceToHost,
                                   (cudaStream_t)stream),
                 "memcpy2d d2h");
}

int poporon_cuda_memset(int device, void *dst, int value, size_t bytes, void *stream)
{
    GUARD(device);
    if (bytes == 0)
        return 0;
    return check(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream), "memset");
}

int poporon_cuda_pointer_kind(const void *p, int *device)
{
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    if (device)
        *device = a.device;
    switch (a.type) {
    case cudaMemoryTypeHost:
        return 1;
    case cudaMemoryTypeDevice:
    case cudaMemoryTypeManaged:
        return 2;
    default:
        return 0;
    }
}

int poporon_cuda_event_create(int device, void **event)
{
    GUARD(device);
    cudaEvent_t ev;
    PP_CUDA_TRY(cudaEventCreate(&ev));
    *event = (void *)ev;
    return 0;
}

int poporon_cuda_event_destroy(int device, void *event)
{
    GUARD(device);
    return check(cudaEventDestroy((cudaEvent_t)event), "cudaEventDestroy");
}

int poporon_cuda_event_record(int device, void *event, void *stream)
{
    GUARD(device);
    return check(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream), "cudaEventRecord");
}

int poporon_cuda_event_elapsed_ms(int device, void *start, void *stop, float *ms)
{
    GUARD(device);
    PP_CUDA_TRY(cudaEventSynchronize((cudaEvent_t)stop));
    return check(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop), "cudaEventElapsedTime");
}

int poporon_cuda_stream_wait_event(int device, void *stream, void *event)
{
    GUARD(device);
    return check(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0), "cudaStreamWaitEvent");
}

/* ------------------------------------------------------------------------------------------ RS */

namespace {
/* A context owns ONE set of scratch (dirty queue, syndrome rows, counters, bit-plane slab, syndrome planes), so two
 * calls on the same handle must not overlap on the device even when the caller puts them on different streams: the
 * stream of a call first waits for the event recorded behind the previous call's last kernel (a no-op on the same
 * stream), and records it again behind its own.  Host-side the calls of one handle must still come from one thread
 * at a time (include/poporon_cuda.h). */
struct ScratchOrder {
    poporon_cuda_rs *ctx;
    cudaStream_t stream;
    ScratchOrder(poporon_cuda_rs *c, cudaStream_t s) : ctx(c), stream(s) { cudaStreamWaitEvent(stream, ctx->busy, 0); }
    ~ScratchOrder() { cudaEventRecord(ctx->busy, stream); }
};
} // namespace

int poporon_cuda_rs_create(const poporon_cuda_rs_desc_t *d, int device, poporon_cuda_rs_t **out)
{
    if (!d || !out || !d->alpha_to || !d->index_of || !d->genpoly_index)
        return POPORON_CUDA_ERR_ARG;
    *out = nullptr;
    if (d->symbol_size < 1 || d->symbol_size > 8 || d->nn != (1u << d->symbol_size) - 1u)
        return POPORON_CUDA_ERR_UNSUPP;
    if (d->nroots < 1 || d->nroots >= d->nn || d->prim == 0)
        return POPORON_CUDA_ERR_UNSUPP;
    GUARD(device);

    RsTables h;
    memset(&h, 0, sizeof h);
    for (unsigned i = 0; i <= d->nn; i++) {
        h.alpha_to[i] = (uint8_t)d->alpha_to[i];
        h.index_of[i] = (uint8_t)d->index_of[i];
    }
    for (unsigned i = 0; i <= d->nroots; i++)
        h.genpoly[i] = (uint8_t)d->genpoly_index[i];

    poporon_cuda_rs *ctx = (poporon_cuda_rs *)calloc(1, sizeof *ctx);
    if (!ctx)
        return POPORON_CUDA_ERR_NOMEM;
    ctx->device = device;
    ctx->sm_count = sm_count_of(device);
    ctx->code.m = d->symbol_size;
    ctx->code.nn = d->nn;
    ctx->code.nroots = d->nroots;
    ctx->code.fcr = d->fcr;
    ctx->code.prim = d->prim;
    ctx->code.iprim = d->iprim;
    ctx->fast255 = rs255_matches(d);
    int rc = check(cudaMalloc((void **)&ctx->d_tables, sizeof h), "cudaMalloc tables");
    if (rc == 0)
        rc = check(cudaMemcpy(ctx->d_tables, &h, sizeof h, cudaMemcpyHostToDevice), "upload tables");
    if (rc == 0)
        rc = check(cudaEventCreateWithFlags(&ctx->busy, cudaEventDisableTiming), "cudaEventCreate");
    if (rc == 0 && ctx->fast255)
        rc = rs255_init(ctx);
    if (rc != 0) {
        poporon_cuda_rs_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return 0;
}

void poporon_cuda_rs_destroy(poporon_cuda_rs_t *ctx)
{
    if (!ctx)
        return;
    DeviceGuard g(ctx->device);
    prof_release(ctx);
    if (ctx->busy)
        cudaEventDestroy(ctx->busy);
    cudaFree(ctx->d_tables);
    cudaFree(ctx->d_queue);
    cudaFree(ctx->d_rem);
    cudaFree(ctx->d_slab);
    cudaFree(ctx->d_synplanes);
    cudaFree(ctx->d_counters);
    free(ctx);
}

int poporon_cuda_rs_device(const poporon_cuda_rs_t *ctx)
{
    return ctx ? ctx->device : POPORON_CUDA_ERR_ARG;
}

void poporon_cuda_rs_force_generic(poporon_cuda_rs_t *ctx, int on)
{
    if (ctx)
        ctx->force_generic = on != 0;
}

uint64_t poporon_cuda_rs_launch_count(const poporon_cuda_rs_t *ctx)
{
    return ctx ? ctx->launches : 0;
}

int poporon_cuda_rs_profile(poporon_cuda_rs_t *ctx, int on)
{
    if (!ctx)
        return POPORON_CUDA_ERR_ARG;
    GUARD(ctx->device);
    if (on && !ctx->prof_ev) {
        ctx->prof_ev = (cudaEvent_t *)calloc(2 * kProfCap, sizeof(cudaEvent_t));
        ctx->prof_id = (uint8_t *)calloc(kProfCap, 1);
        if (!ctx->prof_ev || !ctx->prof_id) {
            prof_release(ctx);
            return POPORON_CUDA_ERR_NOMEM;
        }
        for (int i = 0; i < 2 * kProfCap; i++)
            if (cudaEventCreate(&ctx->prof_ev[i]) != cudaSuccess) {
                prof_release(ctx);
                return check(cudaGetLastError(), "cudaEventCreate (profile)");
            }
    }
    if (on) { /* (re)start: forget what was accumulated */
        prof_drain(ctx);
        memset(ctx->prof_ms, 0, sizeof ctx->prof_ms);
        memset(ctx->prof_n, 0, sizeof ctx->prof_n);
    }
    ctx->prof_on = on != 0;
    return 0;
}

int poporon_cuda_rs_profile_read(poporon_cuda_rs_t *ctx, int kernel, uint64_t *launches, double *total_ms)
{
    if (!ctx || kernel < 0 || kernel >= PROF_KINDS)
        return POPORON_CUDA_ERR_ARG;
    GUARD(ctx->device);
    prof_drain(ctx);
    if (launches)
        *launches = ctx->prof_n[kernel];
    if (total_ms)
        *total_ms = ctx->prof_ms[kernel];
    return 0;
}

const char *poporon_cuda_rs_profile_name(int kernel)
{
    static const char *names[PROF_KINDS] = {"rs255_fft_kernel<ENCODE>", "rs255_fft_kernel<SCREEN>",
                                            "rs255_compact_kernel",     "rs255_solve_kernel",
                                            "rs255_parity_kernel",      "rs255_correct_kernel",
                                            "rs_generic kernels",       "rs255_rows_kernel",
                                            "rs255_erasure_kernel",     "rs255_ee_kernel"};
    return kernel >= 0 && kernel < PROF_KINDS ? names[kernel] : NULL;
}

int poporon_cuda_rs_path(const poporon_cuda_rs_t *ctx, int op, const uint8_t *data, size_t data_stride, size_t size,
                         const uint8_t *parity, size_t parity_stride)
{
    if (!ctx)
        return POPORON_CUDA_ERR_ARG;
    return (ctx->fast255 && !ctx->force_generic && rs255_shape_ok(ctx, op, data, data_stride, size, parity, parity_stride))
               ? 1
               : 0;
}

int poporon_cuda_rs_encode_dev(poporon_cuda_rs_t *ctx, const uint8_t *data, size_t data_stride, size_t size,
                               uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, void *stream)
{
    if (!ctx || (n && (!data || !parity)))
        return POPORON_CUDA_ERR_ARG;
    if (size >= 65536) /* the reference's loop counter is a uint16_t (src/encode.c:121) */
        return POPORON_CUDA_ERR_UNSUPP;
    GUARD(ctx->device);
    if (poporon_cuda_rs_path(ctx, 0, data, data_stride, size, parity, parity_stride) == 1) {
        ScratchOrder order(ctx, (cudaStream_t)stream);
        return rs255_encode(ctx, data, data_stride, size, parity, parity_stride, n, ok, (cudaStream_t)stream);
    }
    return launch_rs_encode_generic(ctx, data, data_stride, size, parity, parity_stride, n, ok, (cudaStream_t)stream);
}

int poporon_cuda_rs_decode_dev(poporon_cuda_rs_t *ctx, int mode, uint8_t *data, size_t data_stride, size_t size,
                               uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *pos,
                               size_t pos_stride, const uint32_t *cnt, const uint16_t *syn, uint8_t *ok,
                               uint32_t *corrected, void *stream)
{
    if (!ctx || (n && (!data || !parity)))
        return POPORON_CUDA_ERR_ARG;
    if (mode == POPORON_CUDA_MODE_ERASURE && n && (!pos || !cnt || pos_stride < (size_t)ctx->code.nroots))
        return POPORON_CUDA_ERR_ARG;
    if (mode == POPORON_CUDA_MODE_EXTSYN && n && !syn)
        return POPORON_CUDA_ERR_ARG;
    if (mode < 0 || mode > 2)
        return POPORON_CUDA_ERR_ARG;
    GUARD(ctx->device);
    /* caller-supplied syndromes also take the bitsliced solver (their rows must be 16-byte aligned for it) */
    if ((mode != POPORON_CUDA_MODE_EXTSYN || ((uintptr_t)syn & 15u) == 0) && n < ((size_t)1 << 30) &&
        poporon_cuda_rs_path(ctx, 1, data, data_stride, size, parity, parity_stride) == 1) {
        ScratchOrder order(ctx, (cudaStream_t)stream);
        return rs255_decode(ctx, mode, data, data_stride, size, parity, parity_stride, n, pos, pos_stride, cnt, syn, ok,
                            corrected, (cudaStream_t)stream);
    }
    return launch_rs_decode_generic(ctx, mode, data, data_stride, size, parity, parity_stride, n, pos, pos_stride, cnt,
                                    syn, ok, corrected, (cudaStream_t)stream);
}

int poporon_cuda_rs_syndromes_dev(poporon_cuda_rs_t *ctx, const uint8_t *data, size_t data_stride, size_t size,
                                  const uint8_t *parity, size_t parity_stride, size_t n, uint16_t *syndromes,
                                  void *stream)
{
    if (!ctx || (n && (!data || !parity || !syndromes)))
        return POPORON_CUDA_ERR_ARG;
    /* the reference computes syndromes only for sizes its padding rule accepts (src/decode.c:418-429) */
    if (size < 1 || size > (size_t)(ctx->code.nn - ctx->code.nroots))
        return POPORON_CUDA_ERR_ARG;
    GUARD(ctx->device);
    if (n < ((size_t)1 << 30) && ((uintptr_t)syndromes & 7u) == 0 &&
        poporon_cuda_rs_path(ctx, 1, data, data_stride, size, parity, parity_stride) == 1) {
        ScratchOrder order(ctx, (cudaStream_t)stream);
        const int rc = rs255_syndromes(ctx, data, data_stride, size, parity, parity_stride, n, syndromes,
                                       (cudaStream_t)stream);
        if (rc != POPORON_CUDA_ERR_UNSUPP)
            return rc;
    }
    return launch_rs_decode_generic(ctx, POPORON_CUDA_MODE_PLAIN, const_cast<uint8_t *>(data), data_stride, size,
                                    const_cast<uint8_t *>(parity), parity_stride, n, nullptr, 0, nullptr, nullptr,
                                    nullptr, nullptr, (cudaStream_t)stream, syndromes);
}

/* ------------------------------------------------------------------------------------------ BCH */

int poporon_cuda_bch_create(const poporon_cuda_bch_desc_t *d, int device, poporon_cuda_bch_t **out)
{
    if (!d || !out || !d->alpha_to || !d->index_of)
        return POPORON_CUDA_ERR_ARG;
    *out = nullptr;
    /* the reference's word arithmetic is only defined for n <= 31 (src/bch.c:39,396) */
    if (d->symbol_size < 3 || d->symbol_size > 5 || d->n != (1u << d->symbol_size) - 1u || d->n > 31)
        return POPORON_CUDA_ERR_UNSUPP;
    if (d->t < 1 || d->t > 16 || d->parity_bits > d->n || d->k + d->parity_bits != d->n)
        return POPORON_CUDA_ERR_UNSUPP;
    GUARD(device);
    uint8_t h[512];
    memset(h, 0, sizeof h);
    for (unsigned i = 0; i <= d->nn; i++) {
        h[i] = (uint8_t)d->alpha_to[i];
        h[256 + i] = (uint8_t)d->index_of[i];
    }
    poporon_cuda_bch *ctx = (poporon_cuda_bch *)calloc(1, sizeof *ctx);
    if (!ctx)
        return POPORON_CUDA_ERR_NOMEM;
    ctx->device = device;
    ctx->sm_count = sm_count_of(device);
    ctx->m = d->symbol_size;
    ctx->nn = d->nn;
    ctx->n = d->n;
    ctx->k = d->k;
    ctx->parity_bits = d->parity_bits;
    ctx->t = d->t;
    ctx->gen = d->gen_poly;
    int rc = check(cudaMalloc((void **)&ctx->d_tables, sizeof h), "cudaMalloc bch tables");
    if (rc == 0)
        rc = check(cudaMemcpy(ctx->d_tables, h, sizeof h, cudaMemcpyHostToDevice), "upload bch tables");
    if (rc == 0)
        rc = bch_build_decode_map(ctx);
    if (rc != 0) {
        poporon_cuda_bch_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return 0;
}

void poporon_cuda_bch_destroy(poporon_cuda_bch_t *ctx)
{
    if (!ctx)
        return;
    DeviceGuard g(ctx->device);
    cudaFree(ctx->d_tables);
    cudaFree(ctx->d_map);
    cudaFree(ctx->d_smap);
    free(ctx);
}

uint64_t poporon_cuda_bch_launch_count(const poporon_cuda_bch_t *ctx)
{
    return ctx ? ctx->launches : 0;
}

int poporon_cuda_bch_encode_dev(poporon_cuda_bch_t *ctx, const uint8_t *data, size_t data_stride, size_t size,
                                uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, void *stream)
{
    if (!ctx || (n && (!data || !parity)))
        return POPORON_CUDA_ERR_ARG;
    GUARD(ctx->device);
    return launch_bch_encode(ctx, data, data_stride, size, parity, parity_stride, n, ok, (cudaStream_t)stream);
}

int poporon_cuda_bch_decode_dev(poporon_cuda_bch_t *ctx, uint8_t *data, size_t data_stride, size_t size,
                                const uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok,
                                uint32_t *corrected, void *stream)
{
    if (!ctx || (n && (!data || !parity)))
        return POPORON_CUDA_ERR_ARG;
    GUARD(ctx->device);
    return launch_bch_decode(ctx, data, data_stride, size, parity, parity_stride, n, ok, corrected,
                             (cudaStream_t)stream);
}

} /* extern "C" */
