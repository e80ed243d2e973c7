/*
 * poporon_cuda.h — C ABI of libpoporon_cuda.so, the sm_100a device layer of poporon-b200.
 *
 * This is the drop-in boundary of the hot path: plain pointers and sizes, `extern "C"`, no C++ or
 * framework types.  The C99 host library (libpoporon.so, include/poporon.h) is the only intended
 * caller; benchmarks and tests may also bind it directly (ctypes) to work on device-resident data.
 *
 * Each entry point replaces one piece of the reference's CPU path (paths relative to the
 * reference tree):
 *
 *   poporon_cuda_rs_create ........ uploads what poporon_rs_create / poporon_gf_create compute on the
 *                                   host (src/gf.c:29-86, src/rs.c:29-83, src/poporon.c:84-93)
 *   poporon_cuda_rs_encode_dev .... rs_encode, src/encode.c:17-145, applied to n codewords
 *   poporon_cuda_rs_decode_dev .... rs_decode = calculate_syndrome_u8 + error_correction_u8,
 *                                   src/decode.c:17-487, all three modes (plain / erasure object /
 *                                   external syndrome), applied to n codewords
 *   poporon_cuda_bch_create ....... poporon_bch_create, src/bch.c:288-322
 *   poporon_cuda_bch_encode_dev ... bch_encode wrapper + poporon_bch_encode, src/encode.c:199-234,
 *                                   src/bch.c:352-380
 *   poporon_cuda_bch_decode_dev ... bch_decode wrapper + poporon_bch_decode, src/decode.c:542-594,
 *                                   src/bch.c:382-435
 *
 * There is no CPU fallback anywhere behind this ABI: if no CUDA device is usable every call
 * returns POPORON_CUDA_ERR_DEVICE and the host library reports failure.
 *
 * Conventions
 *   - All `*_dev` functions take DEVICE pointers and enqueue work on `stream` (a cudaStream_t passed
 *     as void*; NULL = the legacy default stream) of the context's device.  They do not synchronise.
 *   - Codeword i lives at data + i*data_stride (size bytes) and parity + i*parity_stride.
 *     Packed records are expressed as parity = data + size with both strides = size + parity bytes.
 *   - ok[i] receives the bool the reference's poporon_encode/poporon_decode would return for
 *     codeword i; corrected[i] receives what *corrected_num would hold had it been 0 before the call.
 *   - Return value: 0 on success, a negative POPORON_CUDA_ERR_* otherwise.
 */
#ifndef POPORON_CUDA_H
#define POPORON_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define POPORON_CUDA_OK          0
#define POPORON_CUDA_ERR_DEVICE  (-1) /* no device / CUDA runtime error (see poporon_cuda_last_error) */
#define POPORON_CUDA_ERR_ARG     (-2) /* invalid argument */
#define POPORON_CUDA_ERR_NOMEM   (-3) /* device or pinned allocation failed */
#define POPORON_CUDA_ERR_UNSUPP  (-4) /* parameters outside the supported domain */

/* decode modes, same precedence as src/decode.c:446-477 */
#define POPORON_CUDA_MODE_PLAIN   0
#define POPORON_CUDA_MODE_ERASURE 1
#define POPORON_CUDA_MODE_EXTSYN  2

typedef struct poporon_cuda_rs poporon_cuda_rs_t;
typedef struct poporon_cuda_bch poporon_cuda_bch_t;

/* Host-computed description of one RS code (what the reference keeps in poporon_rs_t + poporon_gf_t +
 * decoder_buffer_t::primitive_inverse).  Tables use the reference's encoding: alpha_to = log2exp,
 * index_of = exp2log, both nn+1 entries, zero's log is nn. */
typedef struct {
    uint8_t symbol_size;           /* m, 1..8 */
    uint8_t nroots;                /* 1..nn-1 */
    uint16_t nn;                   /* 2^m - 1 */
    uint16_t gfpoly;
    uint16_t fcr;
    uint16_t prim;
    uint16_t iprim;
    const uint16_t *alpha_to;      /* [nn+1] */
    const uint16_t *index_of;      /* [nn+1] */
    const uint16_t *genpoly_index; /* [nroots+1], index form (src/rs.c:78-80) */
} poporon_cuda_rs_desc_t;

/* Host-computed description of one binary BCH code (struct _poporon_bch_t, src/bch.c:15-23). */
typedef struct {
    uint8_t symbol_size; /* m, 3..5 for device execution (n <= 31) */
    uint8_t t;
    uint16_t n;
    uint16_t k;
    uint16_t parity_bits;
    uint32_t gen_poly;
    uint16_t nn;
    const uint16_t *alpha_to; /* [nn+1] */
    const uint16_t *index_of; /* [nn+1] */
} poporon_cuda_bch_desc_t;

/* ---- device management ------------------------------------------------------------------- */
int poporon_cuda_device_count(void);                 /* >= 0, or a negative error */
const char *poporon_cuda_last_error(void);           /* thread-local text of the last failure */
void poporon_cuda_note_error(const char *text);      /* lets the host library leave its own text there */
int poporon_cuda_device_sm_count(int device);

/* ---- memory / stream helpers so a C host never needs cuda_runtime.h ------------------------ */
int poporon_cuda_malloc(int device, size_t bytes, void **dptr);
int poporon_cuda_free(int device, void *dptr);
int poporon_cuda_host_alloc(size_t bytes, void **hptr); /* page-locked */
int poporon_cuda_host_free(void *hptr);
int poporon_cuda_stream_create(int device, void **stream);
int poporon_cuda_stream_destroy(int device, void *stream);
int poporon_cuda_stream_sync(int device, void *stream);
int poporon_cuda_memcpy_h2d(int device, void *dst, const void *src, size_t bytes, void *stream);
int poporon_cuda_memcpy_d2h(int device, void *dst, const void *src, size_t bytes, void *stream);
int poporon_cuda_memcpy2d_h2d(int device, void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                              size_t height, void *stream);
int poporon_cuda_memcpy2d_d2h(int device, void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                              size_t height, void *stream);
int poporon_cuda_memset(int device, void *dst, int value, size_t bytes, void *stream);
/* 0 = pageable host / unknown, 1 = pinned host, 2 = device memory (ordinal in *device), <0 error */
int poporon_cuda_pointer_kind(const void *p, int *device);
int poporon_cuda_event_create(int device, void **event);
int poporon_cuda_event_destroy(int device, void *event);
int poporon_cuda_event_record(int device, void *event, void *stream);
int poporon_cuda_event_elapsed_ms(int device, void *start, void *stop, float *ms); /* synchronises on stop */
int poporon_cuda_stream_wait_event(int device, void *stream, void *event);          /* cudaStreamWaitEvent */

/* ---- Reed-Solomon -------------------------------------------------------------------------- */
int poporon_cuda_rs_create(const poporon_cuda_rs_desc_t *desc, int device, poporon_cuda_rs_t **out);
void poporon_cuda_rs_destroy(poporon_cuda_rs_t *ctx);
int poporon_cuda_rs_device(const poporon_cuda_rs_t *ctx);

/* Which kernel family a call with these arguments would take: 0 = general-parameter table kernels,
 * 1 = RS(255,223)/0x11D bitsliced fast path: any base alignment, message rows and parity rows as separate arrays
 * (the two-pointer shape of reference include/poporon.h:90-91) or interleaved in one record, as long as 32
 * consecutive rows of each fit a ~8 KB staging buffer (strides up to ~260 bytes); longer strides take it with a
 * 16-byte aligned base (decode: packed records only).  op: 0 encode, 1 decode.  Introspection for tests and the bench. */
int poporon_cuda_rs_path(const poporon_cuda_rs_t *ctx, int op, const uint8_t *data, size_t data_stride, size_t size,
                         const uint8_t *parity, size_t parity_stride);
/* Force the generic kernels even where the fast path applies (tests compare the two). */
void poporon_cuda_rs_force_generic(poporon_cuda_rs_t *ctx, int on);

/* Calls on one context share its scratch: they are ordered on the device by an internal event whatever stream each
 * is given (two calls never overlap), and must be issued from one host thread at a time. */
int poporon_cuda_rs_encode_dev(poporon_cuda_rs_t *ctx, const uint8_t *data, size_t data_stride, size_t size,
                               uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok /* nullable */,
                               void *stream);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * poporon_cuda_rs_encode_dev on the bitsliced path (RS(255,223/239/247), 0x11D, fcr = 1, prim = 1; row layouts for
 * which poporon_cuda_rs_path(ctx, 0, ...) == 1 through the transform kernels), but the scratch between its three
 * kernels is a file-scope device array of 20 905 984 bytes (64 slabs of 255 KB + 64 bytes of planes per codeword)
 * that every GPU holds from the moment it loads the library until the process exits.  The call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine); its first use on a device looks the array's address up
 * with cudaGetSymbolAddress.  Limits: n <= POPORON_CUDA_NOALLOC_MAX_CODEWORDS per call, else POPORON_CUDA_ERR_UNSUPP
 * with the reason in poporon_cuda_last_error (other codes / layouts likewise).  ALL calls on one device share the
 * array, whatever context or stream: unlike poporon_cuda_rs_encode_dev they are NOT ordered against each other or
 * against the allocating entry points, so two calls that overlap on the GPU race - keep them on one stream or order
 * them with events.  Nothing in the array needs clearing between calls (every word read was written by the same
 * call).  ok is nullable. */
#define POPORON_CUDA_NOALLOC_MAX_CODEWORDS 65536
int poporon_cuda_rs_encode_noalloc_dev(poporon_cuda_rs_t *ctx, const uint8_t *data, size_t data_stride, size_t size,
                                       uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok /* nullable */,
                                       void *stream);

/* mode PLAIN:   pos/cnt/syn ignored.
 * mode ERASURE: pos = [n][pos_stride] uint32 data-relative positions, EVERY slot defined by the caller,
 *               pos_stride >= nroots; cnt = [n], each <= nroots.  Reproduces the reference's
 *               slot-order application (src/decode.c:211-214); positions >= size are not written.
 * mode EXTSYN:  syn = [n][nroots] uint16 index-form syndromes (nn = zero), src/decode.c:446-464.
 * ok / corrected are nullable. */
int poporon_cuda_rs_decode_dev(poporon_cuda_rs_t *ctx, int mode, uint8_t *data, size_t data_stride, size_t size,
                               uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *pos,
                               size_t pos_stride, const uint32_t *cnt, const uint16_t *syn, uint8_t *ok,
                               uint32_t *corrected, void *stream);
/* Syndromes only (reference calculate_syndrome_u8, src/decode.c:232-416, a stage the reference keeps internal):
 * syndromes[i * nroots + j] = index_of[S_(j)] of codeword i, nn = 2^m - 1 meaning zero - the form
 * poporon_cuda_rs_decode_dev(mode = EXTSYN) takes back.  Nothing is corrected.  size must satisfy the reference's
 * padding rule (1 <= size <= nn - nroots), else POPORON_CUDA_ERR_ARG. */
int poporon_cuda_rs_syndromes_dev(poporon_cuda_rs_t *ctx, const uint8_t *data, size_t data_stride, size_t size,
                                  const uint8_t *parity, size_t parity_stride, size_t n, uint16_t *syndromes,
                                  void *stream);

/* Number of kernels this context has launched since creation (bench.py's gpu_launches). */
uint64_t poporon_cuda_rs_launch_count(const poporon_cuda_rs_t *ctx);

/* Opt-in per-kernel timing of the RS(255,223) fast path: CUDA event pairs recorded around every kernel the
 * context launches, on the stream it launches on.  profile(ctx, 1) starts/restarts accumulation,
 * profile(ctx, 0) stops it; profile_read synchronises on the recorded events and returns launches and the
 * summed duration of kernel id `kernel` (0 .. 9, names from profile_name).  bench.py's roofline leg. */
int poporon_cuda_rs_profile(poporon_cuda_rs_t *ctx, int on);
int poporon_cuda_rs_profile_read(poporon_cuda_rs_t *ctx, int kernel, uint64_t *launches, double *total_ms);
const char *poporon_cuda_rs_profile_name(int kernel);

/* ---- binary BCH ---------------------------------------------------------------------------- */
int poporon_cuda_bch_create(const poporon_cuda_bch_desc_t *desc, int device, poporon_cuda_bch_t **out);
void poporon_cuda_bch_destroy(poporon_cuda_bch_t *ctx);
int poporon_cuda_bch_encode_dev(poporon_cuda_bch_t *ctx, const uint8_t *data, size_t data_stride, size_t size,
                                uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, void *stream);
int poporon_cuda_bch_decode_dev(poporon_cuda_bch_t *ctx, uint8_t *data, size_t data_stride, size_t size,
                                const uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok,
                                uint32_t *corrected, void *stream);
uint64_t poporon_cuda_bch_launch_count(const poporon_cuda_bch_t *ctx);

#ifdef __cplusplus
}
#endif

#endif /* POPORON_CUDA_H */
