/*
 * poporon/batch.h — batched entry points (new in poporon-b200; not part of the reference API).
 *
 * Each call applies exactly the single-instance semantics of poporon_encode / poporon_decode
 * (reference src/encode.c:236-252, src/decode.c:596-612) to n independent codewords:
 *
 *     for i in 0..n-1:
 *         ok[i] = poporon_{en,de}code(pprn, data + i*data_stride, size, parity + i*parity_stride, &c)
 *         corrected[i] = c            (c starts at 0)
 *
 * Layout.  Codeword i occupies data[i*data_stride .. +size) and parity[i*parity_stride .. +P) with
 * P = poporon_get_parity_size().  Packed records (message immediately followed by its parity) are
 * expressed as parity = data + size, data_stride = parity_stride = size + P.
 *
 * Pointers.  All buffer arguments of one call must be of the same kind:
 *   - host memory: the batch is split contiguously over the handle's devices (one host thread per
 *     device, no inter-GPU traffic), each slice streamed host -> device -> host in chunks on two
 *     CUDA streams.  Page-locked buffers (poporon_host_alloc) overlap copies with kernels; pageable
 *     buffers work but copy synchronously.
 *   - device memory (all on one device): kernels are launched directly on `stream`
 *     (poporon_*_batch_async) or on the default stream followed by a synchronise (the plain forms).
 *
 * Return value: true when the batch was processed (per-codeword results are in ok[]); false on
 * NULL/invalid arguments or a device failure, in which case outputs are unspecified.
 */
#ifndef POPORON_BATCH_H
#define POPORON_BATCH_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

struct _poporon_t;

/* Devices a handle may use for host-pointer batches.  Default: device 0 only.  count = 0 means
 * "all visible devices".  Device-pointer calls always run where the data lives. */
bool poporon_use_devices(struct _poporon_t *pprn, const int *devices, size_t count);
size_t poporon_device_count(void);

/* page-locked host buffers for full copy/compute overlap */
void *poporon_host_alloc(size_t bytes);
void poporon_host_free(void *ptr);

bool poporon_encode_batch(struct _poporon_t *pprn, const uint8_t *data, size_t data_stride, size_t size,
                          uint8_t *parity, size_t parity_stride, size_t n);

/* ok and corrected_num may be NULL */
bool poporon_decode_batch(struct _poporon_t *pprn, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                          size_t parity_stride, size_t n, uint8_t *ok, uint32_t *corrected_num);

/* RS erasure decoding with per-codeword position rows: positions is [n][pos_stride] uint32
 * (data-relative), counts is [n].  Semantics are those of a handle created with an erasure object
 * (reference src/decode.c:466-473), including the slot-order application of corrections
 * (src/decode.c:211-214): EVERY slot of a row must be defined (zero-fill unused slots),
 * pos_stride >= num_roots, counts[i] <= num_roots, positions < size.  RS handles only. */
bool poporon_decode_batch_erasure(struct _poporon_t *pprn, uint8_t *data, size_t data_stride, size_t size,
                                  uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *positions,
                                  size_t pos_stride, const uint32_t *counts, uint8_t *ok, uint32_t *corrected_num);

/* RS decoding from caller-supplied syndromes: syndromes is [n][num_roots] uint16 in index (log)
 * form, 2^m-1 meaning zero (reference src/decode.c:446-464).  RS handles only. */
bool poporon_decode_batch_syndrome(struct _poporon_t *pprn, uint8_t *data, size_t data_stride, size_t size,
                                   uint8_t *parity, size_t parity_stride, size_t n, const uint16_t *syndromes,
                                   uint8_t *ok, uint32_t *corrected_num);

/* Syndromes only: the stage the reference keeps internal (calculate_syndrome_u8, src/decode.c:232-416).  Writes
 * syndromes[i * num_roots + j] = index (log) form of S_j of codeword i, 2^m-1 meaning zero - exactly the array a
 * handle configured with an external syndrome pointer, or poporon_decode_batch_syndrome, takes back.  Nothing is
 * corrected.  1 <= size <= 2^m - 1 - num_roots.  RS handles only; host or device pointers as above. */
bool poporon_syndromes_batch(struct _poporon_t *pprn, const uint8_t *data, size_t data_stride, size_t size,
                             const uint8_t *parity, size_t parity_stride, size_t n, uint16_t *syndromes);

/* Device-pointer forms that enqueue on a caller-owned cudaStream_t (passed as void*) and return
 * without synchronising.  All pointers must be device memory on the same device. */
bool poporon_encode_batch_async(struct _poporon_t *pprn, const uint8_t *data, size_t data_stride, size_t size,
                                uint8_t *parity, size_t parity_stride, size_t n, void *stream);
bool poporon_decode_batch_async(struct _poporon_t *pprn, uint8_t *data, size_t data_stride, size_t size,
                                uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *positions,
                                size_t pos_stride, const uint32_t *counts, uint8_t *ok, uint32_t *corrected_num,
                                void *stream);

/* kernels launched on behalf of this handle so far (all devices) */
uint64_t poporon_kernel_launches(const struct _poporon_t *pprn);
/* Opt-in per-kernel timing (RS handles; CUDA events around every kernel launched for this handle).
 * profile(h, true) starts/restarts accumulation on the devices the handle has touched so far;
 * profile_read returns the kernel's name (NULL past the last id), its launches and summed milliseconds. */
bool poporon_kernel_profile(struct _poporon_t *pprn, bool on);
const char *poporon_kernel_profile_read(struct _poporon_t *pprn, int kernel, uint64_t *launches, double *total_ms);
/* text of the last device-layer failure on this thread ("" if none) */
const char *poporon_last_error(void);

#ifdef __cplusplus
}
#endif

#endif /* POPORON_BATCH_H */
