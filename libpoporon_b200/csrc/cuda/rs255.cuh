// rs255.cuh — the RS(255,223) / 0x11D / fcr=1 / prim=1 fast path (bitsliced kernels, rs255_*.cu).
#pragma once
#include "common.cuh"

namespace ppcu {
// true when the described code is exactly the reference's default RS config (src/poporon.c:281-284)
bool rs255_matches(const poporon_cuda_rs_desc_t *d);
// shapes the fast path accepts (everything else falls to the generic kernels)
bool rs255_shape_ok(const poporon_cuda_rs *ctx, int op, const uint8_t *data, size_t data_stride, size_t size,
                    const uint8_t *parity, size_t parity_stride);
int rs255_init(poporon_cuda_rs *ctx);
int rs255_encode(poporon_cuda_rs *ctx, const uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                 size_t parity_stride, size_t n, uint8_t *ok, cudaStream_t stream);
int rs255_decode(poporon_cuda_rs *ctx, int mode, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                 size_t parity_stride, size_t n, const uint32_t *pos, size_t pos_stride, const uint32_t *cnt,
                 const uint16_t *extsyn, uint8_t *ok, uint32_t *corrected, cudaStream_t stream);
// the 32 index-form syndromes of every codeword (255 = zero), [n][32] uint16; POPORON_CUDA_ERR_UNSUPP for shapes the
// transform kernels do not take
int rs255_syndromes(poporon_cuda_rs *ctx, const uint8_t *data, size_t data_stride, size_t size, const uint8_t *parity,
                    size_t parity_stride, size_t n, uint16_t *out, cudaStream_t stream);
} // namespace ppcu
