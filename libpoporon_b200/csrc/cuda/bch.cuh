// bch.cuh — launchers of the binary BCH kernels (bch.cu).
#pragma once
#include "common.cuh"

namespace ppcu {
int launch_bch_encode(poporon_cuda_bch *ctx, const uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                      size_t parity_stride, size_t n, uint8_t *ok, cudaStream_t stream);
int launch_bch_decode(poporon_cuda_bch *ctx, uint8_t *data, size_t data_stride, size_t size, const uint8_t *parity,
                      size_t parity_stride, size_t n, uint8_t *ok, uint32_t *corrected, cudaStream_t stream);
// n <= 15 only: tabulates the exact decoder over all 2^n received words (synchronous, once per context)
int bch_build_decode_map(poporon_cuda_bch *ctx);
} // namespace ppcu
