// rs_generic.cuh — launchers of the general-parameter RS kernels (rs_generic.cu).
#pragma once
#include "common.cuh"

namespace ppcu {
int launch_rs_encode_generic(poporon_cuda_rs *ctx, const uint8_t *data, size_t data_stride, size_t size,
                             uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, cudaStream_t stream);
int launch_rs_decode_generic(poporon_cuda_rs *ctx, int mode, uint8_t *data, size_t data_stride, size_t size,
                             uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *pos, size_t pos_stride,
                             const uint32_t *cnt, const uint16_t *syn, uint8_t *ok, uint32_t *corrected,
                             cudaStream_t stream, uint16_t *synout = nullptr);
// synout != NULL: the call only computes the syndromes of every codeword (index form, [n][nroots]); nothing is corrected
} // namespace ppcu
