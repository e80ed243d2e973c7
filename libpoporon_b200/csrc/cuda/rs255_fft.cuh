// rs255_fft.cuh — RS(255,223) syndromes / parity as a 15 x 17 transform on bit-planes (rs255_fft.cu).
#pragma once
#include "common.cuh"

namespace ppcu {
struct FftShape {
    int npar;     // parity bytes per codeword (32, 16 or 8)
    int packed;   // SCREEN: parity rows interleaved with the message rows (one stride): one span per sub-tile
    int prow_off; // packed: parity - data
    int bufbytes; // shared-memory bytes of one 32-codeword sub-tile
    int poff;     // offset of the parity span in a sub-tile buffer (split layout)
};
// op 0 = encode (message rows only), 1 = screen (message + parity rows).  Any base alignment, any strides whose
// 32-row span fits a sub-tile buffer.
bool rs255_fft_shape_ok(int npar, int op, const uint8_t *data, size_t data_stride, size_t size, const uint8_t *parity,
                        size_t parity_stride, FftShape *shape);
size_t rs255_fft_slab_bytes(int sm_count);
int rs255_fft_init(poporon_cuda_rs *ctx);
// dirty: bitmap over codewords (bit cw & 31 of word cw >> 5); syn: [n][32], rows of dirty codewords only
int rs255_fft_screen(poporon_cuda_rs *ctx, const FftShape &shape, const uint8_t *data, size_t data_stride, size_t size,
                     const uint8_t *parity, size_t parity_stride, size_t n, uint8_t *slab, uint32_t *dirty,
                     uint8_t *syn, cudaStream_t stream, uint32_t *queue = nullptr, uint32_t *counter = nullptr,
                     uint8_t *ok = nullptr, uint32_t *corrected = nullptr, bool fused = false);
// fused = the kernel also does the zero-syndrome exit (ok / corrected of clean codewords, queue of dirty ones); needs
// ok and corrected 16-byte aligned (or NULL).  Otherwise rs255_fft_compact does it from the dirty bitmap.
int rs255_fft_compact(poporon_cuda_rs *ctx, const uint32_t *dirty, size_t n, uint32_t *queue, uint32_t *counter,
                      uint8_t *ok, uint32_t *corrected, cudaStream_t stream);
// synplanes: scratch of rs255_fft_synplane_bytes(n) bytes (the syndromes of message * x^32, tower planes)
size_t rs255_fft_synplane_bytes(size_t n);
int rs255_fft_encode(poporon_cuda_rs *ctx, const FftShape &shape, const uint8_t *data, size_t data_stride, size_t size,
                     uint8_t *parity, size_t parity_stride, size_t n, uint8_t *slab, uint32_t *synplanes,
                     cudaStream_t stream);
} // namespace ppcu
