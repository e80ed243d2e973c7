// rs255.cu — placeholder until the bitsliced kernels land: reports "no fast path".
#include "rs255.cuh"

namespace ppcu {
bool rs255_matches(const poporon_cuda_rs_desc_t *) { return false; }
bool rs255_shape_ok(size_t, size_t, size_t) { return false; }
int rs255_init(poporon_cuda_rs *) { return 0; }
int rs255_encode(poporon_cuda_rs *, const uint8_t *, size_t, size_t, uint8_t *, size_t, size_t, uint8_t *,
                 cudaStream_t)
{
    return POPORON_CUDA_ERR_UNSUPP;
}
int rs255_decode(poporon_cuda_rs *, int, uint8_t *, size_t, size_t, uint8_t *, size_t, size_t, const uint32_t *,
                 size_t, const uint32_t *, uint8_t *, uint32_t *, cudaStream_t)
{
    return POPORON_CUDA_ERR_UNSUPP;
}
} // namespace ppcu
