// noalloc.cu — poporon_cuda_rs_encode_noalloc_dev: the bitsliced RS(255,k) encoder with its scratch in a file-scope
// device array instead of cudaMalloc'd context memory (see include/poporon_cuda.h).
//
// The encoder is three kernels (rs255_fft.cu): message rows -> syndrome planes, syndrome planes -> parity planes,
// parity planes -> parity rows.  Between them sit two pieces of scratch that rs255_encode (rs255.cu) allocates on
// first use: one 255 KB bit-plane slab per CTA of the first kernel, and 64 bytes of syndrome / parity planes per
// codeword.  Here both live in g_noalloc_scratch, sized for POPORON_CUDA_NOALLOC_MAX_CODEWORDS codewords per call:
//
//   [ kMaxTiles slabs of 255 KB | syndrome planes of kMaxTiles tiles | parity planes of kMaxTiles tiles ]
//
// (a call of n codewords launches min(tiles, SMs) <= kMaxTiles CTAs of the first kernel, one slab each).  Nothing in
// it has to start at zero: every kernel writes the words it later reads (the allocating path hands the same kernels
// uninitialised cudaMalloc memory), so the previous call's contents are simply overwritten.  All calls on one device
// share the array whatever context or stream they use: calls that overlap on the GPU race.
#include "common.cuh"
#include "rs255_fft.cuh"

#include <mutex>

namespace {

constexpr size_t kTileCodewords = 1024;
constexpr size_t kMaxTiles = POPORON_CUDA_NOALLOC_MAX_CODEWORDS / kTileCodewords;
constexpr size_t kSlabBytesPerCta = 255 * 1024;             // rs255_fft.cu kSlabBytes
constexpr size_t kPlaneBytesPerTile = 32 * 32 * 8 * 4;      // rs255_fft.cu kTilePlaneWords * 4
constexpr size_t kSlabArea = kMaxTiles * kSlabBytesPerCta;  // 16 711 680
constexpr size_t kScratchBytes = kSlabArea + 2 * kMaxTiles * kPlaneBytesPerTile; // + 4 194 304 = 20 905 984
static_assert(POPORON_CUDA_NOALLOC_MAX_CODEWORDS % kTileCodewords == 0, "whole tiles");
static_assert(kSlabArea % 256 == 0, "plane area keeps the array's alignment");

// Held by every GPU that loads the library, from module load until the process exits.
static __device__ __align__(256) uint8_t g_noalloc_scratch[kScratchBytes];

// In host code g_noalloc_scratch is only a handle for the runtime's symbol table, not a device address: the kernels
// get the address cudaGetSymbolAddress returns for the current device, looked up once per device.
constexpr int kMaxDevices = 64;
std::once_flag g_once[kMaxDevices];
uint8_t *g_scratch_ptr[kMaxDevices];
cudaError_t g_scratch_err[kMaxDevices];

// device must be current
int scratch_of(int device, uint8_t **out)
{
    if (device < 0 || device >= kMaxDevices) {
        ppcu::set_error("noalloc: device ordinal %d outside 0..%d", device, kMaxDevices - 1);
        return POPORON_CUDA_ERR_UNSUPP;
    }
    std::call_once(g_once[device], [device] {
        void *p = nullptr;
        g_scratch_err[device] = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        g_scratch_ptr[device] = g_scratch_err[device] == cudaSuccess ? (uint8_t *)p : nullptr;
    });
    if (g_scratch_err[device] != cudaSuccess || !g_scratch_ptr[device])
        return ppcu::check(g_scratch_err[device] != cudaSuccess ? g_scratch_err[device] : cudaErrorInvalidSymbol,
                           "cudaGetSymbolAddress(g_noalloc_scratch)");
    *out = g_scratch_ptr[device];
    return 0;
}

} // namespace

extern "C" int poporon_cuda_rs_encode_noalloc_dev(poporon_cuda_rs_t *ctx, const uint8_t *data, size_t data_stride,
                                                  size_t size, uint8_t *parity, size_t parity_stride, size_t n,
                                                  uint8_t *ok, void *stream)
{
    using namespace ppcu;
    if (!ctx || (n && (!data || !parity)))
        return POPORON_CUDA_ERR_ARG;
    if (n > (size_t)POPORON_CUDA_NOALLOC_MAX_CODEWORDS) {
        set_error("poporon_cuda_rs_encode_noalloc_dev: n = %zu codewords exceeds the static scratch "
                  "(POPORON_CUDA_NOALLOC_MAX_CODEWORDS = %d per call); split the batch or use "
                  "poporon_cuda_rs_encode_dev",
                  n, POPORON_CUDA_NOALLOC_MAX_CODEWORDS);
        return POPORON_CUDA_ERR_UNSUPP;
    }
    FftShape shape;
    if (!ctx->fast255 ||
        !rs255_fft_shape_ok(ctx->code.nroots, 0, data, data_stride, size, parity, parity_stride, &shape)) {
        set_error("poporon_cuda_rs_encode_noalloc_dev: only the bitsliced RS(255,223/239/247) encoder has a "
                  "static-scratch variant (code or row layout outside it)");
        return POPORON_CUDA_ERR_UNSUPP;
    }
    if (n == 0)
        return 0;
    DeviceGuard guard(ctx->device);
    if (!guard.ok)
        return check(cudaErrorInvalidDevice, "cudaSetDevice");
    uint8_t *scratch = nullptr;
    int rc = scratch_of(ctx->device, &scratch);
    if (rc != 0)
        return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (ok)
        PP_CUDA_TRY(cudaMemsetAsync(ok, 1, n, s)); // the encoder cannot fail (reference src/encode.c:17-145 returns true)
    // rs255_fft_encode places the parity planes right after the syndrome planes of THIS call's tiles, so both fit the
    // 2 * kMaxTiles tiles reserved behind the slabs
    rc = rs255_fft_encode(ctx, shape, data, data_stride, size, parity, parity_stride, n, scratch,
                          (uint32_t *)(scratch + kSlabArea), s);
    ctx->launches += 3;
    return rc;
}
