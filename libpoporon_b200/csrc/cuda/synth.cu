// synth.cu — deterministic synthetic workload generators (bench / test harness, not the codec).
//
// Counter-based: every byte is a pure function of (seed, codeword index, byte index), so any slice
// of a workload can be regenerated anywhere (each GPU of a sharded run fills its own slice; the CPU
// baseline receives a device-to-host copy of a sample).  SURVEY §8(d) configs 2, 3 and 5.
// Built into its own library (libpoporon_synth.so): the product library libpoporon_cuda.so holds the codec only.
#include "common.cuh"
#include "synth.h"

namespace ppcu {

namespace {

// (this library does not share the product library's error text)
int scheck(cudaError_t e, const char *what)
{
    (void)what;
    return e == cudaSuccess ? 0 : POPORON_CUDA_ERR_DEVICE;
}

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__host__ __device__ __forceinline__ uint64_t hash3(uint64_t seed, uint64_t i, uint64_t j)
{
    return mix64(mix64(seed ^ (i * 0xD6E8FEB86659FD93ull)) + j);
}

// One thread per (codeword, 8-byte word).
__global__ void synth_fill_kernel(uint8_t *data, size_t stride, size_t size, size_t n, uint64_t seed, uint64_t i0)
{
    const size_t words = (size + 7) / 8;
    const size_t total = n * words;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x) {
        const size_t cw = g / words, w = g % words;
        uint64_t h = hash3(seed, i0 + cw, w);
        uint8_t *p = data + cw * stride + w * 8;
        const size_t lim = size - w * 8 < 8 ? size - w * 8 : 8;
        for (size_t b = 0; b < lim; b++, h >>= 8)
            p[b] = (uint8_t)h;
    }
}

// One thread per codeword: picks distinct positions with a 256-bit occupancy map.
__global__ void synth_inject_kernel(uint8_t *cw, size_t stride, size_t len, size_t n, uint64_t seed, uint64_t i0,
                                    int kind, uint32_t emin, uint32_t emax, uint32_t p_ppm, uint32_t symmask)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    uint8_t *p = cw + i * stride;
    if (kind == 1) {
        for (size_t j = 0; j < len; j++) {
            const uint64_t h = hash3(seed, i0 + i, j);
            if ((uint32_t)(h % 1000000ull) < p_ppm)
                p[j] ^= (uint8_t)(1u + (uint32_t)((h >> 32) % symmask));
        }
        return;
    }
    uint32_t taken[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t e = emin + (uint32_t)(hash3(seed, i0 + i, 0xE0000000ull) % (uint64_t)(emax - emin + 1u));
    if (e > len)
        e = (uint32_t)len;
    uint64_t ctr = 0;
    for (uint32_t k = 0; k < e; k++) {
        uint32_t pos;
        uint64_t h;
        do {
            h = hash3(seed, i0 + i, ctr++);
            pos = (uint32_t)(h % len);
        } while (taken[pos >> 5] & (1u << (pos & 31)));
        taken[pos >> 5] |= 1u << (pos & 31);
        p[pos] ^= (uint8_t)(1u + (uint32_t)((h >> 32) % symmask));
    }
}

// Erasure workload (config 3): c ~ U{0..cmax} erased positions in [0,size), written ascending into a
// zero-filled row of `pos_stride` slots; erased symbols get a random magnitude (zero allowed); plus
// v in {0,1,2,3} unknown errors (p = .90/.06/.03/.01) anywhere in the codeword.
__global__ void synth_erasures_kernel(uint8_t *cw, size_t stride, size_t size, size_t len, size_t n, uint64_t seed,
                                      uint64_t i0, uint32_t cmax, uint32_t *pos, size_t pos_stride, uint32_t *cnt,
                                      uint32_t symmask)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    uint8_t *p = cw + i * stride;
    uint32_t taken[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t c = (uint32_t)(hash3(seed, i0 + i, 0xE0000000ull) % (uint64_t)(cmax + 1u));
    if (c > size)
        c = (uint32_t)size;
    uint64_t ctr = 0;
    for (uint32_t k = 0; k < c; k++) {
        uint32_t q;
        uint64_t h;
        do {
            h = hash3(seed, i0 + i, ctr++);
            q = (uint32_t)(h % size);
        } while (taken[q >> 5] & (1u << (q & 31)));
        taken[q >> 5] |= 1u << (q & 31);
        p[q] ^= (uint8_t)((h >> 32) % (symmask + 1u)); // magnitude 0 allowed
    }
    uint32_t *row = pos + i * pos_stride;
    uint32_t w = 0;
    for (uint32_t q = 0; q < size && w < pos_stride; q++)
        if (taken[q >> 5] & (1u << (q & 31)))
            row[w++] = q;
    for (; w < pos_stride; w++)
        row[w] = 0;
    cnt[i] = c;
    const uint32_t u = (uint32_t)(hash3(seed, i0 + i, 0xE0000001ull) % 100ull);
    const uint32_t v = u < 90 ? 0 : (u < 96 ? 1 : (u < 99 ? 2 : 3));
    for (uint32_t k = 0; k < v; k++) {
        uint32_t q;
        uint64_t h;
        do {
            h = hash3(seed, i0 + i, ctr++);
            q = (uint32_t)(h % len);
        } while (taken[q >> 5] & (1u << (q & 31)));
        taken[q >> 5] |= 1u << (q & 31);
        p[q] ^= (uint8_t)(1u + (uint32_t)((h >> 32) % symmask));
    }
}

} // namespace

} // namespace ppcu

using namespace ppcu;

extern "C" int poporon_cuda_synth_fill(int device, uint8_t *data, size_t stride, size_t size, size_t n, uint64_t seed,
                                       uint64_t i0, void *stream)
{
    DeviceGuard g(device);
    if (!g.ok)
        return scheck(cudaErrorInvalidDevice, "set device");
    if (n == 0 || size == 0)
        return 0;
    const size_t total = n * ((size + 7) / 8);
    size_t blocks = (total + 255) / 256;
    const size_t cap = (size_t)sm_count_of(device) * 32;
    if (blocks > cap)
        blocks = cap;
    synth_fill_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(data, stride, size, n, seed, i0);
    return scheck(cudaGetLastError(), "synth_fill launch");
}

extern "C" int poporon_cuda_synth_inject(int device, uint8_t *cw, size_t stride, size_t len, size_t n, uint64_t seed,
                                         uint64_t i0, int kind, uint32_t emin, uint32_t emax, uint32_t p_ppm,
                                         void *stream)
{
    DeviceGuard g(device);
    if (!g.ok)
        return scheck(cudaErrorInvalidDevice, "set device");
    if (len == 0 || len > 256 || emax < emin)
        return POPORON_CUDA_ERR_ARG;
    if (n == 0)
        return 0;
    synth_inject_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(cw, stride, len, n, seed, i0,
                                                                                     kind, emin, emax, p_ppm, 255u);
    return scheck(cudaGetLastError(), "synth_inject launch");
}

extern "C" int poporon_cuda_synth_erasures(int device, uint8_t *cw, size_t stride, size_t size, size_t len, size_t n,
                                           uint64_t seed, uint64_t i0, uint32_t cmax, uint32_t *pos,
                                           size_t pos_stride, uint32_t *cnt, void *stream)
{
    DeviceGuard g(device);
    if (!g.ok)
        return scheck(cudaErrorInvalidDevice, "set device");
    if (len == 0 || len > 256 || size > len || size == 0)
        return POPORON_CUDA_ERR_ARG;
    if (n == 0)
        return 0;
    synth_erasures_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        cw, stride, size, len, n, seed, i0, cmax, pos, pos_stride, cnt, 255u);
    return scheck(cudaGetLastError(), "synth_erasures launch");
}
