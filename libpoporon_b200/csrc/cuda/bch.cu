// bch.cu — binary BCH(n <= 31) encode/decode kernels, one thread per codeword.
//
// The reference keeps the whole codeword in a uint32_t (src/bch.c:352-435) and wraps it in
// big-endian byte packing (src/encode.c:199-234, src/decode.c:542-594).  A codeword is 1-4 data
// bytes + 1-4 parity bytes, so these kernels are pure ALU work on registers; field tables sit in
// shared memory (512 B per CTA).
//
// Decode mirrors the reference step by step so that miscorrections of > t errors and the
// "found != L" / re-syndrome rejections come out identical:
//   syndromes  src/bch.c:25-51     BM  src/bch.c:78-142    Chien  src/bch.c:144-166
//   accept/reject rules  src/bch.c:403-426
#include "common.cuh"
#include "bch.cuh"

namespace ppcu {

namespace {

constexpr int kArr = 64; // fixed work-array length of the reference (src/bch.c:12)

struct BchParams {
    int m, nn, n, k, parity_bits, t;
    uint32_t gen;
    int gen_deg;
};

struct BchSmem {
    uint8_t alpha_to[256];
    uint8_t index_of[256];
};

__device__ __forceinline__ void bch_load_tables(BchSmem &s, const uint8_t *g)
{
    const uint32_t *src = reinterpret_cast<const uint32_t *>(g);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&s);
    for (int i = threadIdx.x; i < 128; i += blockDim.x)
        dst[i] = src[i];
    __syncthreads();
}

__device__ __forceinline__ uint32_t pack_be(const uint8_t *p, int nbytes)
{
    uint32_t v = 0;
    for (int i = 0; i < nbytes && i < 4; i++)
        v |= (uint32_t)p[i] << (8 * (nbytes - 1 - i));
    return v;
}

__device__ __forceinline__ uint32_t low_mask(int bits)
{
    return bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1u);
}

// rem[b][v] = (v * x^(8b + p)) mod g for data byte b (b = 0 the least significant byte of the k data bits): the
// remainder of d * x^p by g is the XOR of four look-ups (division by g is linear over GF(2)), src/bch.c:352-380
__device__ __forceinline__ void bch_build_rem_tables(const BchParams &P, uint32_t (*rem)[256])
{
    for (int e = threadIdx.x; e < 1024; e += blockDim.x) {
        const int b = e >> 8;
        uint32_t r = 0;
        if (8 * b < P.k) {
            uint64_t w = ((uint64_t)(((uint32_t)(e & 255) << (8 * b)) & low_mask(P.k))) << P.parity_bits;
            for (int i = P.k + P.parity_bits - 1; i >= P.gen_deg; i--)
                if (w & (1ull << i))
                    w ^= (uint64_t)P.gen << (i - P.gen_deg);
            r = (uint32_t)w;
        }
        rem[b][e & 255] = r;
    }
    __syncthreads();
}
__device__ __forceinline__ uint32_t bch_rem(const uint32_t (*rem)[256], uint32_t dv)
{
    return rem[0][dv & 255] ^ rem[1][(dv >> 8) & 255] ^ rem[2][(dv >> 16) & 255] ^ rem[3][dv >> 24];
}

__global__ void __launch_bounds__(256)
bch_encode_kernel(BchParams P, const uint8_t *__restrict__ data, size_t data_stride, size_t size,
                  uint8_t *__restrict__ parity, size_t parity_stride, size_t n, uint8_t *__restrict__ ok)
{
    __shared__ uint32_t rem[4][256];
    bch_build_rem_tables(P, rem);
    const int db = (P.k + 7) / 8, pb = (P.parity_bits + 7) / 8;
    for (size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x; cw < n; cw += (size_t)gridDim.x * blockDim.x) {
        bool good = false;
        if (size >= (size_t)db) {
            const uint32_t d = pack_be(data + cw * data_stride, db) & low_mask(P.k);
            const uint32_t pv = bch_rem(rem, d); // systematic encoding: remainder of d * x^p by g
            uint8_t *p = parity + cw * parity_stride;
            for (int i = 0; i < pb; i++)
                p[i] = 0;
            for (int i = 0; i < pb && i < 4; i++)
                p[pb - 1 - i] = (uint8_t)(pv >> (8 * i));
            good = true;
        }
        if (ok)
            ok[cw] = good ? 1 : 0;
    }
}

__device__ bool bch_syndromes(const BchParams &P, const BchSmem &T, uint32_t w, uint8_t *s)
{
    bool any = false;
    for (int i = 0; i < 2 * P.t; i++) {
        uint8_t acc = 0;
        uint32_t bits = w;
        while (bits) {
            const int j = __ffs(bits) - 1;
            bits &= bits - 1;
            acc ^= T.alpha_to[((i + 1) * j) % P.nn];
        }
        s[i] = acc;
        any |= (acc != 0);
    }
    return any;
}

__device__ bool bch_decode_word(const BchParams &P, const BchSmem &T, uint32_t received, uint32_t &fixed_out,
                                int &num_errors)
{
    uint8_t s[kArr], cur[kArr], prev[kArr], keep[kArr];
    received &= low_mask(P.n);
    fixed_out = received;
    num_errors = 0;
    for (int i = 0; i < kArr; i++)
        s[i] = 0;
    if (!bch_syndromes(P, T, received, s))
        return true;

    for (int i = 0; i < kArr; i++)
        cur[i] = prev[i] = 0;
    cur[0] = prev[0] = 1;
    int L = 0, shift = 1;
    uint8_t dprev = 1;
    for (int it = 0; it < 2 * P.t; it++) {
        uint8_t d = s[it];
        for (int i = 1; i <= L; i++)
            if (cur[i] != 0 && s[it - i] != 0)
                d ^= T.alpha_to[(T.index_of[cur[i]] + T.index_of[s[it - i]]) % P.nn];
        if (d == 0) {
            shift++;
            continue;
        }
        const uint8_t mult = T.alpha_to[(P.nn - T.index_of[dprev] + T.index_of[d]) % P.nn];
        const bool grow = (2 * L <= it);
        if (grow)
            for (int i = 0; i < kArr; i++)
                keep[i] = cur[i];
        for (int i = 0; i < kArr - shift; i++)
            if (prev[i] != 0)
                cur[i + shift] ^= T.alpha_to[(T.index_of[prev[i]] + T.index_of[mult]) % P.nn];
        if (grow) {
            for (int i = 0; i < kArr; i++)
                prev[i] = keep[i];
            L = it + 1 - L;
            dprev = d;
            shift = 1;
        } else {
            shift++;
        }
    }
    if (L > P.t)
        return false;

    uint32_t flips = 0;
    int found = 0;
    for (int i = 0; i < P.n; i++) {
        // sigma evaluated at alpha^{-i}; position i is in error when it vanishes
        const int lx = T.index_of[T.alpha_to[(P.nn - i) % P.nn]];
        uint8_t acc = 0;
        for (int j = 0; j <= L; j++)
            if (cur[j] != 0)
                acc ^= T.alpha_to[(T.index_of[cur[j]] + (lx * j) % P.nn) % P.nn];
        if (acc == 0) {
            flips ^= 1u << i;
            if (++found >= L)
                break;
        }
    }
    if (found != L)
        return false;
    const uint32_t fixed = received ^ flips;
    if (bch_syndromes(P, T, fixed, s))
        return false;
    fixed_out = fixed;
    num_errors = found;
    return true;
}

__global__ void __launch_bounds__(128)
bch_decode_kernel(BchParams P, const uint8_t *__restrict__ gtab, uint8_t *data, size_t data_stride, size_t size,
                  const uint8_t *__restrict__ parity, size_t parity_stride, size_t n, uint8_t *__restrict__ ok,
                  uint32_t *__restrict__ corrected)
{
    __shared__ BchSmem T;
    bch_load_tables(T, gtab);
    const size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (cw >= n)
        return;
    const int db = (P.k + 7) / 8, pb = (P.parity_bits + 7) / 8;
    bool good = false;
    uint32_t nc = 0;
    if (size != 0 && size >= (size_t)db) {
        uint8_t *d = data + cw * data_stride;
        const uint32_t dv = pack_be(d, db) & low_mask(P.k);
        const uint32_t pv = pack_be(parity + cw * parity_stride, pb) & low_mask(P.parity_bits);
        uint32_t fixed;
        int ne;
        if (bch_decode_word(P, T, (dv << P.parity_bits) | pv, fixed, ne)) {
            const uint32_t out = (fixed >> P.parity_bits) & low_mask(P.k);
            for (int i = 0; i < db && i < 4; i++)
                d[db - 1 - i] = (uint8_t)(out >> (8 * i));
            nc = ne > 0 ? (uint32_t)ne : 0;
            good = true;
        }
    }
    if (ok)
        ok[cw] = good ? 1 : 0;
    if (corrected)
        corrected[cw] = nc;
}

// ---- complete decode map for short codes (n <= 15) ------------------------------------------------------------
// The outcome of poporon_bch_decode (src/bch.c:382-435) is a pure function of the n received bits, so for n <= 15
// the exact step-by-step decoder above is run ONCE over all 2^n words when the context is created and batches are
// decoded by look-up: word = fixed codeword | num_errors << 24 | ok << 31.  The 128 KB table stays in L1/L2, the
// batch kernel is then bound by its 9 bytes per codeword of HBM traffic instead of ~900 table-driven field
// operations per codeword.  Bit-exact by construction (every entry IS the reference-shaped decoder's answer).
constexpr int kMapMaxN = 15;

__global__ void __launch_bounds__(128)
bch_build_map_kernel(BchParams P, const uint8_t *__restrict__ gtab, uint32_t *__restrict__ map)
{
    __shared__ BchSmem T;
    bch_load_tables(T, gtab);
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= (1u << P.n))
        return;
    uint32_t fixed;
    int ne;
    const bool good = bch_decode_word(P, T, w, fixed, ne);
    map[w] = good ? (fixed | ((uint32_t)(ne > 0 ? ne : 0) << 24) | 0x80000000u) : 0u;
}

__global__ void __launch_bounds__(256)
bch_decode_map_kernel(BchParams P, const uint32_t *__restrict__ map, uint8_t *data, size_t data_stride, size_t size,
                      const uint8_t *__restrict__ parity, size_t parity_stride, size_t n, uint8_t *__restrict__ ok,
                      uint32_t *__restrict__ corrected)
{
    const int db = (P.k + 7) / 8, pb = (P.parity_bits + 7) / 8;
    const bool sized = size != 0 && size >= (size_t)db;
    for (size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x; cw < n; cw += (size_t)gridDim.x * blockDim.x) {
        bool good = false;
        uint32_t nc = 0;
        if (sized) {
            uint8_t *d = data + cw * data_stride;
            const uint32_t dv = pack_be(d, db) & low_mask(P.k);
            const uint32_t pv = pack_be(parity + cw * parity_stride, pb) & low_mask(P.parity_bits);
            const uint32_t e = __ldg(map + ((dv << P.parity_bits) | pv));
            if (e & 0x80000000u) {
                const uint32_t out = ((e & 0xFFFFu) >> P.parity_bits) & low_mask(P.k);
                for (int i = 0; i < db && i < 4; i++)
                    d[db - 1 - i] = (uint8_t)(out >> (8 * i));
                nc = (e >> 24) & 0x7Fu;
                good = true;
            }
        }
        if (ok)
            ok[cw] = good ? 1 : 0;
        if (corrected)
            corrected[cw] = nc;
    }
}

// ---- decode map indexed by the remainder, for 15 < n <= 31 with at most kSmapMaxBits parity bits ---------------
// Everything the decoder of src/bch.c:382-435 decides - the syndromes S_i = r(alpha^i), the locator, the roots, the
// re-check - depends on the received word r only through r mod g (g vanishes at alpha^1 .. alpha^2t, so r and
// s = r mod g have the same syndromes), and it answers "r with these bit positions flipped".  The exact decoder is
// therefore run once per context on the 2^p words w = s (data bits zero), and a batch is decoded as
//   s = (data * x^p mod g) ^ parity      (the encoder's remainder, byte tables in shared memory)
//   entry = smap[s]:  bit 31 ok, bits 26..29 num_errors, bits 0..25 the flips that fall on data bits
// BCH(31,16) decode of 64 Mi codewords: 66 ms with the per-thread decoder.
constexpr int kSmapMaxBits = 25; // BCH(31,6) t = 7: 2^25 entries = 128 MB

__global__ void __launch_bounds__(128)
bch_build_smap_kernel(BchParams P, const uint8_t *__restrict__ gtab, uint32_t *__restrict__ smap)
{
    __shared__ BchSmem T;
    bch_load_tables(T, gtab);
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= (1u << P.parity_bits))
        return;
    uint32_t fixed;
    int ne;
    const bool good = bch_decode_word(P, T, s, fixed, ne);
    smap[s] = good ? ((((fixed ^ s) >> P.parity_bits) & low_mask(P.k)) | ((uint32_t)(ne > 0 ? ne : 0) << 26) | 0x80000000u)
                   : 0u;
}

__global__ void __launch_bounds__(256)
bch_decode_smap_kernel(BchParams P, const uint32_t *__restrict__ smap, uint8_t *data, size_t data_stride, size_t size,
                       const uint8_t *__restrict__ parity, size_t parity_stride, size_t n, uint8_t *__restrict__ ok,
                       uint32_t *__restrict__ corrected)
{
    __shared__ uint32_t rem[4][256];
    bch_build_rem_tables(P, rem);
    const int db = (P.k + 7) / 8, pb = (P.parity_bits + 7) / 8;
    const bool sized = size != 0 && size >= (size_t)db;
    for (size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x; cw < n; cw += (size_t)gridDim.x * blockDim.x) {
        bool good = false;
        uint32_t nc = 0;
        if (sized) {
            uint8_t *d = data + cw * data_stride;
            const uint32_t dv = pack_be(d, db) & low_mask(P.k);
            const uint32_t pv = pack_be(parity + cw * parity_stride, pb) & low_mask(P.parity_bits);
            const uint32_t s = bch_rem(rem, dv) ^ pv;
            const uint32_t e = __ldg(smap + s);
            if (e & 0x80000000u) {
                const uint32_t out = dv ^ (e & low_mask(P.k));
                for (int i = 0; i < db && i < 4; i++)
                    d[db - 1 - i] = (uint8_t)(out >> (8 * i));
                nc = (e >> 26) & 15u;
                good = true;
            }
        }
        if (ok)
            ok[cw] = good ? 1 : 0;
        if (corrected)
            corrected[cw] = nc;
    }
}

// Dense batches of any n <= 31 code (data rows of db bytes and parity rows of pb bytes back to back, db, pb <= 4): four
// codewords per thread, everything moved as aligned 32-bit words (4 db bytes = db words, 4 pb bytes = pb words), status
// as one word and one uint4.  Byte j of codeword u of a group sits at byte u * w + j of the group's words.
template <int width>
__device__ __forceinline__ uint32_t group_be(const uint32_t (&w)[4], int u)
{
    uint32_t v = 0;
#pragma unroll
    for (int j = 0; j < width; j++) {
        const int b = u * width + j;
        v = (v << 8) | ((w[b >> 2] >> (8 * (b & 3))) & 0xFFu);
    }
    return v;
}
template <int width>
__device__ __forceinline__ void group_put_be(uint32_t (&w)[4], int u, uint32_t v)
{
#pragma unroll
    for (int j = 0; j < width; j++) {
        const int b = u * width + j, sh = 8 * (b & 3);
        w[b >> 2] = (w[b >> 2] & ~(0xFFu << sh)) | (((v >> (8 * (width - 1 - j))) & 0xFFu) << sh);
    }
}

template <int db, int pb> // bytes per data row / parity row (compile-time: every shift becomes a constant)
__global__ void __launch_bounds__(256)
bch_decode_smap4_kernel(BchParams P, const uint32_t *__restrict__ smap, uint32_t *data, const uint32_t *__restrict__ parity,
                        size_t n4, uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected)
{
    __shared__ uint32_t rem[4][256];
    bch_build_rem_tables(P, rem);
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += (size_t)gridDim.x * blockDim.x) {
        uint32_t dw[4] = {0, 0, 0, 0}, pw[4] = {0, 0, 0, 0}, nc[4], okw = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (i < db)
                dw[i] = data[g * db + i];
            if (i < pb)
                pw[i] = parity[g * pb + i];
        }
        uint32_t s[4], dv[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            dv[u] = group_be<db>(dw, u) & low_mask(P.k);
            s[u] = bch_rem(rem, dv[u]) ^ (group_be<pb>(pw, u) & low_mask(P.parity_bits));
        }
        uint32_t e[4];
#pragma unroll
        for (int u = 0; u < 4; u++)
            e[u] = __ldg(smap + s[u]);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const bool good = (e[u] & 0x80000000u) != 0;
            if (good) // failure leaves the data bytes alone; success also clears the pad bits
                group_put_be<db>(dw, u, dv[u] ^ (e[u] & low_mask(P.k)));
            okw |= (good ? 1u : 0u) << (8 * u);
            nc[u] = good ? ((e[u] >> 26) & 15u) : 0u;
        }
#pragma unroll
        for (int i = 0; i < 4; i++)
            if (i < db)
                data[g * db + i] = dw[i];
        if (ok)
            reinterpret_cast<uint32_t *>(ok)[g] = okw;
        if (corrected)
            reinterpret_cast<uint4 *>(corrected)[g] = make_uint4(nc[0], nc[1], nc[2], nc[3]);
    }
}

template <int db, int pb>
__global__ void __launch_bounds__(256)
bch_encode_dense4_kernel(BchParams P, const uint32_t *__restrict__ data, uint32_t *__restrict__ parity, size_t n4,
                         uint8_t *__restrict__ ok)
{
    __shared__ uint32_t rem[4][256];
    bch_build_rem_tables(P, rem);
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += (size_t)gridDim.x * blockDim.x) {
        uint32_t dw[4] = {0, 0, 0, 0}, pw[4] = {0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < 4; i++)
            if (i < db)
                dw[i] = data[g * db + i];
#pragma unroll
        for (int u = 0; u < 4; u++)
            group_put_be<pb>(pw, u, bch_rem(rem, group_be<db>(dw, u) & low_mask(P.k)));
#pragma unroll
        for (int i = 0; i < 4; i++)
            if (i < pb)
                parity[g * pb + i] = pw[i];
        if (ok)
            reinterpret_cast<uint32_t *>(ok)[g] = 0x01010101u;
    }
}

// Packed BCH(15,k) batches (1 data byte and 2 parity bytes per codeword, both arrays dense): four codewords per
// thread with 4-, 8- and 16-byte accesses, so that the kernel runs at memory speed (9 B of traffic per codeword).
__global__ void __launch_bounds__(256)
bch_decode_map4_kernel(BchParams P, const uint32_t *__restrict__ map, uint8_t *data, const uint8_t *__restrict__ parity,
                       size_t n4 /* groups of four */, uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected)
{
    const uint32_t kmask = low_mask(P.k), pmask = low_mask(P.parity_bits);
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += (size_t)gridDim.x * blockDim.x) {
        const uint32_t dw = reinterpret_cast<const uint32_t *>(data)[g];
        const uint2 pw = reinterpret_cast<const uint2 *>(parity)[g];
        uint32_t out = 0, okw = 0;
        uint32_t nc[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t dv = (dw >> (8 * u)) & 0xFFu & kmask;
            const uint32_t ph = u < 2 ? pw.x : pw.y; // two bytes per codeword, big endian
            const uint32_t pb2 = (ph >> (16 * (u & 1))) & 0xFFFFu;
            const uint32_t pv = (((pb2 & 0xFFu) << 8) | (pb2 >> 8)) & pmask;
            const uint32_t e = __ldg(map + ((dv << P.parity_bits) | pv));
            const bool good = (e & 0x80000000u) != 0;
            const uint32_t fixed = ((e & 0xFFFFu) >> P.parity_bits) & kmask;
            out |= (good ? fixed : ((dw >> (8 * u)) & 0xFFu)) << (8 * u); // failure leaves the data byte alone
            okw |= (good ? 1u : 0u) << (8 * u);
            nc[u] = good ? ((e >> 24) & 0x7Fu) : 0u;
        }
        reinterpret_cast<uint32_t *>(data)[g] = out;
        if (ok)
            reinterpret_cast<uint32_t *>(ok)[g] = okw;
        if (corrected)
            reinterpret_cast<uint4 *>(corrected)[g] = make_uint4(nc[0], nc[1], nc[2], nc[3]);
    }
}

__global__ void __launch_bounds__(256)
bch_encode4_kernel(BchParams P, const uint8_t *__restrict__ data, uint8_t *__restrict__ parity, size_t n4,
                   uint8_t *__restrict__ ok)
{
    // k <= 8 here: the 2^k parity words (already byte-swapped to big endian) are tabulated per CTA by the
    // reference-shaped division (src/bch.c:352-380), the batch is then one shared-memory look-up per codeword
    __shared__ uint16_t tab[256];
    const uint32_t kmask = low_mask(P.k), pmask = low_mask(P.parity_bits);
    if (threadIdx.x <= kmask) {
        const uint32_t shifted = (uint32_t)threadIdx.x << P.parity_bits;
        uint32_t rem = shifted;
        for (int i = P.n - 1; i >= P.gen_deg; i--)
            if (rem & (1u << i))
                rem ^= P.gen << (i - P.gen_deg);
        const uint32_t pv = (shifted ^ rem) & pmask;
        tab[threadIdx.x] = (uint16_t)(((pv & 0xFFu) << 8) | (pv >> 8));
    }
    __syncthreads();
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += (size_t)gridDim.x * blockDim.x) {
        const uint32_t dw = reinterpret_cast<const uint32_t *>(data)[g];
        const uint32_t p0 = tab[dw & kmask], p1 = tab[(dw >> 8) & kmask], p2 = tab[(dw >> 16) & kmask],
                       p3 = tab[(dw >> 24) & kmask];
        reinterpret_cast<uint2 *>(parity)[g] = make_uint2(p0 | (p1 << 16), p2 | (p3 << 16));
        if (ok)
            reinterpret_cast<uint32_t *>(ok)[g] = 0x01010101u;
    }
}

// one data byte + two parity bytes per codeword, dense arrays, vector-aligned pointers
static bool packed_1_2(const poporon_cuda_bch *ctx, const void *data, size_t data_stride, size_t size, const void *parity,
                       size_t parity_stride, const void *ok, const void *corrected)
{
    const int db = (ctx->k + 7) / 8, pb = (ctx->parity_bits + 7) / 8;
    return db == 1 && pb == 2 && size >= 1 && data_stride == 1 && parity_stride == 2 && ((uintptr_t)data & 3u) == 0 &&
           ((uintptr_t)parity & 7u) == 0 && ((uintptr_t)ok & 3u) == 0 && ((uintptr_t)corrected & 15u) == 0;
}

// (db, pb) of the n <= 31 codes: k + p = n <= 31 bits
#define PP_BCH_DENSE_DISPATCH(KERNEL, db, pb, ...)                                          \
    do {                                                                                    \
        switch ((int)(db) * 8 + (int)(pb)) {                                                \
        case 1 * 8 + 1: KERNEL<1, 1> __VA_ARGS__; break;                                    \
        case 1 * 8 + 2: KERNEL<1, 2> __VA_ARGS__; break;                                    \
        case 1 * 8 + 3: KERNEL<1, 3> __VA_ARGS__; break;                                    \
        case 1 * 8 + 4: KERNEL<1, 4> __VA_ARGS__; break;                                    \
        case 2 * 8 + 1: KERNEL<2, 1> __VA_ARGS__; break;                                    \
        case 2 * 8 + 2: KERNEL<2, 2> __VA_ARGS__; break;                                    \
        case 2 * 8 + 3: KERNEL<2, 3> __VA_ARGS__; break;                                    \
        case 3 * 8 + 1: KERNEL<3, 1> __VA_ARGS__; break;                                    \
        case 3 * 8 + 2: KERNEL<3, 2> __VA_ARGS__; break;                                    \
        case 4 * 8 + 1: KERNEL<4, 1> __VA_ARGS__; break;                                    \
        default: break;                                                                     \
        }                                                                                   \
    } while (0)
static bool dense_widths_ok(size_t db, size_t pb)
{
    return db >= 1 && pb >= 1 && db + pb <= 5 && !(db == 2 && pb == 4) && !(db == 3 && pb == 3);
}

BchParams params_of(const poporon_cuda_bch *ctx)
{
    BchParams P;
    P.m = ctx->m;
    P.nn = ctx->nn;
    P.n = ctx->n;
    P.k = ctx->k;
    P.parity_bits = ctx->parity_bits;
    P.t = ctx->t;
    P.gen = ctx->gen;
    P.gen_deg = ctx->parity_bits;
    return P;
}

} // namespace

int launch_bch_encode(poporon_cuda_bch *ctx, const uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                      size_t parity_stride, size_t n, uint8_t *ok, cudaStream_t stream)
{
    if (n == 0)
        return 0;
    const int threads = 256;
    if (packed_1_2(ctx, data, data_stride, size, parity, parity_stride, ok, nullptr) && n >= 4) {
        const size_t n4 = n / 4, cap = (size_t)ctx->sm_count * 16;
        const size_t blocks = (n4 + threads - 1) / threads;
        bch_encode4_kernel<<<(unsigned)(blocks < cap ? blocks : cap), threads, 0, stream>>>(params_of(ctx), data, parity,
                                                                                      n4, ok);
        ctx->launches++;
        const size_t done = n4 * 4;
        if (done == n)
            return check(cudaGetLastError(), "bch_encode4 launch");
        data += done, parity += done * 2, ok = ok ? ok + done : ok, n -= done;
    }
    {
        // any code, dense rows: four codewords per thread on aligned words
        const size_t db = (size_t)(ctx->k + 7) / 8, pb = (size_t)(ctx->parity_bits + 7) / 8;
        if (n >= 4 && data_stride == db && size == db && parity_stride == pb && dense_widths_ok(db, pb) &&
            (((uintptr_t)data | (uintptr_t)parity | (uintptr_t)ok) & 3u) == 0) {
            const size_t n4 = n / 4, cap = (size_t)ctx->sm_count * 16;
            const size_t blocks = (n4 + threads - 1) / threads;
            PP_BCH_DENSE_DISPATCH(bch_encode_dense4_kernel, db, pb,
                                  <<<(unsigned)(blocks < cap ? blocks : cap), threads, 0, stream>>>(
                                      params_of(ctx), reinterpret_cast<const uint32_t *>(data),
                                      reinterpret_cast<uint32_t *>(parity), n4, ok));
            ctx->launches++;
            const size_t done = n4 * 4;
            if (done == n)
                return check(cudaGetLastError(), "bch_encode (dense4) launch");
            data += done * db, parity += done * pb, ok = ok ? ok + done : ok, n -= done;
        }
    }
    const size_t eblocks = (n + threads - 1) / threads, ecap = (size_t)ctx->sm_count * 8;
    bch_encode_kernel<<<(unsigned)(eblocks < ecap ? eblocks : ecap), threads, 0, stream>>>(
        params_of(ctx), data, data_stride, size, parity, parity_stride, n, ok);
    ctx->launches++;
    return check(cudaGetLastError(), "bch_encode launch");
}

int bch_build_decode_map(poporon_cuda_bch *ctx)
{
    if (ctx->n > kMapMaxN) {
        if (ctx->n > 31 || ctx->parity_bits > kSmapMaxBits || ctx->k > 26 || ctx->t > 15)
            return 0;
        const uint32_t words = 1u << ctx->parity_bits;
        PP_CUDA_TRY(cudaMalloc((void **)&ctx->d_smap, words * sizeof(uint32_t)));
        bch_build_smap_kernel<<<(words + 127) / 128, 128>>>(params_of(ctx), ctx->d_tables, ctx->d_smap);
        ctx->launches++;
        PP_CUDA_TRY(cudaGetLastError());
        return check(cudaDeviceSynchronize(), "bch remainder map");
    }
    const uint32_t words = 1u << ctx->n;
    PP_CUDA_TRY(cudaMalloc((void **)&ctx->d_map, words * sizeof(uint32_t)));
    bch_build_map_kernel<<<(words + 127) / 128, 128>>>(params_of(ctx), ctx->d_tables, ctx->d_map);
    ctx->launches++;
    PP_CUDA_TRY(cudaGetLastError());
    return check(cudaDeviceSynchronize(), "bch decode map");
}

int launch_bch_decode(poporon_cuda_bch *ctx, uint8_t *data, size_t data_stride, size_t size, const uint8_t *parity,
                      size_t parity_stride, size_t n, uint8_t *ok, uint32_t *corrected, cudaStream_t stream)
{
    if (n == 0)
        return 0;
    if (ctx->d_map && packed_1_2(ctx, data, data_stride, size, parity, parity_stride, ok, corrected) && n >= 4) {
        const size_t n4 = n / 4, cap = (size_t)ctx->sm_count * 16;
        const size_t blocks = (n4 + 255) / 256;
        bch_decode_map4_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, stream>>>(params_of(ctx), ctx->d_map,
                                                                                      data, parity, n4, ok, corrected);
        ctx->launches++;
        const size_t done = n4 * 4;
        if (done == n)
            return check(cudaGetLastError(), "bch_decode (map4) launch");
        data += done, parity += done * 2, ok = ok ? ok + done : ok, corrected = corrected ? corrected + done : corrected;
        n -= done;
    }
    if (ctx->d_map) {
        size_t blocks = (n + 255) / 256;
        const size_t cap = (size_t)ctx->sm_count * 32;
        bch_decode_map_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, stream>>>(
            params_of(ctx), ctx->d_map, data, data_stride, size, parity, parity_stride, n, ok, corrected);
        ctx->launches++;
        return check(cudaGetLastError(), "bch_decode (map) launch");
    }
    if (ctx->d_smap) {
        const size_t db = (size_t)(ctx->k + 7) / 8, pb = (size_t)(ctx->parity_bits + 7) / 8;
        if (n >= 4 && data_stride == db && size == db && parity_stride == pb && dense_widths_ok(db, pb) &&
            (((uintptr_t)data | (uintptr_t)parity | (uintptr_t)ok) & 3u) == 0 && ((uintptr_t)corrected & 15u) == 0) {
            const size_t n4 = n / 4, cap4 = (size_t)ctx->sm_count * 16;
            const size_t blocks4 = (n4 + 255) / 256;
            PP_BCH_DENSE_DISPATCH(bch_decode_smap4_kernel, db, pb,
                                  <<<(unsigned)(blocks4 < cap4 ? blocks4 : cap4), 256, 0, stream>>>(
                                      params_of(ctx), ctx->d_smap, reinterpret_cast<uint32_t *>(data),
                                      reinterpret_cast<const uint32_t *>(parity), n4, ok, corrected));
            ctx->launches++;
            const size_t done = n4 * 4;
            if (done == n)
                return check(cudaGetLastError(), "bch_decode (remainder map, dense4) launch");
            data += done * db, parity += done * pb, ok = ok ? ok + done : ok;
            corrected = corrected ? corrected + done : corrected;
            n -= done;
        }
        size_t blocks = (n + 255) / 256;
        const size_t cap = (size_t)ctx->sm_count * 8;
        bch_decode_smap_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, stream>>>(
            params_of(ctx), ctx->d_smap, data, data_stride, size, parity, parity_stride, n, ok, corrected);
        ctx->launches++;
        return check(cudaGetLastError(), "bch_decode (remainder map) launch");
    }
    const int threads = 128;
    bch_decode_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, stream>>>(
        params_of(ctx), ctx->d_tables, data, data_stride, size, parity, parity_stride, n, ok, corrected);
    ctx->launches++;
    return check(cudaGetLastError(), "bch_decode launch");
}

} // namespace ppcu
