// rs_generic.cu — table-driven Reed-Solomon kernels for ANY code the reference accepts
// (GF(2^m), m <= 8, arbitrary primitive polynomial, fcr, prim, nroots, shortened sizes).
//
// One thread owns one codeword.  These kernels are the general-parameter path behind
// poporon_cuda_rs_encode_dev / poporon_cuda_rs_decode_dev; the RS(255,223) default code takes the
// bitsliced kernels in rs255_*.cu instead.  Log/antilog tables and the generator live in shared
// memory (768 B per CTA); per-codeword polynomials live in thread-local arrays.
//
// Semantics follow the reference exactly, including its integer widths where they are observable
// (paths relative to the reference tree):
//   encoder ................. src/encode.c:120-144
//   syndromes ............... src/decode.c:375-415
//   locator init / BM ....... src/decode.c:31-96
//   degree, Chien ........... src/decode.c:98-145
//   omega, Forney ........... src/decode.c:147-191
//   re-syndrome check ....... src/decode.c:193-209
//   apply ................... src/decode.c:211-227
//   glue / outcome table .... src/decode.c:418-487, 596-600   
#include "common.cuh"
#include "rs_correct.cuh"
#include "rs_generic.cuh"

namespace ppcu {

namespace {

template <int MAXR>
__global__ void __launch_bounds__(128)
rs_encode_generic_kernel(RsCode c, const RsTables *__restrict__ gtab, const uint8_t *__restrict__ data,
                         size_t data_stride, size_t size, uint8_t *__restrict__ parity, size_t parity_stride,
                         size_t n, uint8_t *__restrict__ ok)
{
    __shared__ SmemTables T;
    __shared__ uint8_t mtab[MAXR <= 32 ? 32 * 256 : 4]; // mtab[k][v] = v * g_k (src/encode.c:120-144 arithmetic)
    load_tables(T, gtab);
    if constexpr (MAXR <= 32) {
        for (int e = threadIdx.x; e < c.nroots * 256; e += blockDim.x) {
            const uint32_t lg = T.index_of[e & 255];
            uint32_t x = lg + T.genpoly[e >> 8];
            x = x >= (uint32_t)c.nn ? x - c.nn : x;
            mtab[e] = lg == (uint32_t)c.nn ? 0 : T.alpha_to[x];
        }
        __syncthreads();
    }
    const size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (cw >= n)
        return;
    const int nr = c.nroots, nn = c.nn;
    if constexpr (MAXR <= 32) {
        // parity register in registers (static indices, rows past nroots stay zero); fb * g_k by one look-up in a
        // per-CTA product table built with the reference's own log arithmetic
        uint32_t par[MAXR + 1];
#pragma unroll
        for (int j = 0; j <= MAXR; j++)
            par[j] = 0;
        const uint8_t *d = data + cw * data_stride;
        const uint8_t *row0 = mtab + (nr - 1) * 256; // row of g_{nroots-1-q} for q = 0
        for (size_t i = 0; i < size; i++) {
            const uint32_t fb = (uint32_t)(d[i] & nn) ^ par[0];
#pragma unroll
            for (int q = 0; q < MAXR; q++)
                par[q] = q < nr ? par[q + 1] ^ (uint32_t)row0[(int)fb - q * 256] : 0u;
        }
        uint8_t *p = parity + cw * parity_stride;
#pragma unroll
        for (int j = 0; j < MAXR; j++)
            if (j < nr)
                p[j] = (uint8_t)par[j];
        if (ok)
            ok[cw] = 1;
        return;
    }
    uint8_t ring[MAXR]; // circular LFSR state: logical p[j] = ring[(head + j) mod nr]
    for (int j = 0; j < nr; j++)
        ring[j] = 0;
    int head = 0;
    const uint8_t *d = data + cw * data_stride;
    for (size_t i = 0; i < size; i++) {
        const uint32_t fb = T.index_of[(d[i] & nn) ^ ring[head]];
        if (fb != (uint32_t)nn) {
            int pos = head;
            for (int j = 1; j < nr; j++) {
                pos = (pos + 1 == nr) ? 0 : pos + 1;
                uint32_t e = fb + T.genpoly[nr - j];
                e = e >= (uint32_t)nn ? e - nn : e;
                ring[pos] ^= T.alpha_to[e];
            }
            uint32_t e0 = fb + T.genpoly[0];
            e0 = e0 >= (uint32_t)nn ? e0 - nn : e0;
            ring[head] = T.alpha_to[e0];
        } else {
            ring[head] = 0;
        }
        head = (head + 1 == nr) ? 0 : head + 1;
    }
    uint8_t *p = parity + cw * parity_stride;
    for (int j = 0; j < nr; j++) {
        int pos = head + j;
        pos = pos >= nr ? pos - nr : pos;
        p[j] = ring[pos];
    }
    if (ok)
        ok[cw] = 1;
}

template <int MAXR>
__global__ void __launch_bounds__(128)
rs_decode_generic_kernel(RsCode c, const RsTables *__restrict__ gtab, int mode, uint8_t *data, size_t data_stride,
                         size_t size, uint8_t *parity, size_t parity_stride, size_t n,
                         const uint32_t *__restrict__ pos, size_t pos_stride, const uint32_t *__restrict__ cnt,
                         const uint16_t *__restrict__ extsyn, uint8_t *__restrict__ ok,
                         uint32_t *__restrict__ corrected, uint16_t *__restrict__ synout)
{
    __shared__ SmemTables T;
    __shared__ uint8_t mtab[MAXR <= 32 ? 32 * 256 : 4]; // mtab[i][v] = v * alpha^((fcr + i) * prim), src/decode.c:375-415
    load_tables(T, gtab);
    if constexpr (MAXR <= 32) {
        if (mode != POPORON_CUDA_MODE_EXTSYN) {
            for (int e = threadIdx.x; e < c.nroots * 256; e += blockDim.x) {
                const uint32_t v = e & 255;
                mtab[e] = v == 0 ? 0
                                 : T.alpha_to[redc((uint32_t)T.index_of[v] + (uint32_t)(c.fcr + (e >> 8)) * (uint32_t)c.prim,
                                                   c.nn)];
            }
        }
        __syncthreads();
    }
    const size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool present = cw < n; // threads past the batch still walk through the phase barriers of rs_correct
    const int nr = c.nroots, nn = c.nn;
    const uint8_t A0 = (uint8_t)nn;
    uint8_t *d = data + cw * data_stride;
    uint8_t *p = parity + cw * parity_stride;
    uint32_t ncorr = 0;
    bool success = false;
    // nroots <= 32: work arrays in shared memory ([array][index][thread]) and the warps of the CTA phase-locked,
    // as in rs255_correct_kernel; otherwise thread-local arrays
    extern __shared__ __align__(16) uint8_t gwork[];
    [[maybe_unused]] uint8_t syn_local[MAXR <= 32 ? 1 : MAXR + 1];
    [[maybe_unused]] LocalWork<MAXR <= 32 ? 0 : MAXR> wk_local;
    auto run = [&](auto &wk, auto syn) {
    bool need = false, erasure = false; // rs_correct is called by every thread, need = this one has work for it
    uint32_t ec = 0;

    const int16_t pad = (int16_t)(uint16_t)((size_t)nn - (size_t)nr - size); // src/decode.c:418-429
    if (present && size != 0 && !(pad < 0 || pad >= (int16_t)(nn - nr))) {
        if (mode == POPORON_CUDA_MODE_EXTSYN) {
            bool dirty = false;
            for (int i = 0; i < nr; i++) {
                const uint16_t s = extsyn[cw * (size_t)nr + i];
                // index-form values above nn cannot be produced by the reference's own tables; clamp
                // them into the table (they would index out of bounds there)
                syn[i] = (uint8_t)(s > (uint16_t)nn ? nn : s);
                dirty |= (s != (uint16_t)nn);
            }
            success = !dirty;
            need = dirty;
        } else {
            // Horner over data || parity at the generator roots
            const uint32_t first = d[0] & nn;
            const size_t total = size + (size_t)nr;
            if constexpr (MAXR <= 32) {
                // the 32 running sums in registers, one product look-up per root and symbol
                uint32_t sr[MAXR];
#pragma unroll
                for (int i = 0; i < MAXR; i++)
                    sr[i] = first;
                for (size_t j = 1; j < total; j++) {
                    const uint32_t sym = (j < size ? d[j] : p[j - size]) & nn;
#pragma unroll
                    for (int i = 0; i < MAXR; i++)
                        if (i < nr)
                            sr[i] = sym ^ (uint32_t)mtab[i * 256 + sr[i]];
                }
#pragma unroll
                for (int i = 0; i < MAXR; i++)
                    if (i < nr)
                        syn[i] = (uint8_t)sr[i];
            } else {
            for (int i = 0; i < nr; i++)
                syn[i] = (uint8_t)first;
            for (size_t j = 1; j < total; j++) {
                const uint32_t sym = (j < size ? d[j] : p[j - size]) & nn;
                for (int i = 0; i < nr; i++) {
                    const uint8_t s = syn[i];
                    syn[i] = (s == 0) ? (uint8_t)sym
                                      : (uint8_t)(sym ^ T.alpha_to[redc((uint32_t)T.index_of[s] +
                                                                            (uint32_t)(c.fcr + i) * (uint32_t)c.prim,
                                                                        nn)]);
                }
            }
            }
            uint32_t any = 0;
            for (int i = 0; i < nr; i++) {
                any |= syn[i];
                syn[i] = T.index_of[syn[i]];
            }
            if (synout) { // syndromes-only call: index form, as calculate_syndrome_u8 leaves them (src/decode.c:405-413)
                for (int i = 0; i < nr; i++)
                    synout[cw * (size_t)nr + i] = syn[i];
                success = true;
            } else if (!any) {
                success = true;
            } else if (mode == POPORON_CUDA_MODE_ERASURE) {
                ec = cnt[cw];
                need = erasure = ec <= (uint32_t)nr; // > nroots erasures overrun the reference's scratch: failure
            } else {
                need = true;
            }
        }
    }
    if (synout != nullptr) {
        // (uniform over the launch: nobody enters the correction stage)
    } else if constexpr (MAXR <= 32) {
        const bool r = rs_correct<MAXR, CtaPhaseSync>(c, T, d, size, p, wk, syn, erasure ? ec : 0u,
                                                      erasure ? pos + cw * pos_stride : nullptr, pos_stride, erasure, pad,
                                                      ncorr, need);
        if (need)
            success = r;
    } else if (need) {
        success = rs_correct<MAXR>(c, T, d, size, p, wk, syn, erasure ? ec : 0u,
                                   erasure ? pos + cw * pos_stride : nullptr, pos_stride, erasure, pad, ncorr);
    }
    };
    if constexpr (MAXR <= 32) {
        SmemWork<32> wk{gwork + threadIdx.x};
        run(wk, wk.arr(7));
    } else {
        run(wk_local, (uint8_t *)syn_local);
    }
    (void)A0;
    if (!present)
        return;
    if (ok)
        ok[cw] = success ? 1 : 0;
    if (corrected)
        corrected[cw] = ncorr;
}

} // namespace

int launch_rs_encode_generic(poporon_cuda_rs *ctx, const uint8_t *data, size_t data_stride, size_t size,
                             uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, cudaStream_t stream)
{
    if (n == 0)
        return 0;
    const int threads = 128;
    const unsigned blocks = (unsigned)((n + threads - 1) / threads);
    if (ctx->code.nroots <= 32)
        rs_encode_generic_kernel<32><<<blocks, threads, 0, stream>>>(ctx->code, ctx->d_tables, data, data_stride, size,
                                                                     parity, parity_stride, n, ok);
    else
        rs_encode_generic_kernel<255><<<blocks, threads, 0, stream>>>(ctx->code, ctx->d_tables, data, data_stride,
                                                                      size, parity, parity_stride, n, ok);
    ctx->launches++;
    return check(cudaGetLastError(), "rs_encode_generic launch");
}

int launch_rs_decode_generic(poporon_cuda_rs *ctx, int mode, uint8_t *data, size_t data_stride, size_t size,
                             uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *pos, size_t pos_stride,
                             const uint32_t *cnt, const uint16_t *syn, uint8_t *ok, uint32_t *corrected,
                             cudaStream_t stream, uint16_t *synout)
{
    if (n == 0)
        return 0;
    const int threads = 128;
    const unsigned blocks = (unsigned)((n + threads - 1) / threads);
    if (ctx->code.nroots <= 32)
        rs_decode_generic_kernel<32><<<blocks, threads, SmemWork<32>::kBytes, stream>>>(ctx->code, ctx->d_tables, mode, data, data_stride,
                                                                     size, parity, parity_stride, n, pos, pos_stride,
                                                                     cnt, syn, ok, corrected, synout);
    else
        rs_decode_generic_kernel<255><<<blocks, threads, 0, stream>>>(ctx->code, ctx->d_tables, mode, data,
                                                                      data_stride, size, parity, parity_stride, n, pos,
                                                                      pos_stride, cnt, syn, ok, corrected, synout);
    ctx->launches++;
    return check(cudaGetLastError(), "rs_decode_generic launch");
}

} // namespace ppcu
