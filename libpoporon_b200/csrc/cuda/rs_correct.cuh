// rs_correct.cuh — the per-codeword error/erasure correction stage (locator, Berlekamp-Massey, Chien,
// Forney, re-syndrome verification, in-place apply) shared by the general-parameter decode kernel
// (rs_generic.cu) and the RS(255,223) dirty-queue kernel (rs255.cu).  One thread owns one codeword.
//
// Semantics follow the reference exactly, including its integer widths where they are observable
// (paths relative to the reference tree):
//   locator init / BM ....... src/decode.c:31-96
//   degree, Chien ........... src/decode.c:98-145
//   omega, Forney ........... src/decode.c:147-191
//   re-syndrome check ....... src/decode.c:193-209
//   apply ................... src/decode.c:211-227        outcome table: SURVEY Appendix A.5
#pragma once
#include "common.cuh"

namespace ppcu {

struct SmemTables {
    uint8_t alpha_to[256];
    uint8_t index_of[256];
    uint8_t genpoly[256];
};

__device__ __forceinline__ void load_tables(SmemTables &s, const RsTables *g)
{
    const uint32_t *src = reinterpret_cast<const uint32_t *>(g);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&s);
    for (int i = threadIdx.x; i < (int)(sizeof(SmemTables) / 4); i += blockDim.x)
        dst[i] = src[i];
    __syncthreads();
}

// v mod nn for a 16-bit v (what the reference's fold loop computes, src/internal/common.h:102-110).
// nn = 255 (the common field) folds bytes instead of dividing by a run-time value: 2^8 = 1 (mod 255).
__device__ __forceinline__ uint32_t redc(uint32_t v16, int nn)
{
    v16 &= 0xFFFFu;
    if (nn == 255) {
        v16 = (v16 & 0xFFu) + (v16 >> 8); // <= 510
        v16 = (v16 & 0xFFu) + (v16 >> 8); // <= 255
        return v16 == 255u ? 0u : v16;
    }
    return v16 % (uint32_t)nn;
}

// the same for a sum of two logarithms (each < nn, so the sum is < 2*nn): one conditional subtraction
__device__ __forceinline__ uint32_t redc2(uint32_t v, int nn)
{
    return v >= (uint32_t)nn ? v - (uint32_t)nn : v;
}

// Branch-free forms for the register-resident paths (MAXR <= 32).  wrapn: t mod nn for t < 2*nn, and a huge value
// stays huge: a dead term (zero factor) is coded as kDead and clamps to alpha_to[A0] = 0 at the look-up.
constexpr uint32_t kDead = 0xFFFF0000u;
__device__ __forceinline__ uint32_t wrapn(uint32_t t, int nn)
{
    return min(t, t - (uint32_t)nn);
}

// Work arrays of the correction stage (7 arrays of MAXR + 1 bytes, indexed at run time).
// LocalWork: per-thread arrays (local memory) - any MAXR.  SmemWork: one column per thread of a shared-memory
// block laid out [array][index][thread] (consecutive threads = consecutive bytes: conflict-free, and no
// local-memory round trips), for CTAs of exactly kWorkThreads threads and small MAXR.
#ifndef PP_CORRECT_THREADS
#define PP_CORRECT_THREADS 128
#endif
constexpr int kWorkThreads = PP_CORRECT_THREADS;

template <int MAXR>
struct LocalWork {
    uint8_t a[7][MAXR + 1];
    struct Ref {
        uint8_t *p;
        __device__ __forceinline__ uint8_t &operator[](int i) const { return p[i]; }
    };
    __device__ __forceinline__ Ref arr(int k) { return Ref{a[k]}; }
};

template <int MAXR>
struct SmemWork {
    uint8_t *base; // this thread's column: smem block + threadIdx.x
    struct Ref {
        uint8_t *p;
        __device__ __forceinline__ uint8_t &operator[](int i) const { return p[i * kWorkThreads]; }
    };
    __device__ __forceinline__ Ref arr(int k) { return Ref{base + k * (MAXR + 1) * kWorkThreads}; }
    static constexpr int kBytes = 8 * (MAXR + 1) * kWorkThreads; // 7 arrays + the syndromes (array 7)
};

// The correction stage, shared by all three decode modes.  Returns the reference's bool and sets
// ncorr when the reference would have set *errors_corrected.
//
// PSYNC is called between the phases (locator, root search, evaluator, magnitudes, check + apply) by EVERY thread of
// the CTA, also by those that have already failed or were not given a codeword (live = false): with CtaPhaseSync the
// warps of a CTA run the same phase at the same time and share its code in the instruction cache.
struct NoPhaseSync {
    __device__ __forceinline__ void operator()() const {}
};
struct CtaPhaseSync {
    __device__ __forceinline__ void operator()() const { __syncthreads(); }
};

template <int MAXR, class PSYNC = NoPhaseSync, class WORK, class SYN>
__device__ bool rs_correct(const RsCode &c, const SmemTables &T, uint8_t *data, size_t size, uint8_t *parity,
                           WORK &wk, const SYN syn /* index form, [i] */, uint32_t ecount, const uint32_t *epos,
                           size_t pos_stride, bool erasure_apply, int16_t pad, uint32_t &ncorr, bool live = true)
{
    const PSYNC phase_sync;
    bool failed = !live;
    const int nr = c.nroots, nn = c.nn;
    const uint8_t A0 = (uint8_t)nn;
    const auto lambda = wk.arr(0), b = wk.arr(1), t = wk.arr(2), omega = wk.arr(3);
    const auto root = wk.arr(4), loc = wk.arr(5), reg = wk.arr(6);

    int deg = 0;
    if (!failed) {
    for (int i = 0; i <= nr; i++)
        lambda[i] = 0;
    lambda[0] = 1;
    if (ecount > 0) {
        lambda[1] = T.alpha_to[redc((uint32_t)c.prim * (uint32_t)(nn - 1 - (epos[0] + (uint32_t)(int32_t)pad)), nn)];
        for (int i = 1; (uint32_t)i < ecount; i++) {
            const uint32_t term =
                redc((uint32_t)c.prim * (uint32_t)(nn - 1 - (epos[i] + (uint32_t)(int32_t)pad)), nn) & 0xFF;
            for (int j = i + 1; j > 0; j--) {
                const uint8_t lg = T.index_of[lambda[j - 1]];
                if (lg != A0)
                    lambda[j] ^= T.alpha_to[redc2(term + lg, nn)];
            }
        }
    }
    for (int i = 0; i <= nr; i++)
        b[i] = T.index_of[lambda[i]];

    // r runs over 1 .. nr for every thread and a thread joins at r = ecount + 1 (src/decode.c:49: r = no_eras; while
    // (++r <= nroots)): the threads of a warp are then at the same r, so the inner loops have one trip count.
    uint32_t el = ecount;
    for (uint32_t r = 1; r <= (uint32_t)nr; r++) {
        if (r <= ecount)
            continue;
        uint32_t dsum = 0;
        for (uint32_t i = 0; i < r; i++) {
            const uint32_t s = syn[r - i - 1], il = T.index_of[lambda[i]];
            const uint32_t v = T.alpha_to[wrapn(il + s, nn)];
            dsum ^= (il != A0 && s != A0) ? v : 0u;
        }
        const uint8_t disc = T.index_of[dsum];
        if (disc == A0) {
            for (int i = nr; i > 0; i--)
                b[i] = b[i - 1];
            b[0] = A0;
        } else {
            t[0] = lambda[0];
            for (int i = 0; i < nr; i++) {
                const uint32_t bi = b[i], li = lambda[i + 1];
                const uint32_t v = T.alpha_to[wrapn((uint32_t)disc + bi, nn)];
                t[i + 1] = (uint8_t)(li ^ (bi != A0 ? v : 0u));
            }
            if (2 * el <= r + ecount - 1) {
                el = r + ecount - el;
                for (int i = 0; i <= nr; i++)
                    b[i] = (lambda[i] == 0) ? A0
                                            : (uint8_t)redc((uint32_t)T.index_of[lambda[i]] - disc + nn, nn);
            } else {
                for (int i = nr; i > 0; i--)
                    b[i] = b[i - 1];
                b[0] = A0;
            }
            for (int i = 0; i <= nr; i++)
                lambda[i] = t[i];
        }
    }

    for (int i = 0; i <= nr; i++) {
        lambda[i] = T.index_of[lambda[i]];
        if (lambda[i] != A0)
            deg = i;
    }
    if (deg == 0)
        failed = true;
    }
    phase_sync();

    // a failed thread walks through the remaining phases with deg = count = 0 (empty loops) and returns before the check
    int count = 0;
    if (failed) {
    } else if constexpr (MAXR <= 32) {
        // the running terms stay in registers (fully unrolled over the at most 32 coefficients): one table look-up
        // per live term and location instead of three shared-memory accesses
        // A term that is zero (or past the degree) is coded as kDead: it stays huge under the update and looks up
        // alpha_to[A0] = 0, so a term is five integer instructions and one look-up, without a branch.
        uint32_t rg[MAXR + 1];
#pragma unroll
        for (int j = 1; j <= MAXR; j++)
            rg[j] = (j <= deg && lambda[j] != A0) ? (uint32_t)lambda[j] : kDead;
        int16_t k = (int16_t)(c.iprim - 1);
        for (int i = 1; i <= nn; i++, k = (int16_t)redc((uint32_t)(int32_t)(k + c.iprim), nn)) {
            uint32_t q = 1;
#pragma unroll
            for (int j = MAXR; j > 0; j--) {
                rg[j] = wrapn(rg[j] + (uint32_t)j, nn);
                q ^= T.alpha_to[min(rg[j], (uint32_t)A0)];
            }
            if (q != 0)
                continue;
            if (k < pad) {
                failed = true;
                break;
            }
            root[count] = (uint8_t)i;
            loc[count] = (uint8_t)k;
            if (++count == deg)
                break;
        }
    } else {
        for (int i = 1; i <= nr; i++)
            reg[i] = lambda[i];
        int16_t k = (int16_t)(c.iprim - 1);
        for (int i = 1; i <= nn; i++, k = (int16_t)redc((uint32_t)(int32_t)(k + c.iprim), nn)) {
            uint8_t q = 1;
            for (int j = deg; j > 0; j--) {
                if (reg[j] != A0) {
                    reg[j] = (uint8_t)redc2((uint32_t)reg[j] + j, nn);
                    q ^= T.alpha_to[reg[j]];
                }
            }
            if (q != 0)
                continue;
            if (k < pad) {
                failed = true;
                break;
            }
            root[count] = (uint8_t)i;
            loc[count] = (uint8_t)k;
            if (++count == deg)
                break;
        }
    }
    if (deg != count)
        failed = true;
    if (failed)
        deg = count = 0;
    phase_sync();

    if constexpr (MAXR <= 32) {
        // omega = S * lambda mod x^deg as lambda_j times the syndromes shifted up by j places: all 32 product terms of a
        // step in registers with static indices, dead-coded, no branch (the triangular loop nest ran at 7 of 32 lanes)
        uint32_t sh[MAXR], oacc[MAXR];
#pragma unroll
        for (int i = 0; i < MAXR; i++) {
            sh[i] = (i < nr && syn[i] != A0) ? (uint32_t)syn[i] : kDead;
            oacc[i] = 0;
        }
        for (int j = 0; j < deg; j++) {
            const uint32_t lj = lambda[j] != A0 ? (uint32_t)lambda[j] : kDead;
#pragma unroll
            for (int i = 0; i < MAXR; i++)
                oacc[i] ^= T.alpha_to[min(wrapn(sh[i] + lj, nn), (uint32_t)A0)];
#pragma unroll
            for (int i = MAXR - 1; i > 0; i--)
                sh[i] = sh[i - 1];
            sh[0] = kDead;
        }
#pragma unroll
        for (int i = 0; i < MAXR; i++)
            if (i < deg)
                omega[i] = T.index_of[oacc[i]];
    } else
    for (int i = 0; i < deg; i++) {
        uint8_t acc = 0;
        for (int j = i; j >= 0; j--) {
            if (syn[i - j] != A0 && lambda[j] != A0)
                acc ^= T.alpha_to[redc2((uint32_t)syn[i - j] + lambda[j], nn)];
        }
        omega[i] = T.index_of[acc];
    }

    phase_sync();
    if (!failed)
        ncorr = 0;
    if constexpr (MAXR <= 32) {
        // evaluator and odd locator coefficients in registers, dead-coded; i * root (mod nn) kept as a running sum
        // (i * root <= 31 * 255 never reaches the 16-bit truncation of the reference's gf_mod argument)
        uint32_t og[MAXR], ld[MAXR / 2];
#pragma unroll
        for (int i = 0; i < MAXR; i++)
            og[i] = (i < deg && omega[i] != A0) ? (uint32_t)omega[i] : kDead;
#pragma unroll
        for (int i = 0; i < MAXR; i += 2)
            ld[i / 2] = (i <= ((deg < nr - 1 ? deg : nr - 1) & ~1) && lambda[i + 1] != A0) ? (uint32_t)lambda[i + 1] : kDead;
        for (int j = count - 1; j >= 0; j--) {
            const uint32_t rt = root[j];
            const uint32_t rtn = rt == (uint32_t)nn ? 0u : rt; // root mod nn
            uint32_t m = 0, num1 = 0, den = 0;
#pragma unroll
            for (int i = 0; i < MAXR; i++) {
                num1 ^= T.alpha_to[min(wrapn(og[i] + m, nn), (uint32_t)A0)];
                if ((i & 1) == 0)
                    den ^= T.alpha_to[min(wrapn(ld[i / 2] + m, nn), (uint32_t)A0)];
                m = wrapn(m + rtn, nn);
            }
            if (num1 == 0) {
                b[j] = 0;
                continue;
            }
            const uint8_t num2 = T.alpha_to[redc(rt * ((uint32_t)c.fcr - 1u) + (uint32_t)nn, nn)];
            b[j] = T.alpha_to[redc((uint32_t)T.index_of[num1] + T.index_of[num2] + nn - T.index_of[den], nn)];
            ncorr++;
        }
    } else
    for (int j = count - 1; j >= 0; j--) {
        uint8_t num1 = 0, den = 0;
        for (int i = deg - 1; i >= 0; i--) {
            if (omega[i] != A0)
                num1 ^= T.alpha_to[redc((uint32_t)omega[i] + (uint32_t)(i * root[j]), nn)];
        }
        if (num1 == 0) {
            b[j] = 0;
            continue;
        }
        const uint8_t num2 = T.alpha_to[redc((uint32_t)root[j] * ((uint32_t)c.fcr - 1u) + (uint32_t)nn, nn)];
        for (int i = (deg < nr - 1 ? deg : nr - 1) & ~1; i >= 0; i -= 2) {
            if (lambda[i + 1] != A0)
                den ^= T.alpha_to[redc((uint32_t)lambda[i + 1] + (uint32_t)(i * root[j]), nn)];
        }
        b[j] = T.alpha_to[redc((uint32_t)T.index_of[num1] + T.index_of[num2] + nn - T.index_of[den], nn)];
        ncorr++;
    }

    phase_sync();
    if (failed)
        return false;

    // Re-syndrome check.  When no exponent can reach 2^16 (the reference truncates its gf_mod argument to 16 bits)
    // the exponent of correction j at syndrome i is ib_j + (fcr + i) * step_j (mod nn): one running sum per j.
    const bool plain_exponents =
        (uint64_t)((uint32_t)c.fcr + (uint32_t)nr) * (uint64_t)c.prim * (uint64_t)(nn - 1) + (uint64_t)nn < 65536ull; // (64-bit: a large prim must not wrap)
    if (MAXR <= 32 && plain_exponents) {
        uint32_t acc[MAXR <= 32 ? MAXR : 1];
#pragma unroll
        for (int i = 0; i < (MAXR <= 32 ? MAXR : 1); i++)
            acc[i] = 0;
        for (int j = 0; j < count; j++) {
            const uint32_t bj = b[j];
            if (bj == 0)
                continue;
            const uint32_t step = redc((uint32_t)c.prim * (uint32_t)(nn - loc[j] - 1), nn);
            uint32_t e = redc((uint32_t)T.index_of[bj] + (uint32_t)c.fcr * (uint32_t)c.prim * (uint32_t)(nn - loc[j] - 1), nn);
#pragma unroll
            for (int i = 0; i < (MAXR <= 32 ? MAXR : 1); i++) {
                acc[i] ^= T.alpha_to[e];
                e = wrapn(e + step, nn);
            }
        }
        bool same = true;
#pragma unroll
        for (int i = 0; i < (MAXR <= 32 ? MAXR : 1); i++)
            if (i < nr && acc[i] != T.alpha_to[syn[i]])
                same = false;
        if (!same)
            return false;
    } else
    for (int i = 0; i < nr; i++) {
        uint8_t acc = 0;
        for (int j = 0; j < count; j++) {
            if (b[j] == 0)
                continue;
            // the reference forms this product in int, stores it to an int16_t and adds it back in int
            // before truncating to 16 bits: only the low 16 bits of the sum survive, so wrap-around
            // unsigned arithmetic gives the same value
            const uint32_t k16 = (uint32_t)(c.fcr + i) * (uint32_t)c.prim * (uint32_t)(nn - loc[j] - 1);
            acc ^= T.alpha_to[redc((uint32_t)T.index_of[b[j]] + k16, nn)];
        }
        if (acc != T.alpha_to[syn[i]])
            return false;
    }

    if (erasure_apply) {
        for (int i = 0; i < count; i++) {
            // slot i of the caller's row, in Chien order (src/decode.c:211-214); slots past the row or
            // positions past the message are outside the reference's defined behaviour and are skipped
            if ((size_t)i < pos_stride) {
                const uint32_t p = epos[i];
                if (p < size)
                    data[p] ^= b[i];
            }
        }
    } else {
        for (int i = 0; i < count; i++) {
            const int32_t p = (int32_t)loc[i] - (int32_t)pad;
            if (p >= 0 && p < (int32_t)size)
                data[p] ^= b[i];
            else if (p >= (int32_t)size && p < (int32_t)(size + nr))
                parity[p - size] ^= b[i];
            else
                return false;
        }
    }
    return true;
}

} // namespace ppcu
