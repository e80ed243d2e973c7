// rs255_solve_core.h — bitsliced error solver for RS(255,223)/0x11D, fcr = 1, prim = 1, plain mode.
//
// Replaces the reference's error_correction_u8 (src/decode.c:17-230) for the common case: the key-equation
// solver (Berlekamp-Massey, src/decode.c:53-96), the Chien search (:117-141), the error evaluator
// (:147-158) and Forney's formula (:159-191), for 32 codewords at once per thread ("lane"): every value is
// held as 8 bit-planes, one 32-bit word per plane, bit c of a word belonging to codeword c of the lane.
// There are no table look-ups and no data-dependent branches in the arithmetic:
//
//   * a variable x variable GF(2^8) product is 64 AND-XOR (LOP3) into 15 partial planes + a 16-op reduction
//     modulo x^8 + x^4 + x^3 + x^2 + 1; products that are summed share one reduction;
//   * a product by a compile-time constant is a fixed XOR network;
//   * the reference's Karn-style BM divides by the discrepancy (b = lambda / delta, src/decode.c:84-90).
//     Here the inversion-free form is used: T = gamma*Lambda + Delta*x*B;  on a length change B = Lambda,
//     gamma = Delta.  With identical branch conditions (Delta != 0, 2*el <= r - 1) every polynomial is a
//     non-zero scalar multiple of the reference's, so degree, roots, L and the Forney ratio are identical;
//   * the length test runs on a bitsliced counter k = r - 1 - 2*el (k >= 0  <=>  2*el <= r - 1;
//     length change: k <- -k - 1 = ~k; otherwise k <- k + 1);
//   * the Chien search advances the 16 locator terms by constant multipliers location by location and
//     tests the sum for zero; the root masks are transposed in blocks of 32 locations and the roots
//     x = alpha^(k+1) appended to a byte list per codeword (about one root per lane and location);
//   * Forney's formula then runs bitsliced on the list slots (slot e of all 32 codewords at once):
//     x*Omega(x) by Horner, x*Lambda'(x) from the odd terms, the quotient through a^254.
//
// Domain.  The solver keeps 17 locator coefficients: it is exact for every codeword whose final register
// length L is <= 16 and whose locator has degree L.  Anything else (L > 16, or a vanishing leading
// coefficient - both only happen for uncorrectable words) is reported as "declined" and is decoded by the
// general per-codeword stage (rs_correct.cuh), which restates the reference literally.  For accepted
// codewords the reference's outcome is: roots found == L and none in the padding ->
// true, corrected = number of non-zero magnitudes; otherwise false, 0.  (The re-syndrome verification,
// src/decode.c:193-209, cannot fail when L distinct roots of a length-L connection polynomial were found.)
//
// The code is plain C++ (uint32_t logic only) so that the same functions run under nvcc for sm_100a and,
// for unit tests, on the host (tests/host_solver.cpp emulates one lane at a time).
#pragma once
#include <stdint.h>

#include <utility>

#include "gf256_consts.cuh"
#include "rs255_fft_gen.cuh"

#if defined(__CUDACC__)
#define BS_FN __host__ __device__ __forceinline__
#else
#define BS_FN inline __attribute__((always_inline))
#endif

// Compiler-level fence: keeps nvcc from hoisting every shared-memory load of a long unrolled sequence to the
// top (which costs more registers than the kernel has); loads are software-pipelined one term ahead instead.
#if defined(__CUDA_ARCH__)
#define BS_FENCE() asm volatile("" ::: "memory")
#else
#define BS_FENCE() ((void)0)
#endif

namespace bs {

constexpr int kT = 16;      // correctable errors
constexpr int kNL = kT + 1; // locator coefficients kept
constexpr int kNS = 32;     // syndromes

// "memory" = an array of 32-bit words with a fixed word stride: shared memory [word][lane] on the device
// (stride 32), a plain array on the host (stride 1)
template <int STRIDE>
struct Mem {
    uint32_t *base;
    BS_FN uint32_t ld(int idx) const { return base[idx * STRIDE]; }
    BS_FN void st(int idx, uint32_t v) const { base[idx * STRIDE] = v; }
    // half-word h (0/1) and byte b (0..3) of word idx (little endian, as the GPU sees them)
    BS_FN void st_half(int idx, int h, uint32_t v) const
    {
        reinterpret_cast<uint16_t *>(base + idx * STRIDE)[h] = (uint16_t)v;
    }
    BS_FN uint32_t ld_byte(int idx, int b) const { return reinterpret_cast<const uint8_t *>(base + idx * STRIDE)[b]; }
    BS_FN void st_byte(int idx, int b, uint32_t v) const
    {
        reinterpret_cast<uint8_t *>(base + idx * STRIDE)[b] = (uint8_t)v;
    }
};

template <int STRIDE>
BS_FN void ld8(const Mem<STRIDE> M, int blk, uint32_t (&v)[8])
{
#pragma unroll
    for (int t = 0; t < 8; t++)
        v[t] = M.ld(8 * blk + t);
}
template <int STRIDE>
BS_FN void st8(const Mem<STRIDE> M, int blk, const uint32_t (&v)[8])
{
#pragma unroll
    for (int t = 0; t < 8; t++)
        M.st(8 * blk + t, v[t]);
}
BS_FN void cp8(uint32_t (&d)[8], const uint32_t (&s)[8])
{
#pragma unroll
    for (int t = 0; t < 8; t++)
        d[t] = s[t];
}

// ---- variable x variable -------------------------------------------------------------------------------
BS_FN void mac(uint32_t (&acc)[15], const uint32_t (&a)[8], const uint32_t (&b)[8])
{
#pragma unroll
    for (int s = 0; s < 8; s++)
#pragma unroll
        for (int t = 0; t < 8; t++)
            acc[s + t] ^= a[s] & b[t];
}

BS_FN void clear15(uint32_t (&acc)[15])
{
#pragma unroll
    for (int i = 0; i < 15; i++)
        acc[i] = 0;
}

// x^8 = x^4 + x^3 + x^2 + 1  (0x11D)
BS_FN void reduce(uint32_t (&acc)[15], uint32_t (&out)[8])
{
#pragma unroll
    for (int k = 14; k >= 8; k--) {
        const uint32_t v = acc[k];
        acc[k - 8] ^= v;
        acc[k - 6] ^= v;
        acc[k - 5] ^= v;
        acc[k - 4] ^= v;
    }
#pragma unroll
    for (int i = 0; i < 8; i++)
        out[i] = acc[i];
}

BS_FN void mul(const uint32_t (&a)[8], const uint32_t (&b)[8], uint32_t (&out)[8])
{
    uint32_t acc[15];
    clear15(acc);
    mac(acc, a, b);
    reduce(acc, out);
}

// ---- fixed linear maps -----------------------------------------------------------------------------------
// out[t] = xor of in[s] over the bits s set in M[t]
template <unsigned M>
BS_FN uint32_t xor_subset(const uint32_t (&in)[8])
{
    uint32_t v = 0;
    if constexpr ((M & 1u) != 0) v ^= in[0];
    if constexpr ((M & 2u) != 0) v ^= in[1];
    if constexpr ((M & 4u) != 0) v ^= in[2];
    if constexpr ((M & 8u) != 0) v ^= in[3];
    if constexpr ((M & 16u) != 0) v ^= in[4];
    if constexpr ((M & 32u) != 0) v ^= in[5];
    if constexpr ((M & 64u) != 0) v ^= in[6];
    if constexpr ((M & 128u) != 0) v ^= in[7];
    return v;
}

// multiply by the constant C
template <uint8_t C>
BS_FN void cmul(const uint32_t (&in)[8], uint32_t (&out)[8])
{
    uint32_t r[8];
    r[0] = xor_subset<gf256::rowmask(C, 0)>(in);
    r[1] = xor_subset<gf256::rowmask(C, 1)>(in);
    r[2] = xor_subset<gf256::rowmask(C, 2)>(in);
    r[3] = xor_subset<gf256::rowmask(C, 3)>(in);
    r[4] = xor_subset<gf256::rowmask(C, 4)>(in);
    r[5] = xor_subset<gf256::rowmask(C, 5)>(in);
    r[6] = xor_subset<gf256::rowmask(C, 6)>(in);
    r[7] = xor_subset<gf256::rowmask(C, 7)>(in);
#pragma unroll
    for (int i = 0; i < 8; i++)
        out[i] = r[i];
}

// squaring is GF(2)-linear: row t collects the input bits s whose square (2^s)^2 has bit t
constexpr uint8_t sq_rowmask(int t)
{
    unsigned m = 0;
    for (int s = 0; s < 8; s++) {
        const uint8_t e = (uint8_t)(1u << s);
        if (gf256::mul(e, e) & (1u << t))
            m |= 1u << s;
    }
    return (uint8_t)m;
}

BS_FN void sqr(const uint32_t (&in)[8], uint32_t (&out)[8])
{
    uint32_t r[8];
    r[0] = xor_subset<sq_rowmask(0)>(in);
    r[1] = xor_subset<sq_rowmask(1)>(in);
    r[2] = xor_subset<sq_rowmask(2)>(in);
    r[3] = xor_subset<sq_rowmask(3)>(in);
    r[4] = xor_subset<sq_rowmask(4)>(in);
    r[5] = xor_subset<sq_rowmask(5)>(in);
    r[6] = xor_subset<sq_rowmask(6)>(in);
    r[7] = xor_subset<sq_rowmask(7)>(in);
#pragma unroll
    for (int i = 0; i < 8; i++)
        out[i] = r[i];
}

// a^254 = a^-1 (0 -> 0): 4 products, 7 squarings
BS_FN void inv(const uint32_t (&a)[8], uint32_t (&out)[8])
{
    uint32_t a2[8], a3[8], a12[8], a15[8], t[8];
    sqr(a, a2);
    mul(a2, a, a3);
    sqr(a3, t);
    sqr(t, a12);
    mul(a12, a3, a15);
    sqr(a15, t);
    sqr(t, t);
    sqr(t, t);
    sqr(t, t); // a^240
    mul(t, a12, t); // a^252
    mul(t, a2, out); // a^254
}

BS_FN uint32_t or8(const uint32_t (&a)[8])
{
    return (a[0] | a[1] | a[2]) | (a[3] | a[4] | a[5]) | (a[6] | a[7]);
}

// (a & m) | (b & ~m): one LOP3 on the device (the optimiser would otherwise unfold the masked merge)
BS_FN uint32_t merge_bits(uint32_t a, uint32_t b, uint32_t m)
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(d) : "r"(a), "r"(b), "r"(m));
    return d;
#else
    return (a & m) | (b & ~m);
#endif
}

// smallest of the 32 seven-bit two's-complement counters held in the planes K (bit-plane b = bit b)
BS_FN int min_counter(const uint32_t (&K)[7])
{
    uint32_t cand = 0xFFFFFFFFu;
    int v = 0;
    {
        const uint32_t neg = cand & K[6]; // any negative value is smaller than every non-negative one
        if (neg) {
            cand = neg;
            v = -64;
        }
    }
#pragma unroll
    for (int b = 5; b >= 0; b--) {
        const uint32_t zero = cand & ~K[b];
        if (zero)
            cand = zero;
        else
            v += 1 << b;
    }
    return v;
}

#if defined(__CUDA_ARCH__)
#define BS_WARP_MIN(x) __reduce_min_sync(0xFFFFFFFFu, (x))
#else
#define BS_WARP_MIN(x) (x)
#endif

// ---- inversion-free Berlekamp-Massey ---------------------------------------------------------------------
// S: 32 syndromes, word (8*m + t) = plane t of S_{m+1}.  B: scratch of kT*8 words.
// On return L[0..16] is the locator (a non-zero multiple of the reference's lambda; L[17] is padding), K the
// counter planes after the last iteration (k = 32 - 2*el, two's complement, 7 planes).
//
// One pass over the coefficients per iteration r: T_i = gamma*L_i + Delta_r*B_{i-1} replaces L_i, B_i becomes
// L_i or B_{i-1}, and T_i * S_{r+1-i} is accumulated straight into the NEXT discrepancy Delta_{r+1}.
// Instruction-cache shape: the 18 coefficient slots are kNG groups of kG; the loop body is written for the
// group held in L[0..kG-1] and the groups are rotated through those registers (moves, which run on the FMA pipe).
// The straight-line version (17 coefficient steps) ran at a third of the ALU rate, fetch-bound
// (tools/ubench_icache.cu); with eight desynchronised warps per SM sharing a 32 KB instruction cache even the
// 6-step body (28 KB) missed constantly: measured on B200, 16 Mi codewords, kG = 6 / 3 / 2 -> 10.3 / 9.5 / 9.4 ms.
constexpr int kLS = 18; // coefficient slots (17 used)
#ifndef BS_BM_GROUP
#define BS_BM_GROUP 2
#endif
constexpr int kG = BS_BM_GROUP;   // coefficients per group (loop body size: ~295 instructions each)
constexpr int kNG = kLS / kG;     // groups
static_assert(kG * kNG == kLS, "the group size must divide the number of slots");

BS_FN void rotate_groups(uint32_t (&L)[kLS][8])
{
#pragma unroll
    for (int j = 0; j < kG; j++)
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const uint32_t tmp = L[j][t];
#pragma unroll
            for (int g = 0; g + 1 < kNG; g++)
                L[j + kG * g][t] = L[j + kG * (g + 1)][t];
            L[j + kG * (kNG - 1)][t] = tmp;
        }
}

// ns = number of syndromes = BM iterations: 32 for RS(255,223); 16 / 8 for RS(255,239) / RS(255,247), whose register
// lengths stay <= ns / 2 (the counter k = r - 2 el declines everything longer, as for t = 16).
template <int STRIDE, int SO>
BS_FN void ribm(const Mem<SO> S, const Mem<STRIDE> B, uint32_t (&L)[kLS][8], uint32_t (&K)[7], int ns)
{
    uint32_t gamma[8], delta[8];
#pragma unroll
    for (int i = 0; i < kLS; i++)
#pragma unroll
        for (int t = 0; t < 8; t++) {
            L[i][t] = (i == 0 && t == 0) ? 0xFFFFFFFFu : 0u;
            if (i < kT)
                B.st(8 * i + t, (i == 0 && t == 0) ? 0xFFFFFFFFu : 0u);
        }
#pragma unroll
    for (int t = 0; t < 8; t++)
        gamma[t] = t == 0 ? 0xFFFFFFFFu : 0u;
#pragma unroll
    for (int b = 0; b < 7; b++)
        K[b] = 0;
    ld8(S, 0, delta); // Delta_1 = L_0 * S_1 = S_1

#pragma unroll 1
    for (int r = 1; r <= ns; r++) {
        const uint32_t chg = or8(delta) & ~K[6]; // delta != 0 and k >= 0
        const int top = r < kT ? r : kT;         // highest coefficient that can be non-zero after this iteration
        // the counters first (they depend on chg only): afterwards k = r - 2*el with el the NEW register length
        {
            uint32_t carry = 0xFFFFFFFFu;
#pragma unroll
            for (int b = 0; b < 7; b++) {
                const uint32_t inc = K[b] ^ carry;
                carry &= K[b];
                K[b] = merge_bits(~K[b], inc, chg);
            }
        }
        // Largest new register length among the warp's 1024 codewords: T_i = 0 beyond it for every one of them
        // (deg T <= el), so coefficient groups past it only shift B and skip the three products.
        const int elmax = (r - BS_WARP_MIN(min_counter(K))) >> 1;
        uint32_t acc[15], bcur[8];
        int full = 0; // coefficients 0 .. full-1 went through the complete update
        clear15(acc);
#pragma unroll
        for (int t = 0; t < 8; t++)
            bcur[t] = 0u; // "B_{-1}"
#pragma unroll 1
        for (int g = 0; g < kNG; g++) {
            // Every coefficient of an active group is updated, without branches (so that the six steps can be
            // scheduled into each other): beyond `top` both L_i and B_{i-1} are still zero and stay zero.
            if (kG * g <= top && kG * g <= elmax) {
#pragma unroll
                for (int j = 0; j < kG; j++) {
                    const int i = kG * g + j;
                    uint32_t bnext[8], s[8], tacc[15], T[8], nb[8];
                    const int d = r - i; // S_{r+1-i}; clamped where T_i is zero (i > r) or the sum is not used any more (r = 32)
                    const int sidx = d < 0 ? 0 : (d > kNS - 1 ? kNS - 1 : d);
#pragma unroll
                    for (int t = 0; t < 8; t++) {
                        bnext[t] = i < kT ? B.ld(8 * i + t) : 0u; // old B_i, wanted by the next step ("B_16" = 0)
                        s[t] = S.ld(8 * sidx + t);
                    }
                    clear15(tacc);
                    mac(tacc, gamma, L[j]);
                    mac(tacc, delta, bcur);
#pragma unroll
                    for (int t = 0; t < 8; t++)
                        nb[t] = merge_bits(L[j][t], bcur[t], chg);
                    if (i < kT)
                        st8(B, i, nb);
                    reduce(tacc, T);
                    cp8(L[j], T);
                    mac(acc, T, s);
                    cp8(bcur, bnext);
                }
                full = kG * g + kG;
            }
            rotate_groups(L);
        }
        // Coefficients full .. top were skipped: L_i = T_i = 0 there for all codewords, only B moves,
        // B_i <- (length change ? L_i = 0 : B_{i-1}).  A small rolled loop, highest index first (the loop body of the
        // groups above must stay inside the 32 KB instruction cache), bcur = old B_{full-1}.
        {
            const int hi = top < kT - 1 ? top : kT - 1;
#pragma unroll 1
            for (int i = hi; i > full; i--) {
#pragma unroll
                for (int t = 0; t < 8; t++)
                    B.st(8 * i + t, B.ld(8 * (i - 1) + t) & ~chg);
            }
            if (full <= hi) {
#pragma unroll
                for (int t = 0; t < 8; t++)
                    B.st(8 * full + t, bcur[t] & ~chg);
            }
        }
#pragma unroll
        for (int t = 0; t < 8; t++)
            gamma[t] = merge_bits(delta[t], gamma[t], chg);
        reduce(acc, delta);
    }
}

// ---- the same recurrence with the locator in memory -------------------------------------------------------------
// LM: 17 word blocks, L_i in block i.  One rolled loop over the coefficients 0 .. min(top, elmax) of an iteration:
// no coefficient registers to rotate (the register form spends ~15 % of the kernel's instructions on those moves),
// a loop body of one coefficient step, and ~100 live registers instead of ~250.  L_i and the syndrome of the NEXT
// step are fetched while the current step computes; the old B_i is only needed by the next step anyway.
#ifndef BS_BM_MEM
#define BS_BM_MEM 1
#endif
// gamma and Delta multiply every coefficient of an iteration: with the eight multiples gamma * x^s (and Delta * x^s)
// held as planes, a product is 64 AND-XOR with nothing left to reduce
#ifndef BS_BM_TABLES
#define BS_BM_TABLES 1
#endif
#ifndef BS_BM_UNROLL
#define BS_BM_UNROLL 1 // coefficient steps per trip of the loop
#endif
constexpr int kBmUnroll = BS_BM_UNROLL;
// times x modulo x^8 + x^4 + x^3 + x^2 + 1
BS_FN void xtime(const uint32_t (&a)[8], uint32_t (&o)[8])
{
    o[0] = a[7];
    o[1] = a[0];
    o[2] = a[1] ^ a[7];
    o[3] = a[2] ^ a[7];
    o[4] = a[3] ^ a[7];
    o[5] = a[4];
    o[6] = a[5];
    o[7] = a[6];
}
BS_FN void xmultiples(const uint32_t (&a)[8], uint32_t (&m)[8][8])
{
    cp8(m[0], a);
#pragma unroll
    for (int s = 1; s < 8; s++)
        xtime(m[s - 1], m[s]);
}
template <int STRIDE, int SO>
BS_FN void ribm_mem(const Mem<SO> S, const Mem<STRIDE> B, const Mem<SO> LM, uint32_t (&K)[7], int ns)
{
    uint32_t gamma[8], delta[8];
#pragma unroll 1
    for (int i = 0; i < kNL; i++)
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const uint32_t v = (i == 0 && t == 0) ? 0xFFFFFFFFu : 0u;
            LM.st(8 * i + t, v);
            if (i < kT)
                B.st(8 * i + t, v);
        }
#pragma unroll
    for (int t = 0; t < 8; t++)
        gamma[t] = t == 0 ? 0xFFFFFFFFu : 0u;
#pragma unroll
    for (int b = 0; b < 7; b++)
        K[b] = 0;
    ld8(S, 0, delta); // Delta_1 = L_0 * S_1 = S_1

#pragma unroll 1
    for (int r = 1; r <= ns; r++) {
        const uint32_t chg = or8(delta) & ~K[6]; // delta != 0 and k >= 0
        const int top = r < kT ? r : kT;         // highest coefficient that can be non-zero after this iteration
        {
            uint32_t carry = 0xFFFFFFFFu;
#pragma unroll
            for (int b = 0; b < 7; b++) {
                const uint32_t inc = K[b] ^ carry;
                carry &= K[b];
                K[b] = merge_bits(~K[b], inc, chg);
            }
        }
        // largest new register length among the warp's 1024 codewords: T_i = 0 beyond it for every one of them
        const int elmax = (r - BS_WARP_MIN(min_counter(K))) >> 1;
        const int lim = top < elmax ? top : elmax; // coefficients 0 .. lim take the complete update
        uint32_t acc[15], bcur[8], li[8], s[8];
        clear15(acc);
#pragma unroll
        for (int t = 0; t < 8; t++)
            bcur[t] = 0u; // "B_{-1}"
        ld8(LM, 0, li);
        ld8(S, r > kNS - 1 ? kNS - 1 : r, s);
#if BS_BM_TABLES
        uint32_t gx[8][8], dx[8][8];
        xmultiples(gamma, gx);
        xmultiples(delta, dx);
#endif
#pragma unroll kBmUnroll
        for (int i = 0; i <= lim; i++) {
            uint32_t bold[8], ln[8], sn[8], tacc[15], T[8];
            const int d = r - i - 1; // the next step's syndrome S_{r+1-(i+1)}, clamped where it is not used
            const int sidx = d < 0 ? 0 : (d > kNS - 1 ? kNS - 1 : d);
            const int bidx = i < kT ? i : kT - 1; // ("B_16" = 0: masked below)
#pragma unroll
            for (int t = 0; t < 8; t++) {
                ln[t] = LM.ld(8 * (i < kT ? i + 1 : kT) + t);
                sn[t] = S.ld(8 * sidx + t);
                bold[t] = B.ld(8 * bidx + t);
            }
#if BS_BM_TABLES
#pragma unroll
            for (int t = 0; t < 8; t++) {
                uint32_t v = li[0] & gx[0][t];
#pragma unroll
                for (int q = 1; q < 8; q++)
                    v ^= li[q] & gx[q][t];
#pragma unroll
                for (int q = 0; q < 8; q++)
                    v ^= bcur[q] & dx[q][t];
                T[t] = v;
            }
            (void)tacc;
            if (i < kT) {
#pragma unroll
                for (int t = 0; t < 8; t++)
                    B.st(8 * i + t, merge_bits(li[t], bcur[t], chg));
            }
#else
            clear15(tacc);
            mac(tacc, gamma, li);
            mac(tacc, delta, bcur);
            if (i < kT) {
#pragma unroll
                for (int t = 0; t < 8; t++)
                    B.st(8 * i + t, merge_bits(li[t], bcur[t], chg));
            }
            reduce(tacc, T);
#endif
            st8(LM, i, T);
            mac(acc, T, s);
            const uint32_t live = i < kT ? 0xFFFFFFFFu : 0u;
#pragma unroll
            for (int t = 0; t < 8; t++) {
                bcur[t] = bold[t] & live;
                li[t] = ln[t];
                s[t] = sn[t];
            }
        }
        // Coefficients lim+1 .. top were skipped: L_i = T_i = 0 there for all codewords, only B moves,
        // B_i <- (length change ? L_i = 0 : B_{i-1}); highest index first, bcur = old B_lim.
        {
            const int full = lim + 1;
            const int hi = top < kT - 1 ? top : kT - 1;
#pragma unroll 1
            for (int i = hi; i > full; i--) {
#pragma unroll
                for (int t = 0; t < 8; t++)
                    B.st(8 * i + t, B.ld(8 * (i - 1) + t) & ~chg);
            }
            if (full <= hi) {
#pragma unroll
                for (int t = 0; t < 8; t++)
                    B.st(8 * full + t, bcur[t] & ~chg);
            }
        }
#pragma unroll
        for (int t = 0; t < 8; t++)
            gamma[t] = merge_bits(delta[t], gamma[t], chg);
        reduce(acc, delta);
    }
}

// ---- evaluator: Omega_i = sum_{j <= i} S_{i-j+1} * L_j, i = 0..15 (src/decode.c:147-158) -------------------
// Rolled loops: L_0..L_15 are parked in P (16 blocks; the BM scratch, dead by now), Omega_i goes to word block i
// of O.
// (teff = ns / 2: for a code with fewer than 32 roots only Omega_0 .. Omega_(teff-1) exist; the rest is stored as zero)
template <int STRIDE, int SO>
BS_FN void omega_parked(const Mem<SO> S, const Mem<STRIDE> P, const Mem<SO> O, int teff, int jmax = 64);
template <int STRIDE, int SO>
BS_FN void omega(const Mem<SO> S, const uint32_t (&L)[kLS][8], const Mem<STRIDE> P, const Mem<SO> O, int teff)
{
#pragma unroll
    for (int j = 0; j < kT; j++)
        st8(P, j, L[j]);
    omega_parked<STRIDE, SO>(S, P, O, teff);
}
// P: L_0 .. L_15 in word blocks 0 .. 15 (jmax: the highest coefficient of P that can be non-zero, where the caller knows)
template <int STRIDE, int SO>
BS_FN void omega_parked(const Mem<SO> S, const Mem<STRIDE> P, const Mem<SO> O, int teff, int jmax)
{
#pragma unroll 1
    for (int i = teff; i < kT; i++) {
        const uint32_t z[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        st8(O, i, z);
    }
#pragma unroll 1
    for (int i = 0; i < teff; i++) {
        // the syndromes come from the slower memory: fetched two terms ahead (s, s1, s2 = S_{i-j}, S_{i-j-1}, S_{i-j-2})
        uint32_t acc[15], o[8], a[8], s[8], s1[8], s2[8];
        clear15(acc);
        ld8(P, 0, a);
        ld8(S, i, s);
        ld8(S, i >= 1 ? i - 1 : 0, s1);
        const int jtop = i < jmax ? i : jmax;
#pragma unroll 1
        for (int j = 0; j <= jtop; j++) {
            // unconditional loads with clamped indices (values past the end are never used)
            uint32_t an[8];
            ld8(S, j + 2 <= i ? i - j - 2 : 0, s2);
            ld8(P, j < i ? j + 1 : 0, an);
            mac(acc, a, s);
            cp8(a, an);
            cp8(s, s1);
            cp8(s1, s2);
        }
        reduce(acc, o);
        st8(O, i, o);
    }
}

// ---- 32x32 bit-matrix transpose: x[r] bit c <-> x[c] bit r (an involution) ---------------------------------
BS_FN void transpose32(uint32_t (&x)[32])
{
#pragma unroll
    for (int r = 0; r < 16; r++) {
        const uint32_t a = x[r], b = x[r + 16];
#if defined(__CUDA_ARCH__)
        x[r] = __byte_perm(a, b, 0x5410);
        x[r + 16] = __byte_perm(a, b, 0x7632);
#else
        x[r] = (a & 0x0000FFFFu) | (b << 16);
        x[r + 16] = (a >> 16) | (b & 0xFFFF0000u);
#endif
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 8)
            continue;
        const uint32_t a = x[r], b = x[r + 8];
#if defined(__CUDA_ARCH__)
        x[r] = __byte_perm(a, b, 0x6240);
        x[r + 8] = __byte_perm(a, b, 0x7351);
#else
        x[r] = (a & 0x00FF00FFu) | ((b << 8) & 0xFF00FF00u);
        x[r + 8] = ((a >> 8) & 0x00FF00FFu) | (b & 0xFF00FF00u);
#endif
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 4)
            continue;
        const uint32_t a = x[r], b = x[r + 4];
        x[r] = merge_bits(a, b << 4, 0x0F0F0F0Fu);
        x[r + 4] = merge_bits(a >> 4, b, 0x0F0F0F0Fu);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 2)
            continue;
        const uint32_t a = x[r], b = x[r + 2];
        x[r] = merge_bits(a, b << 2, 0x33333333u);
        x[r + 2] = merge_bits(a >> 2, b, 0x33333333u);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 1)
            continue;
        const uint32_t a = x[r], b = x[r + 1];
        x[r] = merge_bits(a, b << 1, 0x55555555u);
        x[r + 1] = merge_bits(a >> 1, b, 0x55555555u);
    }
}

// ---- Chien search as a linear recurring sequence ---------------------------------------------------------------
// Location k = 0..254 corresponds to the root alpha^(k+1) (src/decode.c:117-141 with prim = 1: k = i - 1), and the
// value examined there is  v_k = sum_{j=0..16} L_j * alpha^(j*(k+1)).  As a sequence in k this is a sum of 17
// geometric progressions, so its generating function is rational with a CONSTANT denominator:
//     V(z) = sum_k v_k z^k = sum_j A_j / (1 - b_j z) = N(z) / D(z),   A_j = L_j alpha^j,  b_j = alpha^j,
//     D(z) = prod_{j=0..16} (1 - alpha^j z)  (degree 17, D_0 = 1),   N(z) = sum_j A_j prod_{i != j} (1 - alpha^i z).
// Power-series division of N by D is the classic shift-register circuit: keep r (17 coefficients, r = N at the
// start); each step emits v = r_0 and replaces r by (r + v*D)/z, i.e. r_i <- r_{i+1} + v*D_{i+1} (r_17 = 0).
// One location therefore costs two 16-entry subset tables of the 8 planes of v (22 XOR) and ONE three-input XOR
// per state plane (136) - the same shape as the encoder's LFSR step - instead of 16 in-place constant multipliers
// plus a 17-term sum (~260 LOP3 + ~115 moves).  N is a constant linear map of the locator (17 x 17 field
// constants M[i][j] = alpha^j * [z^i] prod_{i' != j} (1 - alpha^i' z)), applied once per sweep.
struct ChienConsts {
    uint8_t d[18];     // D_0 .. D_17
    uint8_t m[17][17]; // N_i = sum_j m[i][j] * L_j
};

constexpr ChienConsts make_chien_consts()
{
    ChienConsts c{};
    const gf256::Poly all = gf256::roots_poly(0, 17); // prod (x - alpha^j), degree 17: D_i = all_{17-i}
    for (int i = 0; i <= 17; i++)
        c.d[i] = all.c[17 - i];
    for (int j = 0; j <= 16; j++) {
        gf256::Poly q = gf256::zero_poly(); // prod_{i != j} (x - alpha^i), degree 16; [z^i] of the reversed = q_{16-i}
        q.c[0] = 1;
        int deg = 0;
        for (int i = 0; i <= 16; i++) {
            if (i == j)
                continue;
            const uint8_t r = gf256::pow_alpha(i);
            for (int k = deg + 1; k > 0; k--)
                q.c[k] = q.c[k - 1] ^ gf256::mul(q.c[k], r);
            q.c[0] = gf256::mul(q.c[0], r);
            deg++;
        }
        for (int i = 0; i <= 16; i++)
            c.m[i][j] = gf256::mul(gf256::pow_alpha(j), q.c[16 - i]);
    }
    return c;
}
constexpr ChienConsts kChien = make_chien_consts();
static_assert(kChien.d[0] == 1, "D(0) = 1");

// a ^ b ^ c in one LOP3 (spelled out on the device: ptxas otherwise re-associates the table look-ups away)
BS_FN uint32_t xor3(uint32_t a, uint32_t b, uint32_t c)
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    return a ^ b ^ c;
#endif
}

// all 16 subset-XORs of four planes (11 XORs)
BS_FN void subset_table(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3, uint32_t (&T)[16])
{
    T[0] = 0;
    T[1] = b0;
    T[2] = b1;
    T[3] = b0 ^ b1;
    T[4] = b2;
    T[5] = b2 ^ b0;
    T[6] = b2 ^ b1;
    T[7] = b2 ^ T[3];
    T[8] = b3;
    T[9] = b3 ^ b0;
    T[10] = b3 ^ b1;
    T[11] = b3 ^ T[3];
    T[12] = b3 ^ b2;
    T[13] = T[12] ^ b0;
    T[14] = T[12] ^ b1;
    T[15] = T[12] ^ T[3];
}

constexpr int kNR = 17; // state bytes of the sequence generator

template <int IT>
BS_FN void chien_update_one(uint32_t (&R)[kNR][8], const uint32_t (&T1)[16], const uint32_t (&T2)[16])
{
    constexpr int I = IT >> 3, T = IT & 7;
    constexpr unsigned m = gf256::rowmask(kChien.d[I + 1], T);
    if constexpr (I < kNR - 1)
        R[I][T] = xor3(R[I + 1][T], T1[m & 15], T2[m >> 4]);
    else
        R[I][T] = T1[m & 15] ^ T2[m >> 4];
}
template <int... IT>
BS_FN void chien_update_all(uint32_t (&R)[kNR][8], const uint32_t (&T1)[16], const uint32_t (&T2)[16],
                            std::integer_sequence<int, IT...>)
{
    (chien_update_one<IT>(R, T1, T2), ...);
}

// one location: returns the mask of codewords whose locator vanishes there, and advances the generator
BS_FN uint32_t chien_step(uint32_t (&R)[kNR][8])
{
    const uint32_t nz = or8(R[0]);
    uint32_t T1[16], T2[16];
    subset_table(R[0][0], R[0][1], R[0][2], R[0][3], T1);
    subset_table(R[0][4], R[0][5], R[0][6], R[0][7], T2);
    chien_update_all(R, T1, T2, std::make_integer_sequence<int, 8 * kNR>{});
    return ~nz;
}

// R += M[.][J] * L_J for one locator coefficient (8 planes in l)
template <int J, int IT>
BS_FN void chien_init_one(uint32_t (&R)[kNR][8], const uint32_t (&T1)[16], const uint32_t (&T2)[16])
{
    constexpr int I = IT >> 3, T = IT & 7;
    constexpr unsigned m = gf256::rowmask(kChien.m[I][J], T);
    if constexpr (J == 0)
        R[I][T] = T1[m & 15] ^ T2[m >> 4];
    else
        R[I][T] = xor3(R[I][T], T1[m & 15], T2[m >> 4]);
}
template <int J, int... IT>
BS_FN void chien_init_all(uint32_t (&R)[kNR][8], const uint32_t (&T1)[16], const uint32_t (&T2)[16],
                          std::integer_sequence<int, IT...>)
{
    (chien_init_one<J, IT>(R, T1, T2), ...);
}
// LM: the locator, word block j = L_j (17 blocks).  The next coefficient is loaded while the current one is used.
template <int J, int SL>
BS_FN void chien_init_from(const Mem<SL> LM, uint32_t (&R)[kNR][8], const uint32_t (&l)[8])
{
    if constexpr (J < kNR) {
        uint32_t ln[8], T1[16], T2[16];
        if constexpr (J + 1 < kNR)
            ld8(LM, J + 1, ln);
        BS_FENCE();
        subset_table(l[0], l[1], l[2], l[3], T1);
        subset_table(l[4], l[5], l[6], l[7], T2);
        chien_init_all<J>(R, T1, T2, std::make_integer_sequence<int, 8 * kNR>{});
        if constexpr (J + 1 < kNR)
            chien_init_from<J + 1, SL>(LM, R, ln);
    }
}

// ---- memory map of the work area W (200 words per lane) ---------------------------------------------------------
//   [0, 32)     scratch: root masks of 32 locations / x planes of four list slots / magnitude words
//   [32, 160)   root lists: word 32 + 4c + (e >> 2), byte e & 3 = location + 1 of the e-th root of codeword c (0 = none)
//   [160, 168)  root counters, one byte per codeword
//   [168, 200)  parked root masks of the previous block of 32 locations
constexpr int kWordsS = 8 * kNS;  // 256: the syndromes
constexpr int kWordsO = 8 * kT;   // 128: the evaluator coefficients
constexpr int kWordsW = 200;      // the work area
constexpr int kListBase = 32;
constexpr int kCntBase = 160;

struct LaneOut {
    uint32_t K[7];       // k = 32 - 2*L per codeword (bit-planes, two's complement)
    uint32_t nzL[kNL];   // nzL[j] bit c: locator coefficient j of codeword c is non-zero
    uint32_t padfail;    // bit c: a root fell into the zero padding (k < pad), src/decode.c:132-134
};

#if defined(__CUDA_ARCH__)
#define BS_CTZ(x) (__ffs((int)(x)) - 1)
#else
#define BS_CTZ(x) __builtin_ctz(x)
#endif

// The sweep runs in blocks of 32 locations: the 32 root masks of a block are transposed into one word per
// codeword (bit kk = "location 32*blk + kk is a root") and parked in the S area; while the NEXT block is being
// swept, each location step also appends the roots of one codeword of the parked block to that codeword's list
// (the list work is all latency - find-first-set, table look-up, byte store - and hides behind the XOR networks
// of the sweep).  exp_tab[i] = alpha^i, i = 0..255.
constexpr int kParkBase = 168; // 32 parked mask words

// bits: the parked mask word of codeword c, n: its root counter (both loaded by the caller, ahead of use).
// List entries are location + 1 (1..255; 0 = empty slot).
#ifndef BS_APPEND_FLAT
#define BS_APPEND_FLAT 2
#endif
constexpr int kAppendFlat = BS_APPEND_FLAT;
template <int STRIDE>
BS_FN void append_roots(const Mem<STRIDE> S, int c, int blk, uint32_t bits, uint32_t n)
{
    const uint32_t n0 = n;
    // a codeword has about one root per block: the first kAppendFlat are taken without branches
#pragma unroll
    for (int it = 0; it < kAppendFlat; it++) {
        const bool live = bits != 0;
        const int kk = live ? BS_CTZ(bits) : 0;
        if (live && n < (uint32_t)kT)
            S.st_byte(kListBase + 4 * c + (int)(n >> 2), (int)(n & 3u), (uint32_t)(32 * blk + kk + 1));
        n += live ? 1u : 0u;
        bits &= bits - 1;
    }
    while (bits) {
        const int kk = BS_CTZ(bits);
        bits &= bits - 1;
        if (n < (uint32_t)kT)
            S.st_byte(kListBase + 4 * c + (int)(n >> 2), (int)(n & 3u), (uint32_t)(32 * blk + kk + 1));
        n++;
    }
    if (n != n0)
        S.st_byte(kCntBase + (c >> 2), c & 3, n);
}

// A hook that does nothing (host tests).  The device passes one that issues the PREVIOUS tile's corrections in
// small portions: before(step) starts the loads of one portion, after(step) consumes them, 256 steps per sweep.
struct NoHook {
    BS_FN void begin() {}
    BS_FN void before(int) {}
    BS_FN void after(int) {}
    BS_FN void end() {}
    // codeword-wide form (transform-form Chien search): wstep(c), c = 0..32, handles the whole record of codeword c
    BS_FN void wbegin() {}
    BS_FN void wstep(int) {}
    BS_FN void wend() {}
};

template <int STRIDE, int SL, class Hook>
BS_FN void chien_collect(const Mem<SL> LM, const Mem<STRIDE> S, int pad, LaneOut &out, Hook &hook)
{
    uint32_t R[kNR][8];
    {
        uint32_t l0[8];
        ld8(LM, 0, l0);
        chien_init_from<0, SL>(LM, R, l0);
    }
    for (int w = kListBase; w < kParkBase + 32; w++)
        S.st(w, 0u);
    hook.begin();
    uint32_t padfail = 0;
#pragma unroll 1
    for (int blk = 0; blk < 8; blk++) {
#pragma unroll 1
        for (int kk = 0; kk < 32; kk++) {
            const int k = 32 * blk + kk;
            // parked block (all-zero masks while blk == 0): operands fetched before the sweep step, used after it
            const uint32_t bits = S.ld(kParkBase + kk);
            const uint32_t n = S.ld_byte(kCntBase + (kk >> 2), kk & 3);
            hook.before(k);
            uint32_t m = 0;
            if (k < 255) {
                m = chien_step(R);
                if (k < pad)
                    padfail |= m;
            }
            hook.after(k);
            S.st(kk, m);
            append_roots<STRIDE>(S, kk, blk - 1, bits, n);
        }
        uint32_t X[32];
#pragma unroll
        for (int c = 0; c < 32; c++)
            X[c] = S.ld(c);
        transpose32(X);
#pragma unroll
        for (int c = 0; c < 32; c++)
            S.st(kParkBase + c, X[c]);
    }
#pragma unroll 1
    for (int c = 0; c < 32; c++)
        append_roots<STRIDE>(S, c, 7, S.ld(kParkBase + c), S.ld_byte(kCntBase + (c >> 2), c & 3));
    hook.end();
    out.padfail = padfail;
}

// ---- Chien search as a 15 x 17 transform -----------------------------------------------------------------------
// Lambda(alpha^j) for ALL 255 j from the 17 locator coefficients at ~14 K plane-operations per 32 codewords instead of
// 255 x 158 = 40 K for the sequence generator above (tools/gen_rs255_fft.py derives and checks the algebra):
//   j <-> (u = j mod 17, w = j mod 15);   alpha^(j d) = beta^(8 w d) gamma^(8 u d),  beta = alpha^17, gamma = alpha^15
//   Lambda(alpha^j) = sum_{d < 15} beta^(8 w d) Z_d(u),   Z_d(u) = gamma^(8 u d) (Lambda_d + [d < 2] gamma^u Lambda_(d+15))
// so every u is one 15-point DFT with GF(16) twiddles (rsfft::dft15 on the two 4-plane halves of tower-basis symbols;
// its output k belongs to j = 120 u + 17 k mod 255) of fifteen values that advance from u to u + 1 by constant
// multipliers (rsfft::cmulT, 12 operations each on average).  Location k = j - 1 (mod 255) is a root where all eight
// planes vanish.  The 255 root masks go to MK (256 words per lane, word 255 = 0); afterwards they are transposed in
// blocks of 32 locations into the per-codeword root lists exactly as above (the order of a list's entries differs -
// nothing depends on it: every root is corrected independently).  The hook gets its 256 steps during the transform.
static_assert(rsfft::kChienJ[1] == 17 && rsfft::kChienJ[7] == 119 && rsfft::kChienJ[14] == 238, "kChienJ[k] = 17 k");
BS_FN uint32_t or4(const uint32_t (&a)[4])
{
    return (a[0] | a[1]) | (a[2] | a[3]);
}

template <int D>
BS_FN void chien_advance(uint32_t (&Z)[13][8])
{
    if constexpr (D < 13) {
        rsfft::cmulT<rsfft::kChienZ[D]>(Z[D]);
        chien_advance<D + 1>(Z);
    }
}

template <int STRIDE, int SL, class Hook>
BS_FN void chien_transform(const Mem<SL> LM, const Mem<SL> MK, const Mem<STRIDE> S, int pad, LaneOut &out, Hook &hook)
{
    {
        uint32_t Z[13][8], A0[8], Tb[8], Tc[8], Td[8];
        {
            uint32_t l[8];
            ld8(LM, 0, l);
            rsfft::conv_in(l, A0);
            ld8(LM, 15, l);
            rsfft::conv_in(l, Tb);
            ld8(LM, 1, l);
            rsfft::conv_in(l, Tc);
            ld8(LM, 16, l);
            rsfft::conv_in(l, Td);
#pragma unroll
            for (int d = 2; d < 15; d++) {
                ld8(LM, d, l);
                rsfft::conv_in(l, Z[d - 2]);
            }
        }
        for (int w = kListBase; w < kParkBase + 32; w++)
            S.st(w, 0u);
        hook.wbegin();
        int jb = 0; // jb = 120 u mod 255
#pragma unroll 1
        for (int u = 0; u < 17; u++) {
            uint32_t acc[15];
            {
                uint32_t x[15][4];
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    x[0][i] = A0[i] ^ Tb[i];
                    x[1][i] = Tc[i] ^ Td[i];
                }
#pragma unroll
                for (int d = 2; d < 15; d++)
#pragma unroll
                    for (int i = 0; i < 4; i++)
                        x[d][i] = Z[d - 2][i];
                rsfft::dft15(x);
#pragma unroll
                for (int k = 0; k < 15; k++)
                    acc[k] = or4(x[k]);
            }
            hook.wstep(2 * u); // (the previous tile's corrections: two codewords per step, loads one step ahead of use)
            {
                uint32_t x[15][4];
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    x[0][i] = A0[4 + i] ^ Tb[4 + i];
                    x[1][i] = Tc[4 + i] ^ Td[4 + i];
                }
#pragma unroll
                for (int d = 2; d < 15; d++)
#pragma unroll
                    for (int i = 0; i < 4; i++)
                        x[d][i] = Z[d - 2][4 + i];
                rsfft::dft15(x);
#pragma unroll
                for (int k = 0; k < 15; k++)
                    acc[k] |= or4(x[k]);
            }
            hook.wstep(2 * u + 1);
#pragma unroll
            for (int k = 0; k < 15; k++) {
                int j = jb + 17 * k; // rsfft::kChienJ[k]
                j = j >= 255 ? j - 255 : j;
                const int loc = j == 0 ? 254 : j - 1;
                MK.st(loc, ~acc[k]);
            }
            jb += 120;
            jb = jb >= 255 ? jb - 255 : jb;
            chien_advance<0>(Z);
            rsfft::cmulT<rsfft::kChienB>(Tb);
            rsfft::cmulT<rsfft::kChienC>(Tc);
            rsfft::cmulT<rsfft::kChienD>(Td);
        }
        MK.st(255, 0u);
        hook.wend();
    }
    uint32_t padfail = 0;
    uint32_t X[32];
#pragma unroll
    for (int c = 0; c < 32; c++)
        X[c] = MK.ld(c);
#pragma unroll 1
    for (int blk = 0; blk < 8; blk++) {
        const int npad = pad - 32 * blk; // locations below the padding boundary in this block
#pragma unroll
        for (int c = 0; c < 32; c++)
            padfail |= c < npad ? X[c] : 0u;
        transpose32(X);
#pragma unroll
        for (int c = 0; c < 32; c++)
            S.st(kParkBase + c, X[c]);
        // the next block's masks are on their way while this block's roots are appended
        const int nb = blk < 7 ? blk + 1 : 7;
#pragma unroll
        for (int c = 0; c < 32; c++)
            X[c] = MK.ld(32 * nb + c);
        // four codewords per trip: their (latency-bound) list updates overlap
#pragma unroll 1
        for (int c0 = 0; c0 < 32; c0 += 4) {
            uint32_t bits[4], cnt[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                bits[q] = S.ld(kParkBase + c0 + q);
                cnt[q] = S.ld_byte(kCntBase + (c0 >> 2), q);
            }
#pragma unroll
            for (int q = 0; q < 4; q++)
                append_roots<STRIDE>(S, c0 + q, blk, bits[q], cnt[q]);
        }
    }
    out.padfail = padfail;
}

// ---- Forney on four list slots at a time -----------------------------------------------------------------------
// x = the root alpha^(k+1) found for each of the 32 codewords (planes).  Magnitude (src/decode.c:159-191,
// fcr = 1 so num2 = 1):  Omega(x) / Lambda'(x)  =  x*Omega(x) / (sum of the odd-degree locator terms at x).
// A vanishing numerator gives magnitude 0 (the reference does not count such a root either); an empty slot
// (x = 0) gives 0 as well.
//
// List slots 4w .. 4w+3 of all 32 codewords: afterwards scratch words 8y .. 8y+7 of S hold the magnitude planes
// of slot 4w + y, and cnt (5 bit-planes) has counted the non-zero magnitudes.  magnitude_words() then gives
// one word per codeword (byte y = magnitude of slot 4w + y).
// The two Horner recurrences run over the four slots in lock step: one coefficient (fetched from the L2-resident
// scratch one step ahead) feeds four independent multiply-reduce chains, so its latency hides behind ~340
// instructions and the chains overlap each other's dependent steps.
#ifndef BS_FORNEY_TABLES
#define BS_FORNEY_TABLES 1
#endif
// v <- v * m + c, with m given as its multiples m * alpha^s (xmultiples): plane t of the product is the sum over s of
// v[s] & mm[s][t]
BS_FN void mul_add_tab(uint32_t (&v)[8], const uint32_t (&mm)[8][8], const uint32_t (&c)[8])
{
    uint32_t r[8];
#pragma unroll
    for (int t = 0; t < 8; t++) {
        uint32_t a = c[t] ^ (v[0] & mm[0][t]);
#pragma unroll
        for (int q = 1; q < 8; q++)
            a ^= v[q] & mm[q][t];
        r[t] = a;
    }
    cp8(v, r);
}

template <int STRIDE, int SO, int NC>
BS_FN void forney_math(const Mem<SO> LM, const Mem<SO> O, const Mem<STRIDE> S, uint32_t (&cnt)[NC], int teff);
template <int STRIDE, int SO>
BS_FN void forney_slots(const Mem<SO> LM, const Mem<SO> O, const Mem<STRIDE> S, int w,
                        const uint8_t *root_tab /* [0] = 0, [v] = alpha^v */, uint32_t (&cnt)[5], int teff = kT)
{
    {
        uint32_t X[32];
#pragma unroll
        for (int c = 0; c < 32; c++) {
            const uint32_t v = S.ld(kListBase + 4 * c + w); // four list entries (location + 1, 0 = empty)
            X[c] = (uint32_t)root_tab[v & 0xFFu] | ((uint32_t)root_tab[(v >> 8) & 0xFFu] << 8) |
                   ((uint32_t)root_tab[(v >> 16) & 0xFFu] << 16) | ((uint32_t)root_tab[v >> 24] << 24);
        }
        transpose32(X); // X[8*y + t] = plane t of the roots in slot 4w + y
#pragma unroll
        for (int p = 0; p < 32; p++)
            S.st(p, X[p]);
    }
    forney_math<STRIDE, SO, 5>(LM, O, S, cnt, teff);
}
// The roots of four slots are staged as planes in words 0 .. 31 of S (word 8 y + t = plane t of slot y); cnt: NC
// bit-planes counting the non-zero magnitudes.
template <int STRIDE, int SO, int NC>
BS_FN void forney_math(const Mem<SO> LM, const Mem<SO> O, const Mem<STRIDE> S, uint32_t (&cnt)[NC], int teff)
{
    // teff = ns / 2 (16, 8 or 4; 32 for the erasure form): Omega has teff coefficients and the locator's highest odd
    // coefficient is teff - 1, so the two Horner chains start there
#if BS_FORNEY_TABLES
    // Both Horner chains of a slot multiply by the same element at every step (x, then x^2): with its eight multiples
    // x * alpha^s held as planes a step is 64 AND-XOR, the coefficient folded into the first of them and nothing to
    // reduce.  Two slots at a time (128 registers of multiples).
#pragma unroll 1
    for (int y0 = 0; y0 < 4; y0 += 2) {
        uint32_t x[2][8], xm[2][8][8], v[2][8], o[8];
        ld8(O, teff - 1, o);
#pragma unroll
        for (int y = 0; y < 2; y++) {
            ld8(S, y0 + y, x[y]);
            xmultiples(x[y], xm[y]);
            cp8(v[y], o);
        }
        // (the coefficients come from the L2: a step on held multiples is shorter than that round trip, so they are
        // fetched two steps ahead)
        uint32_t o1[8];
        ld8(O, teff - 2, o);
        ld8(O, teff >= 3 ? teff - 3 : 0, o1);
#pragma unroll 1
        for (int j = teff - 2; j >= 0; j--) {
            uint32_t on[8];
            ld8(O, j > 1 ? j - 2 : 0, on);
#pragma unroll
            for (int y = 0; y < 2; y++)
                mul_add_tab(v[y], xm[y], o);
            cp8(o, o1);
            cp8(o1, on);
        }
        {
            const uint32_t z[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
#pragma unroll
            for (int y = 0; y < 2; y++) {
                mul_add_tab(v[y], xm[y], z); // one more factor x
#pragma unroll
                for (int t = 0; t < 8; t++) // parked in the (idle) parked-mask words until the division
                    S.st(kParkBase + 8 * (y0 + y) + t, v[y][t]);
            }
        }
        ld8(LM, teff - 1, o);
#pragma unroll
        for (int y = 0; y < 2; y++) {
            uint32_t x2[8];
            sqr(x[y], x2);
            xmultiples(x2, xm[y]);
            cp8(v[y], o);
        }
        ld8(LM, teff - 3, o);
        ld8(LM, teff >= 6 ? teff - 5 : 1, o1);
#pragma unroll 1
        for (int j = teff - 3; j >= 1; j -= 2) {
            uint32_t on[8];
            ld8(LM, j > 3 ? j - 4 : 1, on);
#pragma unroll
            for (int y = 0; y < 2; y++)
                mul_add_tab(v[y], xm[y], o);
            cp8(o, o1);
            cp8(o1, on);
        }
#pragma unroll
        for (int y = 0; y < 2; y++) {
            mul(v[y], x[y], v[y]);
            st8(S, y0 + y, v[y]);
        }
    }
#else
    {
        uint32_t x[4][8], num[4][8], o[8];
        ld8(O, teff - 1, o);
#pragma unroll
        for (int y = 0; y < 4; y++) {
            ld8(S, y, x[y]);
            cp8(num[y], o);
        }
        // x * Omega(x): Horner from Omega_(teff-1) down, then one more factor x
        ld8(O, teff - 2, o);
#pragma unroll 1
        for (int j = teff - 2; j >= 0; j--) {
            // (unconditional, index clamped: a conditionally loaded array is demoted to local memory and its
            // spill store then waits for the load)
            uint32_t on[8];
            ld8(O, j > 0 ? j - 1 : 0, on);
#pragma unroll
            for (int y = 0; y < 4; y++) {
                uint32_t acc[15];
                clear15(acc);
                cp8(*reinterpret_cast<uint32_t(*)[8]>(&acc[0]), o);
                mac(acc, num[y], x[y]);
                reduce(acc, num[y]);
            }
            cp8(o, on);
        }
#pragma unroll
        for (int y = 0; y < 4; y++) {
            mul(num[y], x[y], num[y]);
            for (int t = 0; t < 8; t++) // parked in the (idle) parked-mask words until the division
                S.st(kParkBase + 8 * y + t, num[y][t]);
        }
        // x * Lambda'(x) = x * sum over odd j of L_j x^(j-1): Horner in x^2 from L_15 down
        uint32_t x2[4][8], odd[4][8];
        ld8(LM, teff - 1, o);
#pragma unroll
        for (int y = 0; y < 4; y++) {
            sqr(x[y], x2[y]);
            cp8(odd[y], o);
        }
        ld8(LM, teff - 3, o);
#pragma unroll 1
        for (int j = teff - 3; j >= 1; j -= 2) {
            uint32_t on[8];
            ld8(LM, j > 1 ? j - 2 : 1, on);
#pragma unroll
            for (int y = 0; y < 4; y++) {
                uint32_t acc[15];
                clear15(acc);
                cp8(*reinterpret_cast<uint32_t(*)[8]>(&acc[0]), o);
                mac(acc, odd[y], x2[y]);
                reduce(acc, odd[y]);
            }
            cp8(o, on);
        }
        // denominators over the x planes
#pragma unroll
        for (int y = 0; y < 4; y++) {
            mul(odd[y], x[y], odd[y]);
            st8(S, y, odd[y]);
        }
    }
#endif
#pragma unroll 1
    for (int y = 0; y < 4; y++) {
        uint32_t num[8], den[8], mag[8];
        ld8(S, y, den);
#pragma unroll
        for (int t = 0; t < 8; t++)
            num[t] = S.ld(kParkBase + 8 * y + t);
        inv(den, den);
        mul(num, den, mag);
        st8(S, y, mag);
        uint32_t m = or8(mag);
#pragma unroll
        for (int b = 0; b < NC; b++) {
            const uint32_t t = cnt[b] & m;
            cnt[b] ^= m;
            m = t;
        }
    }
}

template <int STRIDE>
BS_FN void magnitude_words(const Mem<STRIDE> S, uint32_t (&X)[32])
{
#pragma unroll
    for (int p = 0; p < 32; p++)
        X[p] = S.ld(p);
    transpose32(X);
}

enum { kDeclined = 0, kFail = 1, kOk = 2 };

// Outcome of codeword c of a lane (see the header comment for the rules).  nroots = roots found.
template <int STRIDE>
BS_FN int classify(const Mem<STRIDE> S, const LaneOut &o, int c, int &nroots, int ns = kNS)
{
    int k = 0;
#pragma unroll
    for (int b = 0; b < 7; b++)
        k |= (int)((o.K[b] >> c) & 1u) << b;
    if (k & 64)
        return kDeclined; // k < 0: L > 16
    const int el = (ns - k) >> 1;
    uint32_t lead = 0;
#pragma unroll
    for (int j = 0; j < kNL; j++)
        if (j == el)
            lead = (o.nzL[j] >> c) & 1u;
    if (!lead || el == 0)
        return kDeclined; // degree below the register length: leave it to the literal stage
    nroots = (int)S.ld_byte(kCntBase + (c >> 2), c & 3);
    if (nroots != el || ((o.padfail >> c) & 1u))
        return kFail;
    return kOk;
}

// key equation + root search for one lane: afterwards classify() and forney_slots() may be used.
//   Syn: the 32 syndromes (256 words), read-only.   Om: receives the evaluator (128 words).
//   W: work area of kWordsW words: scratch / root lists / counters / parked masks (map above); its words
//      [kListBase, kListBase + 128) double as the BM scratch and as the parking space of the evaluator loop.
// Syn and Om may live in slower memory than W (on the device: an L2-resident scratch in global memory, so that
// a warp needs 25 KB of shared memory and eight warps fit one SM).
//   Mk: 256 words for the root masks of the transform form of the Chien search (same kind of memory as Syn).
#ifndef BS_CHIEN_TRANSFORM
#define BS_CHIEN_TRANSFORM 1
#endif
constexpr int kWordsM = 256;

template <int STRIDE, int SO, class Hook>
BS_FN void solve_lane(const Mem<SO> Syn, const Mem<SO> Om, const Mem<SO> Mk, const Mem<STRIDE> W, int pad, LaneOut &out,
                      Hook &hook, int ns = kNS)
{
    const Mem<STRIDE> B{W.base + kListBase * STRIDE};
#if BS_BM_MEM
    {
        // the locator lives in the root-mask area (free until the sweep) during BM and the evaluator
        ribm_mem<STRIDE, SO>(Syn, B, Mk, out.K, ns);
        omega_parked<SO, SO>(Syn, Mk, Om, ns >> 1);
#pragma unroll
        for (int j = 0; j < kNL; j++) {
            uint32_t l[8];
            ld8(Mk, j, l);
            out.nzL[j] = or8(l);
            st8(Syn, j, l);
        }
    }
#else
    {
        uint32_t L[kLS][8];
        ribm<STRIDE, SO>(Syn, B, L, out.K, ns);
        omega<STRIDE, SO>(Syn, L, B, Om, ns >> 1);
        // the syndromes are dead now: their first 17 word blocks receive the locator (read by the sweep set-up
        // and by Forney)
#pragma unroll
        for (int j = 0; j < kNL; j++) {
            out.nzL[j] = or8(L[j]);
            st8(Syn, j, L[j]);
        }
    }
#endif
#if BS_CHIEN_TRANSFORM
    chien_transform<STRIDE, SO, Hook>(Syn, Mk, W, pad, out, hook);
#else
    (void)Mk;
    chien_collect<STRIDE, SO, Hook>(Syn, W, pad, out, hook);
#endif
}

} // namespace bs
