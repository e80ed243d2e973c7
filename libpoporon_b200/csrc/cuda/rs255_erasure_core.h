// rs255_erasure_core.h — bitsliced erasure-only decoding for RS(255,223)/0x11D, fcr = 1, prim = 1.
//
// Replaces the reference's error_correction_u8 (src/decode.c:17-230) for rows of erasure mode whose errors all sit
// on the declared erasure positions - whatever their number (1 .. 32).  For such a row the reference's path is:
//   locator initialised to Gamma(x) = prod (1 - X_i x) over the declared positions (src/decode.c:31-47);
//   Berlekamp-Massey from r = count + 1 finds every discrepancy zero and changes nothing (:51-96);
//   the Chien search finds the declared positions, in ascending order of location (:112-145);
//   omega = S * Gamma mod x^count (:147-158), Forney (:159-191), the re-syndrome check passes (:193-209), and the
//   i-th root's magnitude goes to the i-th declared slot (:211-214) - the root's own position when the declared
//   positions are strictly ascending.
// The discrepancy of iteration r is the coefficient r - 1 of the same product S * Gamma, so one triangular product
//   P_n = sum_{j <= n} Gamma_j S_(n-j+1),  n = 0 .. 31
// gives both: P_0 .. P_(count-1) is omega, and the row qualifies exactly when P_count .. P_31 all vanish (P is
// computed in place of the syndromes: P_n replaces S_(n+1)).  For a
// qualifying row those are zero, and so are Gamma's coefficients above count: the Horner chains of Forney's formula
// can run over all 32 coefficients for every codeword of a lane alike.  No Berlekamp-Massey, no Chien search.
//
// 32 codewords per thread on bit-planes, as in rs255_solve_core.h (whose helpers this uses); plain C++, so that the
// same code runs on the host for tests (tests/native/host_solver.cpp).
#pragma once
#include "rs255_solve_core.h"

namespace bs {

constexpr int kES = 32;            // erasure slots per codeword
constexpr int kWordsG = 8 * 33;    // Gamma_0 .. Gamma_32
constexpr int kWordsEL = 8 * 32;   // slot lists: word 8 c + w = slots 4 w .. 4 w + 3 of codeword c, a byte each:
                                   // location + 1 (location = position + padding), 0 = slot not declared
constexpr int kActBase = 32;       // words 32 .. 63 of the work area: act[i] bit c = slot i of codeword c is declared

// the elements tab[list byte] of slots 4 w .. 4 w + 3 as planes in words 0 .. 31 of S (word 8 y + t = plane t of slot y)
template <int STRIDE, int SO>
BS_FN void stage_slots(const Mem<SO> EL, const Mem<STRIDE> S, int w, const uint8_t *tab)
{
    uint32_t X[32];
#pragma unroll
    for (int c = 0; c < 32; c++) {
        const uint32_t v = EL.ld(8 * c + w);
        X[c] = (uint32_t)tab[v & 0xFFu] | ((uint32_t)tab[(v >> 8) & 0xFFu] << 8) |
               ((uint32_t)tab[(v >> 16) & 0xFFu] << 16) | ((uint32_t)tab[v >> 24] << 24);
    }
    transpose32(X);
#pragma unroll
    for (int p = 0; p < 32; p++)
        S.st(p, X[p]);
}

// Gamma(x) = prod over the declared slots of (1 + X_i x), X_i = alpha^(254 - location) = inv_tab[list byte]
// (inv_tab[0] = 0: an undeclared slot multiplies by 1).  G: 33 word blocks.  One factor at a time: Gamma_k +=
// X_i Gamma_(k-1), k = i + 1 .. 1, on the eight held multiples of X_i.
// nslots: no codeword of the lane (of the warp, on the device) declares more slots than that
template <int STRIDE, int SO>
BS_FN void erasure_locator(const Mem<SO> EL, const Mem<SO> G, const Mem<STRIDE> S, const uint8_t *inv_tab, int nslots = kES)
{
#pragma unroll 1
    for (int i = 0; i < kES; i++)
        S.st(kActBase + i, 0u);
#pragma unroll 1
    for (int k = 0; k <= kES; k++)
#pragma unroll
        for (int t = 0; t < 8; t++)
            G.st(8 * k + t, (k == 0 && t == 0) ? 0xFFFFFFFFu : 0u);
#pragma unroll 1
    for (int w = 0; 4 * w < nslots; w++) {
        stage_slots<STRIDE, SO>(EL, S, w, inv_tab);
#pragma unroll 1
        for (int y = 0; y < 4; y++) {
            const int i = 4 * w + y;
            uint32_t X[8], xm[8][8], lo[8];
            ld8(S, y, X);
            S.st(kActBase + i, or8(X));
            xmultiples(X, xm);
            // step k: Gamma_k += X_i Gamma_(k-1).  The next step's Gamma_k is this step's Gamma_(k-1), still unchanged
            // and in registers: one block is loaded per step, two steps ahead of its use.
            uint32_t cur[8], nx1[8];
            ld8(G, i + 1, cur);
            ld8(G, i, lo);
            ld8(G, i >= 1 ? i - 1 : 0, nx1);
#pragma unroll 1
            for (int k = i + 1; k >= 1; k--) {
                uint32_t nx2[8];
                ld8(G, k >= 3 ? k - 3 : 0, nx2); // Gamma_(k-3), the multiplicand of step k - 2
#pragma unroll
                for (int t = 0; t < 8; t++) {
                    uint32_t a = cur[t];
#pragma unroll
                    for (int q = 0; q < 8; q++)
                        a ^= lo[q] & xm[q][t];
                    cur[t] = a;
                }
                st8(G, k, cur);
                cp8(cur, lo);
                cp8(lo, nx1);
                cp8(nx1, nx2);
            }
        }
    }
}

// P_n = sum_{j <= min(n, jmax)} Gamma_j S_(n-j+1), n = 31 .. 0, IN PLACE: P_n replaces S_(n+1) (word block n of SP), which
// no later (smaller) n reads.  Halves the scratch of the product - what keeps the kernel's state inside the L2.
template <int SO>
BS_FN void erasure_product(const Mem<SO> SP, const Mem<SO> G, int jmax, int ns = kES)
{
#pragma unroll 1
    for (int n = ns - 1; n >= 0; n--) {
        const int jtop = n < jmax ? n : jmax;
        uint32_t acc[15], o[8], g[8], s[8], g1[8], s1[8];
        clear15(acc);
        ld8(G, 0, g);
        ld8(SP, n, s);
        ld8(G, jtop >= 1 ? 1 : 0, g1);
        ld8(SP, n >= 1 ? n - 1 : 0, s1);
#pragma unroll 1
        for (int j = 0; j <= jtop; j++) {
            // (two terms ahead, indices clamped: values past the end are never used)
            uint32_t g2[8], s2[8];
            ld8(G, j + 2 <= jtop ? j + 2 : 0, g2);
            ld8(SP, n - j >= 2 ? n - j - 2 : 0, s2);
            mac(acc, g, s);
            cp8(g, g1);
            cp8(s, s1);
            cp8(g1, g2);
            cp8(s1, s2);
        }
        reduce(acc, o);
        st8(SP, n, o);
    }
}

// bit c set: codeword c does NOT qualify (some P_n with n >= its declared count is non-zero)
template <int STRIDE, int SO>
BS_FN uint32_t erasure_residue(const Mem<SO> P, const Mem<STRIDE> S, int ns = kES)
{
    uint32_t bad = 0;
#pragma unroll 1
    for (int n = 0; n < ns; n++) {
        uint32_t p[8];
        ld8(P, n, p);
        bad |= or8(p) & ~S.ld(kActBase + n);
    }
    return bad;
}

// Everything up to the magnitudes' inputs: Gamma into G, P in place of the syndromes (= omega below the declared count);
// returns the mask of codewords that do not qualify.  SP: on entry the 32 syndromes as planes (word 8 m + t = plane t
// of S_(m+1)), on return P_0 .. P_31.
template <int STRIDE, int SO>
BS_FN uint32_t erasure_lane(const Mem<SO> SP, const Mem<SO> EL, const Mem<SO> G, const Mem<STRIDE> S,
                            const uint8_t *inv_tab, int nslots = kES, int ns = kES)
{
    // ns = roots of the code = syndromes per codeword = the most slots a row may declare (32; 16 / 8 for RS(255,239) /
    // RS(255,247)): P_0 .. P_(ns-1) are formed and tested
    erasure_locator<STRIDE, SO>(EL, G, S, inv_tab, nslots);
    erasure_product<SO>(SP, G, nslots, ns);
    return erasure_residue<STRIDE, SO>(SP, S, ns);
}

// Magnitudes of slots 4 w .. 4 w + 3: afterwards words 8 y .. 8 y + 7 of S hold the magnitude planes of slot 4 w + y
// (magnitude_words() turns them into one word per codeword), cnt (6 bit-planes) has counted the non-zero ones.
// root_tab[list byte] = alpha^(location + 1), the root of Gamma that belongs to the slot.
template <int STRIDE, int SO>
// teff: an even number >= 4 and >= the degree of every locator of the lane (its coefficients and those of omega are
// zero from there on): the Horner chains start at teff - 1.
BS_FN void erasure_forney(const Mem<SO> EL, const Mem<SO> G, const Mem<SO> P, const Mem<STRIDE> S, int w,
                          const uint8_t *root_tab, uint32_t (&cnt)[6], int teff = kES)
{
    stage_slots<STRIDE, SO>(EL, S, w, root_tab);
    forney_math<STRIDE, SO, 6>(G, P, S, cnt, teff);
}

// ---- errors on top of the erasures -------------------------------------------------------------------------------
// A row with regular declared positions AND errors elsewhere: the reference's Berlekamp-Massey, started from Gamma
// with el = count (src/decode.c:51-96), is the plain algorithm on the sequence T_m = P_(count+m) (the "Forney
// syndromes": discrepancy, length rule 2 el <= r + count - 1 and updates are the plain ones with r - count and
// el - count), its locator is Gamma * sigma with sigma the plain locator of T.  So for a lane whose codewords all
// declare F positions: sigma by the plain ribm_mem on the word blocks F .. 31 of P (32 - F terms), its roots by the
// plain Chien transform, then Lambda = Gamma * sigma and Omega = P * sigma by the same in-place triangular product,
// and Forney on the merged, location-ordered list of declared positions and found roots - whose i-th magnitude the
// reference applies at the caller's i-th SLOT (src/decode.c:211-214), which the caller of this code must do too.
//
// ee_lane: everything up to the roots of sigma.  SP: the syndromes as planes, with room for 64 word blocks (the
// shifted sequence is read up to one block past its end, as the plain solver does); G: 33 blocks; SG: 17 blocks,
// receives sigma; MK: 256 words; W: the work area (root lists of sigma as after solve_lane).  Afterwards
// classify(W, out, c, nroots, 32 - F) gives the outcome of sigma for codeword c.
template <int STRIDE, int SO, class Hook>
BS_FN void ee_lane(const Mem<SO> SP, const Mem<SO> EL, const Mem<SO> G, const Mem<SO> SG, const Mem<SO> MK,
                   const Mem<STRIDE> W, const uint8_t *inv_tab, int F, int pad, LaneOut &out, Hook &hook,
                   uint32_t *bm_area = nullptr)
{
    erasure_locator<STRIDE, SO>(EL, G, W, inv_tab, F);
    erasure_product<SO>(SP, G, F, kES);
    const Mem<SO> T{SP.base + 8 * F * SO};
    // B of Berlekamp-Massey: 128 words, by default the list area of W (filled only afterwards)
    const Mem<STRIDE> B{bm_area ? bm_area : W.base + kListBase * STRIDE};
    ribm_mem<STRIDE, SO>(T, B, SG, out.K, kES - F);
#pragma unroll
    for (int j = 0; j < kNL; j++) {
        uint32_t l[8];
        ld8(SG, j, l);
        out.nzL[j] = or8(l);
    }
    chien_transform<STRIDE, SO, Hook>(SG, MK, W, pad, out, hook);
}

// Lambda = Gamma * sigma (in place of Gamma, 33 blocks) and Omega = P * sigma (in place of P); lmax: the largest degree
// of sigma among the codewords that will be used
template <int SO>
BS_FN void ee_products(const Mem<SO> SP, const Mem<SO> G, const Mem<SO> SG, int lmax)
{
    erasure_product<SO>(G, SG, lmax, kES + 1);
    erasure_product<SO>(SP, SG, lmax, kES);
}

} // namespace bs
