// host_solver.cpp — runs the bitsliced RS(255,223) solver core (libpoporon_b200/csrc/cuda/rs255_solve_core.h)
// on the host, one lane (32 codewords) at a time, so that its arithmetic can be checked against the oracle
// without a GPU.  Test infrastructure: built on the fly by tests/test_solver_core.py.
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#include "rs255_solve_core.h"
#include "rs255_erasure_core.h"

static int host_solve_ns(const uint8_t *syn, size_t n, int pad, int ns, int8_t *status, uint8_t *nroots, uint16_t *lists);

extern "C" int host_solve(const uint8_t *syn /* [n][32]: S_1..S_32 as field elements */, size_t n, int pad,
                          int8_t *status /* [n] 0 declined, 1 fail, 2 ok */, uint8_t *nroots /* [n] */,
                          uint16_t *lists /* [n][16]: location | magnitude << 8 */)
{
    return host_solve_ns(syn, n, pad, 32, status, nroots, lists);
}

// the same for a code with ns = 16 or 8 roots (RS(255,239), RS(255,247)): only S_1..S_ns of a row are meaningful
extern "C" int host_solve_roots(const uint8_t *syn, size_t n, int pad, int ns, int8_t *status, uint8_t *nroots,
                                uint16_t *lists)
{
    return host_solve_ns(syn, n, pad, ns, status, nroots, lists);
}

static int host_solve_ns(const uint8_t *syn, size_t n, int pad, int ns, int8_t *status, uint8_t *nroots, uint16_t *lists)
{
    uint8_t root_tab[256];
    root_tab[0] = 0;
    for (int i = 1; i < 256; i++)
        root_tab[i] = gf256::pow_alpha(i);
    for (size_t g = 0; g < n; g += 32) {
        uint32_t Syn[bs::kWordsS], B[bs::kWordsO], S[bs::kWordsW], Mk[bs::kWordsM];
        memset(Syn, 0, sizeof Syn);
        memset(B, 0, sizeof B);
        memset(S, 0, sizeof S);
        const size_t cnt = n - g < 32 ? n - g : 32;
        for (size_t c = 0; c < cnt; c++)
            for (int m = 0; m < 32; m++)
                for (int t = 0; t < 8; t++)
                    if ((syn[(g + c) * 32 + m] >> t) & 1)
                        Syn[8 * m + t] |= 1u << c;
        bs::LaneOut out;
        const bs::Mem<1> Sm{S}, Bm{B}, Synm{Syn}, Mkm{Mk};
        bs::NoHook hook;
        bs::solve_lane<1, 1>(Synm, Bm, Mkm, Sm, pad, out, hook, ns);
        for (size_t c = 0; c < cnt; c++) {
            int nr = 0;
            status[g + c] = (int8_t)bs::classify<1>(Sm, out, (int)c, nr, ns);
            nroots[g + c] = (uint8_t)nr;
        }
        uint32_t nzc[5] = {0, 0, 0, 0, 0};
        for (int w = 0; w < 4; w++) {
            bs::forney_slots<1, 1>(Synm, Bm, Sm, w, root_tab, nzc, ns >> 1);
            uint32_t X[32];
            bs::magnitude_words<1>(Sm, X);
            for (size_t c = 0; c < cnt; c++) {
                const uint32_t mags = X[c], xs = S[bs::kListBase + 4 * c + w];
                for (int y = 0; y < 4; y++) {
                    const uint32_t v = (xs >> (8 * y)) & 0xFFu, mag = (mags >> (8 * y)) & 0xFFu;
                    const uint32_t loc = v ? v - 1u : 0u; // list entries are location + 1
                    lists[(g + c) * 16 + 4 * w + y] = (uint16_t)(loc | (mag << 8));
                }
            }
        }
        (void)nzc;
    }
    return 0;
}

// Erasure-only form (rs255_erasure_core.h): lists[n][32] = location + 1 of every declared slot (0 = none), ascending.
// bad[i] = 1: the row does not qualify (errors outside the declared slots); else mags[i][slot] = the magnitudes and
// count[i] = the number of non-zero ones.
static int host_erasure_ns(const uint8_t *syn, const uint8_t *lists, size_t n, int ns, int8_t *bad, uint8_t *mags, uint8_t *count);
extern "C" int host_erasure(const uint8_t *syn /* [n][32] */, const uint8_t *lists /* [n][32] */, size_t n, int8_t *bad,
                            uint8_t *mags /* [n][32] */, uint8_t *count /* [n] */)
{
    return host_erasure_ns(syn, lists, n, 32, bad, mags, count);
}
// a code with ns = 16 or 8 roots: only S_1 .. S_ns of a row and the first ns slots of a list are meaningful
extern "C" int host_erasure_roots(const uint8_t *syn, const uint8_t *lists, size_t n, int ns, int8_t *bad, uint8_t *mags,
                                  uint8_t *count)
{
    return host_erasure_ns(syn, lists, n, ns, bad, mags, count);
}
static int host_erasure_ns(const uint8_t *syn, const uint8_t *lists, size_t n, int ns, int8_t *bad, uint8_t *mags, uint8_t *count)
{
    uint8_t root_tab[256], inv_tab[256];
    root_tab[0] = inv_tab[0] = 0;
    for (int v = 1; v < 256; v++) {
        root_tab[v] = gf256::pow_alpha(v);
        inv_tab[v] = gf256::pow_alpha((255 - v) % 255);
    }
    for (size_t g = 0; g < n; g += 32) {
        static uint32_t Syn[bs::kWordsS], EL[bs::kWordsEL], G[bs::kWordsG], S[bs::kWordsW];
        memset(Syn, 0, sizeof Syn);
        memset(EL, 0, sizeof EL);
        memset(S, 0, sizeof S);
        const size_t cnt = n - g < 32 ? n - g : 32;
        for (size_t c = 0; c < cnt; c++) {
            for (int m = 0; m < 32; m++)
                for (int t = 0; t < 8; t++)
                    if ((syn[(g + c) * 32 + m] >> t) & 1)
                        Syn[8 * m + t] |= 1u << c;
            for (int e = 0; e < 32; e++)
                EL[8 * c + (e >> 2)] |= (uint32_t)lists[(g + c) * 32 + e] << (8 * (e & 3));
        }
        const bs::Mem<1> Synm{Syn}, ELm{EL}, Gm{G}, Sm{S};
        int nslots = 0; // the largest declared count of the lane, as the kernel passes it
        for (size_t c = 0; c < cnt; c++)
            for (int e = 0; e < 32; e++)
                if (lists[(g + c) * 32 + e] && e + 1 > nslots)
                    nslots = e + 1;
        const uint32_t badmask = bs::erasure_lane<1, 1>(Synm, ELm, Gm, Sm, inv_tab, nslots, ns);
        uint32_t nz[6] = {0, 0, 0, 0, 0, 0};
        memset(mags + g * 32, 0, cnt * 32);
        const int teff = nslots <= 4 ? 4 : ((nslots + 1) & ~1) < ns ? ((nslots + 1) & ~1) : ns; // as the kernel bounds it
        for (int w = 0; 4 * w < ns; w++) {
            bs::erasure_forney<1, 1>(ELm, Gm, Synm, Sm, w, root_tab, nz, teff);
            uint32_t X[32];
            bs::magnitude_words<1>(Sm, X);
            for (size_t c = 0; c < cnt; c++)
                for (int y = 0; y < 4; y++)
                    mags[(g + c) * 32 + 4 * w + y] = (uint8_t)(X[c] >> (8 * y));
        }
        for (size_t c = 0; c < cnt; c++) {
            bad[g + c] = (int8_t)((badmask >> c) & 1u);
            unsigned k = 0;
            for (int b = 0; b < 6; b++)
                k |= ((nz[b] >> c) & 1u) << b;
            count[g + c] = (uint8_t)k;
        }
    }
    return 0;
}

// Errors on top of the erasures (rs255_erasure_core.h, ee_lane): every group of 32 rows declares the same number F of
// positions (fcount[group]).  lists[n][32]: location + 1 of the declared slots, ascending.
//   status[i]: 0 declined, 1 fail (the reference returns false without touching anything), 2 ok
//   merged[i][32]: location + 1 of declared positions and found roots in ascending order of location (rank order)
//   mags[i][32]: the magnitude of every rank;  count[i]: the non-zero ones
extern "C" int host_ee(const uint8_t *syn /* [n][32] */, const uint8_t *lists /* [n][32] */, size_t n, const int *fcount,
                       int pad, int8_t *status, uint8_t *merged, uint8_t *mags, uint8_t *count)
{
    uint8_t root_tab[256], inv_tab[256];
    root_tab[0] = inv_tab[0] = 0;
    for (int v = 1; v < 256; v++) {
        root_tab[v] = gf256::pow_alpha(v);
        inv_tab[v] = gf256::pow_alpha((255 - v) % 255);
    }
    for (size_t g = 0; g < n; g += 32) {
        static uint32_t SP[8 * 64], EL[bs::kWordsEL], G[bs::kWordsG], SG[8 * 17], MK[bs::kWordsM], W[bs::kWordsW];
        memset(SP, 0, sizeof SP);
        memset(EL, 0, sizeof EL);
        memset(W, 0, sizeof W);
        const size_t cnt = n - g < 32 ? n - g : 32;
        const int F = fcount[g / 32];
        for (size_t c = 0; c < cnt; c++) {
            for (int m = 0; m < 32; m++)
                for (int t = 0; t < 8; t++)
                    if ((syn[(g + c) * 32 + m] >> t) & 1)
                        SP[8 * m + t] |= 1u << c;
            for (int e = 0; e < 32; e++)
                EL[8 * c + (e >> 2)] |= (uint32_t)lists[(g + c) * 32 + e] << (8 * (e & 3));
        }
        const bs::Mem<1> SPm{SP}, ELm{EL}, Gm{G}, SGm{SG}, MKm{MK}, Wm{W};
        bs::LaneOut out;
        bs::NoHook hook;
        bs::ee_lane<1, 1>(SPm, ELm, Gm, SGm, MKm, Wm, inv_tab, F, pad, out, hook);
        int lmax = 0;
        memset(EL, 0, sizeof EL);
        for (size_t c = 0; c < cnt; c++) {
            int nroots = 0;
            int st = bs::classify<1>(Wm, out, (int)c, nroots, 32 - F);
            uint8_t m[48];
            int k = 0;
            if (st == bs::kOk) {
                // merge (both ascending); a found root on a declared position is a double root of the locator: the
                // reference's search then finds fewer roots than the degree and fails
                int a = 0, b = 0;
                while (a < F || b < nroots) {
                    const int va = a < F ? lists[(g + c) * 32 + a] : 256;
                    const int vb = b < nroots ? (int)((W[bs::kListBase + 4 * c + (b >> 2)] >> (8 * (b & 3))) & 0xFF) : 256;
                    if (va == vb) {
                        st = bs::kFail;
                        break;
                    }
                    if (va < vb) {
                        m[k++] = (uint8_t)va;
                        a++;
                    } else {
                        m[k++] = (uint8_t)vb;
                        b++;
                    }
                }
            }
            status[g + c] = (int8_t)st;
            if (st == bs::kOk) {
                if (nroots > lmax)
                    lmax = nroots;
                for (int e = 0; e < k; e++) {
                    merged[(g + c) * 32 + e] = m[e];
                    EL[8 * c + (e >> 2)] |= (uint32_t)m[e] << (8 * (e & 3));
                }
            }
        }
        bs::ee_products<1>(SPm, Gm, SGm, lmax);
        uint32_t nz[6] = {0, 0, 0, 0, 0, 0};
        const int dmax = F + lmax, teff = dmax <= 4 ? 4 : (dmax + 1) & ~1; // as the kernel bounds it
        for (int w = 0; w < 8; w++) {
            bs::erasure_forney<1, 1>(ELm, Gm, SPm, Wm, w, root_tab, nz, teff);
            uint32_t X[32];
            bs::magnitude_words<1>(Wm, X);
            for (size_t c = 0; c < cnt; c++)
                for (int y = 0; y < 4; y++)
                    mags[(g + c) * 32 + 4 * w + y] = (uint8_t)(X[c] >> (8 * y));
        }
        for (size_t c = 0; c < cnt; c++) {
            unsigned k = 0;
            for (int b = 0; b < 6; b++)
                k |= ((nz[b] >> c) & 1u) << b;
            count[g + c] = (uint8_t)k;
        }
    }
    return 0;
}
