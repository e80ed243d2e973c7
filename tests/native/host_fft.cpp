// host_fft.cpp — runs the generated XOR networks of the RS(255,223) syndrome transform
// (libpoporon_b200/csrc/cuda/rs255_fft_gen.cuh) on the host, 32 codewords per call in bit-plane form, so that the
// arithmetic can be checked against the oracle without a GPU.  Test infrastructure: built by tests/test_fft_core.py.
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#include "rs255_fft_gen.cuh"

// cw: [32][len] bytes (first byte = highest exponent), len <= 255.  syn: [32][32] = S_1..S_32 of every codeword.
extern "C" void host_fft_syndromes(const uint8_t *cw, int len, uint8_t *syn)
{
    // planes of the symbol with exponent e, standard basis -> tower basis
    static uint32_t T[255][8];
    memset(T, 0, sizeof T);
    for (int m = 0; m < len; m++) {
        const int e = len - 1 - m;
        uint32_t s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 32; i++)
            for (int t = 0; t < 8; t++)
                if ((cw[(size_t)i * len + m] >> t) & 1)
                    s[t] |= 1u << i;
        rsfft::conv_in(s, T[e]);
    }
    uint32_t acc[33][8];
    memset(acc, 0, sizeof acc);
    for (int c = 16; c >= 0; c--) {
        uint32_t lo[15][4], hi[15][4], R[15][8];
        for (int d = 0; d < 15; d++)
            for (int t = 0; t < 4; t++) {
                lo[d][t] = T[17 * d + c][t];
                hi[d][t] = T[17 * d + c][4 + t];
            }
        rsfft::dft15(lo);
        rsfft::dft15(hi);
        for (int k = 0; k < 15; k++)
            for (int t = 0; t < 4; t++) {
                R[k][t] = lo[k][t];
                R[k][4 + t] = hi[k][t];
            }
#define HS(J) rsfft::hstep<J>(acc[J], R[(J) % 15]);
        HS(1) HS(2) HS(3) HS(4) HS(5) HS(6) HS(7) HS(8) HS(9) HS(10) HS(11) HS(12) HS(13) HS(14) HS(15) HS(16)
        HS(17) HS(18) HS(19) HS(20) HS(21) HS(22) HS(23) HS(24) HS(25) HS(26) HS(27) HS(28) HS(29) HS(30) HS(31) HS(32)
#undef HS
    }
    for (int j = 1; j <= 32; j++) {
        uint32_t s[8];
        rsfft::conv_out(acc[j], s);
        for (int i = 0; i < 32; i++) {
            uint8_t v = 0;
            for (int t = 0; t < 8; t++)
                v |= (uint8_t)(((s[t] >> i) & 1u) << t);
            syn[i * 32 + (j - 1)] = v;
        }
    }
}

// parity of 32 messages ([32][size] bytes): syndromes of message * x^32 (tower planes), then the mask table
extern "C" void host_fft_encode(const uint8_t *msg, int size, uint8_t *parity /* [32][32] */)
{
    static uint8_t cw[32 * 255];
    const int len = size + 32;
    memset(cw, 0, sizeof cw);
    for (int i = 0; i < 32; i++)
        memcpy(cw + (size_t)i * len, msg + (size_t)i * size, (size_t)size);
    // tower-basis syndromes: redo the transform without conv_out
    static uint32_t T[255][8];
    memset(T, 0, sizeof T);
    for (int m = 0; m < size; m++) {
        const int e = len - 1 - m;
        uint32_t s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 32; i++)
            for (int t = 0; t < 8; t++)
                if ((cw[(size_t)i * len + m] >> t) & 1)
                    s[t] |= 1u << i;
        rsfft::conv_in(s, T[e]);
    }
    uint32_t acc[33][8];
    memset(acc, 0, sizeof acc);
    for (int c = 16; c >= 0; c--) {
        uint32_t lo[15][4], hi[15][4], R[15][8];
        for (int d = 0; d < 15; d++)
            for (int t = 0; t < 4; t++) {
                lo[d][t] = T[17 * d + c][t];
                hi[d][t] = T[17 * d + c][4 + t];
            }
        rsfft::dft15(lo);
        rsfft::dft15(hi);
        for (int k = 0; k < 15; k++)
            for (int t = 0; t < 4; t++) {
                R[k][t] = lo[k][t];
                R[k][4 + t] = hi[k][t];
            }
#define HS(J) rsfft::hstep<J>(acc[J], R[(J) % 15]);
        HS(1) HS(2) HS(3) HS(4) HS(5) HS(6) HS(7) HS(8) HS(9) HS(10) HS(11) HS(12) HS(13) HS(14) HS(15) HS(16)
        HS(17) HS(18) HS(19) HS(20) HS(21) HS(22) HS(23) HS(24) HS(25) HS(26) HS(27) HS(28) HS(29) HS(30) HS(31) HS(32)
#undef HS
    }
    for (int q = 0; q < 32; q++)
        for (int u = 0; u < 8; u++) {
            uint32_t p = 0;
            for (int j = 0; j < 32; j++) {
                const unsigned m = rsfft::EncM::v.m[j][q][u];
                for (int t = 0; t < 8; t++)
                    if ((m >> t) & 1)
                        p ^= acc[j + 1][t];
            }
            for (int i = 0; i < 32; i++)
                if ((p >> i) & 1)
                    parity[i * 32 + q] |= (uint8_t)(1u << u);
        }
}
