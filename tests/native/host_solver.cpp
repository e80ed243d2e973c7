// host_solver.cpp — runs the bitsliced RS(255,223) solver core (libpoporon_b200/csrc/cuda/rs255_solve_core.h)
// on the host, one lane (32 codewords) at a time, so that its arithmetic can be checked against the oracle
// without a GPU.  Test infrastructure: built on the fly by tests/test_solver_core.py.
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#include "rs255_solve_core.h"

extern "C" int host_solve(const uint8_t *syn /* [n][32]: S_1..S_32 as field elements */, size_t n, int pad,
                          int8_t *status /* [n] 0 declined, 1 fail, 2 ok */, uint8_t *nroots /* [n] */,
                          uint16_t *lists /* [n][16]: location | magnitude << 8 */)
{
    for (size_t g = 0; g < n; g += 32) {
        uint32_t S[bs::kWordsS], B[bs::kWordsB];
        memset(S, 0, sizeof S);
        memset(B, 0, sizeof B);
        const size_t cnt = n - g < 32 ? n - g : 32;
        for (size_t c = 0; c < cnt; c++)
            for (int m = 0; m < 32; m++)
                for (int t = 0; t < 8; t++)
                    if ((syn[(g + c) * 32 + m] >> t) & 1)
                        S[8 * m + t] |= 1u << c;
        bs::LaneOut out;
        bs::solve_lane<1>(bs::Mem<1>{S}, bs::Mem<1>{B}, pad, out);
        for (size_t c = 0; c < cnt; c++) {
            int nr = 0;
            status[g + c] = (int8_t)bs::classify<1>(bs::Mem<1>{B}, out, (int)c, nr);
            nroots[g + c] = (uint8_t)nr;
            for (int e = 0; e < 16; e++)
                lists[(g + c) * 16 + e] = (uint16_t)(S[8 * c + (e >> 1)] >> (16 * (e & 1)));
        }
    }
    return 0;
}
