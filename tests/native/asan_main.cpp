// asan_main.cpp — drives the host harness of the bitsliced cores (host_solver.cpp) with random syndromes and slot lists;
// built with -fsanitize=address,undefined by tests/test_solver_core.py::test_cores_under_address_sanitizer.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
extern "C" int host_solve_roots(const uint8_t*, size_t, int, int, int8_t*, uint8_t*, uint16_t*);
extern "C" int host_erasure_roots(const uint8_t*, const uint8_t*, size_t, int, int8_t*, uint8_t*, uint8_t*);
extern "C" int host_ee(const uint8_t*, const uint8_t*, size_t, const int*, int, int8_t*, uint8_t*, uint8_t*, uint8_t*);
int main() {
    srand(1);
    const size_t n = 32 * 40;
    std::vector<uint8_t> syn(n * 32), lists(n * 32), mags(n * 32), merged(n * 32), cnt(n), nroots(n);
    std::vector<int8_t> st(n);
    std::vector<uint16_t> l16(n * 16);
    std::vector<int> fc(n / 32);
    for (int rep = 0; rep < 6; rep++) {
        const int pad = rep % 2 ? 0 : 123;
        for (auto &v : syn) v = rand() & 255;
        for (size_t g = 0; g < n / 32; g++) {
            fc[g] = 1 + rand() % 30;
            for (size_t i = 32 * g; i < 32 * g + 32; i++) {
                int loc = pad;
                for (int e = 0; e < 32; e++) {
                    loc += 1 + rand() % 3;
                    lists[i * 32 + e] = (e < fc[g] && loc < 223) ? (uint8_t)(loc + 1) : 0;
                }
            }
        }
        for (int ns : {32, 16, 8}) {
            host_solve_roots(syn.data(), n, pad, ns, st.data(), nroots.data(), l16.data());
            host_erasure_roots(syn.data(), lists.data(), n, ns, st.data(), mags.data(), cnt.data());
        }
        host_ee(syn.data(), lists.data(), n, fc.data(), pad, st.data(), merged.data(), mags.data(), cnt.data());
    }
    printf("asan run done\n");
    return 0;
}
