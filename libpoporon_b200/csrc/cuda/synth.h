/* synth.h - workload generators of bench.py and the tests (libpoporon_synth.so; NOT part of the product libraries). */
#ifndef POPORON_SYNTH_H
#define POPORON_SYNTH_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* ---- synthetic workload generation (bench / test harness only; not part of the codec) ------- */
/* data[i][j] = synth_byte(seed, i0 + i, j) for j < size.  Same function in oracle-free Python (tests). */
int poporon_cuda_synth_fill(int device, uint8_t *data, size_t stride, size_t size, size_t n, uint64_t seed,
                            uint64_t i0, void *stream);
/* XOR deterministic error patterns into packed codewords (cw = data||parity, `len` symbols each):
 * codeword i gets e_i distinct positions with non-zero magnitudes, where
 *   kind 0: e_i uniform in [emin, emax];   kind 1: each symbol hit independently with prob p_ppm / 1e6.
 * Re-applying the same call removes the errors again (XOR). */
int poporon_cuda_synth_inject(int device, uint8_t *cw, size_t stride, size_t len, size_t n, uint64_t seed,
                              uint64_t i0, int kind, uint32_t emin, uint32_t emax, uint32_t p_ppm, void *stream);

/* Erasure workload on packed codewords: c_i uniform in [0, cmax] erased
 * positions in [0, size), written ASCENDING into row i of pos ([n][pos_stride], unused slots 0) with
 * cnt[i] = c_i; erased symbols get a random magnitude (zero allowed); on top, v in {0,1,2,3}
 * unknown errors (p = .90/.06/.03/.01) anywhere in the `len` symbols. */
int poporon_cuda_synth_erasures(int device, uint8_t *cw, size_t stride, size_t size, size_t len, size_t n,
                                uint64_t seed, uint64_t i0, uint32_t cmax, uint32_t *pos, size_t pos_stride,
                                uint32_t *cnt, void *stream);

#ifdef __cplusplus
}
#endif
#endif
