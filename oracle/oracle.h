/*
 * oracle.h — CPU oracle for the RS/BCH hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Nothing under oracle/ is part of the product: only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load these libraries, and
 * only as the checker or the timed CPU baseline.  The product (libpoporon.so +
 * libpoporon_cuda.so) never links, loads or calls anything in this directory.
 *
 * Two libraries are built from this directory (see oracle/Makefile):
 *
 *   oracle/liboracle.so            "port": our own plain-C restatement of the reference
 *                                  algorithm (oracle_rs.c, oracle_bch.c).  Symbols orc_*.
 *   oracle/_ref/libpoporon_ref.so  "reference": the UNMODIFIED reference sources compiled
 *                                  where they lie under /root/reference/src, plus
 *                                  ref_shim.c (ours) which drives the reference's public
 *                                  API over a batch.  Symbols ref_*.  git-ignored, shipped
 *                                  to the GPU box with the snapshot.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle.py, -m "not gpu")
 * against (i) the known-answer vectors of SURVEY.md Appendix B, (ii) the committed
 * fixtures in tests/golden/ which were produced by the compiled reference
 * (tests/golden/make_golden.py), and (iii) when oracle/_ref exists, live differential
 * runs against the reference itself.
 *
 * Both libraries export the same batch-runner ABI (prefix orc_ / ref_) so tests can
 * swap one for the other.  "Batch" = the single-instance semantics of
 * poporon_encode / poporon_decode (reference src/encode.c:236-252, src/decode.c:596-612)
 * applied to n independent codewords; `threads` > 1 fans the batch out over pthreads with
 * one independent handle per thread (this is the "one poporon_t per host core" CPU
 * baseline of BASELINE.md §3).
 */
#ifndef POPORON_ORACLE_H
#define POPORON_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* RS code parameters, same meaning as poporon_rs_config_create (reference src/poporon.c:214-233). */
typedef struct {
    uint8_t symbol_size;
    uint16_t gfpoly;
    uint16_t fcr;
    uint16_t prim;
    uint8_t nroots;
} orc_rs_params_t;

/* Decode modes (reference src/decode.c:446-477). */
#define ORC_MODE_PLAIN   0 /* no erasure object, no external syndrome */
#define ORC_MODE_ERASURE 1 /* handle created with a non-NULL poporon_erasure_t */
#define ORC_MODE_EXTSYN  2 /* handle created with a non-NULL syndrome pointer */

/*
 * Every runner returns 0 on success, <0 if the codec could not be constructed
 * (constructor returned NULL in the reference).
 *
 * RS encode: parity[i] <- encode(data[i]); ok[i] (nullable) <- return value.
 * RS decode: in place on data/parity; ok[i] <- return value; corrected[i] <- the value
 *   *corrected_num holds after the call when it held 0 before.
 *   erasure mode: pos is [n][pos_stride] uint32 with EVERY slot defined, cnt[n].
 *   ext-syndrome mode: syn is [n][nroots] uint16 in index form.
 */
#define ORC_DECL(prefix)                                                                                            \
    int prefix##_rs_encode_batch(const orc_rs_params_t *p, const uint8_t *data, size_t data_stride, size_t size,    \
                                 uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok, int threads);        \
    int prefix##_rs_decode_batch(const orc_rs_params_t *p, int mode, uint8_t *data, size_t data_stride,             \
                                 size_t size, uint8_t *parity, size_t parity_stride, size_t n, const uint32_t *pos, \
                                 size_t pos_stride, const uint32_t *cnt, const uint16_t *syn, uint8_t *ok,          \
                                 uint32_t *corrected, int threads);                                                 \
    int prefix##_bch_encode_batch(uint8_t m, uint16_t poly, uint8_t t, const uint8_t *data, size_t data_stride,     \
                                  size_t size, uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok,        \
                                  int threads);                                                                     \
    int prefix##_bch_decode_batch(uint8_t m, uint16_t poly, uint8_t t, uint8_t *data, size_t data_stride,           \
                                  size_t size, const uint8_t *parity, size_t parity_stride, size_t n, uint8_t *ok,  \
                                  uint32_t *corrected, int threads);                                                \
    /* introspection: fills out[0..3] = {codeword_length n, data_length k, parity_bits, gen_poly} */               \
    int prefix##_bch_info(uint8_t m, uint16_t poly, uint8_t t, uint32_t out[4]);                                    \
    /* genpoly_index[nroots+1], alpha_to[nn+1], index_of[nn+1] (uint16 each), *iprim */                             \
    int prefix##_rs_info(const orc_rs_params_t *p, uint16_t *genpoly_index, uint16_t *alpha_to,                     \
                         uint16_t *index_of, uint16_t *iprim);                                                      \
    uint8_t prefix##_gf_mod(uint8_t m, uint16_t poly, uint16_t value);                                              \
    const char *prefix##_kind(void);

ORC_DECL(orc)
ORC_DECL(ref)

#ifdef __cplusplus
}
#endif

#endif /* POPORON_ORACLE_H */
