/*
 * poporon.h — public C99 API of poporon-b200, a B200-native batched Reed-Solomon / BCH codec.
 *
 * The declarations in the first half are source-compatible with colopl/libpoporon's
 * include/poporon.h (reference include/poporon.h:21-99): same names, argument meaning and error
 * behaviour, so code written against the reference recompiles against this header unchanged.
 * All arithmetic runs in sm_100a CUDA kernels behind libpoporon_cuda.so (include/poporon_cuda.h);
 * there is no CPU implementation of the codecs in this library and no CPU fallback: on a machine
 * without a usable CUDA device poporon_create() returns NULL.
 *
 * The second half (include/poporon/batch.h) adds the batched entry points, which apply exactly the
 * single-instance semantics to n independent codewords and accept host or device pointers.
 *
 * Scope: Reed-Solomon (PPLN_FEC_RS) and binary BCH (PPLN_FEC_BCH).  The LDPC constructors and the
 * RNG symbols exist for link compatibility only and report failure (NULL / false).
 */
#ifndef POPORON_H
#define POPORON_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "poporon/erasure.h"
#include "poporon/gf.h"
#include "poporon/rng.h"

/* FEC family tags (reference include/poporon.h:22-25) */
#define POPORON_FEC_RS      1
#define POPORON_FEC_LDPC    2
#define POPORON_FEC_BCH     3
#define POPORON_FEC_UNKNOWN 255

/* LDPC enumerators are kept so that callers compile; LDPC itself is not provided. */
#define POPORON_LDPC_RATE_1_3 0
#define POPORON_LDPC_RATE_1_2 1
#define POPORON_LDPC_RATE_2_3 2
#define POPORON_LDPC_RATE_3_4 3
#define POPORON_LDPC_RATE_4_5 4
#define POPORON_LDPC_RATE_5_6 5
#define POPORON_LDPC_MATRIX_RANDOM    1
#define POPORON_LDPC_MATRIX_QC_RANDOM 2

#ifdef __cplusplus
extern "C" {
#endif

typedef uint32_t poporon_buildtime_t;
typedef struct _poporon_t poporon_t;
typedef struct _poporon_config_t poporon_config_t;

typedef enum {
    PPLN_FEC_RS = POPORON_FEC_RS,
    PPLN_FEC_LDPC = POPORON_FEC_LDPC,
    PPLN_FEC_BCH = POPORON_FEC_BCH,
    PPLN_FEC_UNKNOWN = POPORON_FEC_UNKNOWN,
} poporon_fec_type_t;

typedef enum {
    PPRN_LDPC_RATE_1_3 = POPORON_LDPC_RATE_1_3,
    PPRN_LDPC_RATE_1_2 = POPORON_LDPC_RATE_1_2,
    PPRN_LDPC_RATE_2_3 = POPORON_LDPC_RATE_2_3,
    PPRN_LDPC_RATE_3_4 = POPORON_LDPC_RATE_3_4,
    PPRN_LDPC_RATE_4_5 = POPORON_LDPC_RATE_4_5,
    PPRN_LDPC_RATE_5_6 = POPORON_LDPC_RATE_5_6,
} poporon_ldpc_rate_t;

typedef enum {
    PPRN_LDPC_RANDOM = POPORON_LDPC_MATRIX_RANDOM,
    PPRN_LDPC_QC_RANDOM = POPORON_LDPC_MATRIX_QC_RANDOM,
} poporon_ldpc_matrix_type_t;

/* ---- configuration objects (reference src/poporon.c:214-304) -------------------------------- */

/* RS over GF(2^symbol_size) with field polynomial generator_polynomial, roots
 * alpha^{(first_consecutive_root + i) * primitive_element}, i < num_roots.  `erasure` and `syndrome`
 * are BORROWED for the lifetime of any handle created from the config and read on every decode. */
poporon_config_t *poporon_rs_config_create(uint8_t symbol_size, uint16_t generator_polynomial,
                                           uint16_t first_consecutive_root, uint16_t primitive_element,
                                           uint8_t num_roots, poporon_erasure_t *erasure, uint16_t *syndrome);
/* binary BCH of length 2^symbol_size - 1 correcting `correction_capability` bit errors */
poporon_config_t *poporon_bch_config_create(uint8_t symbol_size, uint16_t generator_polynomial,
                                            uint8_t correction_capability);
/* not provided: always NULL */
poporon_config_t *poporon_ldpc_config_create(size_t block_size, poporon_ldpc_rate_t rate,
                                             poporon_ldpc_matrix_type_t matrix_type, uint32_t column_weight,
                                             bool use_soft_decode, bool use_outer_interleave, bool use_inner_interleave,
                                             uint32_t interleave_depth, uint32_t lifting_factor,
                                             uint32_t max_iterations, const int8_t *soft_llr, size_t soft_llr_size,
                                             uint64_t seed);

poporon_config_t *poporon_config_rs_default(void);  /* RS(255,223): (8, 0x11D, 1, 1, 32) */
poporon_config_t *poporon_config_bch_default(void); /* BCH(15,5):   (4, 0x13, 3) */
poporon_config_t *poporon_config_ldpc_default(size_t block_size, poporon_ldpc_rate_t rate);         /* NULL */
poporon_config_t *poporon_config_ldpc_burst_resistant(size_t block_size, poporon_ldpc_rate_t rate); /* NULL */
void poporon_config_destroy(poporon_config_t *config);

/* ---- codec handle ---------------------------------------------------------------------------- */

/* NULL on invalid parameters, non-primitive field polynomial, allocation failure, or when no CUDA
 * device can be used.  A handle is not thread-safe; distinct handles are fully independent. */
poporon_t *poporon_create(const poporon_config_t *config);
void poporon_destroy(poporon_t *pprn);

/* One codeword.  `data` (size bytes) and `parity` (poporon_get_parity_size bytes) are host memory.
 * encode: writes parity only.  decode: corrects data and (RS) parity in place; *corrected_num, when
 * non-NULL, receives the number of corrected symbols/bits under the reference's rules
 * (SURVEY Appendix A.5).  false on NULL arguments (and, for decode, size == 0). */
bool poporon_encode(poporon_t *pprn, uint8_t *data, size_t size, uint8_t *parity);
bool poporon_decode(poporon_t *pprn, uint8_t *data, size_t size, uint8_t *parity, size_t *corrected_num);

poporon_fec_type_t poporon_get_fec_type(const poporon_t *pprn);
uint32_t poporon_get_iterations_used(const poporon_t *pprn); /* 0 for RS and BCH */
size_t poporon_get_parity_size(const poporon_t *pprn);
size_t poporon_get_info_size(const poporon_t *pprn);

uint32_t poporon_version_id(void); /* 20000000, the reference API level this header tracks */
poporon_buildtime_t poporon_buildtime(void);

#ifdef __cplusplus
}
#endif

#include "poporon/batch.h"

#endif /* POPORON_H */
