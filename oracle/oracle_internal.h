/* oracle_internal.h — private types of the CPU restatement.  TEST INFRASTRUCTURE ONLY (see oracle.h). */
#ifndef POPORON_ORACLE_INTERNAL_H
#define POPORON_ORACLE_INTERNAL_H

#include "oracle.h"

#define ORC_MAX_ROOTS 255

typedef struct {
    uint8_t m;
    uint16_t nn; /* 2^m - 1; also the log of zero ("A0") */
    uint16_t poly;
    uint16_t alpha_to[256]; /* reference log2exp */
    uint16_t index_of[256]; /* reference exp2log */
} orc_field_t;

typedef struct {
    orc_field_t f;
    uint16_t nn, fcr, prim, iprim;
    uint8_t nroots;
    uint16_t genpoly[ORC_MAX_ROOTS + 1]; /* index form */
} orc_rs_t;

typedef struct {
    orc_field_t f;
    uint8_t t;
    uint16_t n, k, parity_bits, gen_deg;
    uint32_t gen;
} orc_bch_t;

int orc_field_init(orc_field_t *f, uint8_t m, uint16_t poly);

int orc_rs_init(orc_rs_t *rs, const orc_rs_params_t *p);
void orc_rs_encode(const orc_rs_t *rs, const uint8_t *data, size_t size, uint8_t *parity);
int orc_rs_syndromes(const orc_rs_t *rs, const uint8_t *data, size_t size, const uint8_t *parity, uint16_t *syn);
int orc_rs_correct(const orc_rs_t *rs, uint8_t *data, size_t size, uint8_t *parity, const uint16_t *syn,
                   uint32_t ecount, const uint32_t *epos, int erasure_apply, int16_t pad, size_t *ncorr);
int orc_rs_decode(const orc_rs_t *rs, int mode, uint8_t *data, size_t size, uint8_t *parity, uint32_t ecount,
                  const uint32_t *epos, const uint16_t *ext_syn, size_t *corrected);

int orc_bch_init(orc_bch_t *b, uint8_t m, uint16_t poly, uint8_t t);
int orc_bch_encode_word(const orc_bch_t *b, uint32_t data, uint32_t *codeword);
int orc_bch_decode_word(const orc_bch_t *b, uint32_t received, uint32_t *corrected, int32_t *num_errors);
int orc_bch_encode_bytes(const orc_bch_t *b, const uint8_t *data, size_t size, uint8_t *parity);
int orc_bch_decode_bytes(const orc_bch_t *b, uint8_t *data, size_t size, const uint8_t *parity, size_t *corrected);

#endif
