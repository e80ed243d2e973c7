/*
 * poporon/gf.h — GF(2^m) log/antilog tables (host side).  Source-compatible with the reference's
 * include/poporon/gf.h:21-24 (behaviour: src/gf.c:29-86).  The tables are computed on the host at
 * handle creation and uploaded to the device; the kernels never call back into this object.
 */
#ifndef POPORON_GF_H
#define POPORON_GF_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct _poporon_gf_t poporon_gf_t;

/* NULL unless 1 <= symbol_size <= 8 and generator_polynomial is primitive over GF(2) */
poporon_gf_t *poporon_gf_create(uint8_t symbol_size, uint16_t generator_polynomial);
void poporon_gf_destroy(poporon_gf_t *gf);
/* value mod (2^m - 1) */
uint8_t poporon_gf_mod(poporon_gf_t *gf, uint16_t value);

#ifdef __cplusplus
}
#endif

#endif /* POPORON_GF_H */
