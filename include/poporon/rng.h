/*
 * poporon/rng.h — link-compatibility declarations for the reference's RNG (include/poporon/rng.h).
 * The RNG only serves LDPC matrix construction in the reference (src/ldpc.c) and is outside the
 * RS/BCH hot path this library implements: poporon_rng_create() returns NULL.
 */
#ifndef POPORON_RNG_H
#define POPORON_RNG_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#define POPORON_RNG_TYPE_XOSHIRO128PP 0

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { XOSHIRO128PP = POPORON_RNG_TYPE_XOSHIRO128PP } poporon_rng_type_t;
typedef struct _poporon_rng_t poporon_rng_t;

poporon_rng_t *poporon_rng_create(poporon_rng_type_t type, void *seed, size_t seed_size); /* NULL */
void poporon_rng_destroy(poporon_rng_t *rng);
bool poporon_rng_next(poporon_rng_t *rng, void *dest, size_t size); /* false */

#ifdef __cplusplus
}
#endif

#endif /* POPORON_RNG_H */
