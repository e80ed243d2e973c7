/*
 * poporon/erasure.h — growable list of known-bad symbol positions for RS erasure decoding.
 * Source-compatible with the reference's include/poporon/erasure.h:21-26 (behaviour:
 * src/erasure.c:12-127).  Positions are data-relative symbol indices.  The list is attached to a
 * config with poporon_rs_config_create() and is borrowed, not copied: the caller refills it
 * (reset + add_position) between poporon_decode() calls.
 *
 * Unlike the reference, unused slots are zero-initialised, which makes the reference's
 * "slot order" application deterministic.
 */
#ifndef POPORON_ERASURE_H
#define POPORON_ERASURE_H

#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct _poporon_erasure_t poporon_erasure_t;

/* capacity = initial_capacity, or num_roots when that is 0 */
poporon_erasure_t *poporon_erasure_create(uint16_t num_roots, uint32_t initial_capacity);
/* NULL when erasure_positions is NULL or erasure_count is 0; capacity = max(count, num_roots) */
poporon_erasure_t *poporon_erasure_create_from_positions(uint16_t num_roots, const uint32_t *erasure_positions,
                                                         uint32_t erasure_count);
bool poporon_erasure_add_position(poporon_erasure_t *erasure, uint32_t position); /* grows as needed */
void poporon_erasure_reset(poporon_erasure_t *erasure);                           /* count = 0 */
void poporon_erasure_destroy(poporon_erasure_t *erasure);                         /* NULL is a no-op */

#ifdef __cplusplus
}
#endif

#endif /* POPORON_ERASURE_H */
