/* eraslist.c — the caller-owned erasure position list (public poporon_erasure_* API).
 * Capacity and growth rules follow reference src/erasure.c:27-127; storage is zero-initialised so
 * slots that were never written read as position 0. */
#include "pp_internal.h"

poporon_erasure_t *poporon_erasure_create(uint16_t num_roots, uint32_t initial_capacity)
{
    const uint32_t cap = initial_capacity ? initial_capacity : num_roots;
    poporon_erasure_t *e = (poporon_erasure_t *)calloc(1, sizeof *e);
    if (!e)
        return NULL;
    e->positions = (uint32_t *)calloc(cap ? cap : 1, sizeof(uint32_t));
    e->corrections = (uint16_t *)calloc(cap ? cap : 1, sizeof(uint16_t));
    if (!e->positions || !e->corrections) {
        poporon_erasure_destroy(e);
        return NULL;
    }
    e->capacity = cap;
    return e;
}

poporon_erasure_t *poporon_erasure_create_from_positions(uint16_t num_roots, const uint32_t *erasure_positions,
                                                         uint32_t erasure_count)
{
    poporon_erasure_t *e;
    if (!erasure_positions || erasure_count == 0)
        return NULL;
    e = poporon_erasure_create(num_roots, erasure_count > num_roots ? erasure_count : num_roots);
    if (!e)
        return NULL;
    memcpy(e->positions, erasure_positions, (size_t)erasure_count * sizeof(uint32_t));
    e->count = erasure_count;
    return e;
}

bool poporon_erasure_add_position(poporon_erasure_t *e, uint32_t position)
{
    if (!e)
        return false;
    if (e->count >= e->capacity) {
        uint32_t ncap = e->capacity * 2u;
        uint32_t *np;
        uint16_t *nc;
        if (ncap < e->capacity + 32u)
            ncap = e->capacity + 32u;
        np = (uint32_t *)calloc(ncap, sizeof(uint32_t));
        nc = (uint16_t *)calloc(ncap, sizeof(uint16_t));
        if (!np || !nc) {
            free(np);
            free(nc);
            return false;
        }
        memcpy(np, e->positions, (size_t)e->capacity * sizeof(uint32_t));
        memcpy(nc, e->corrections, (size_t)e->capacity * sizeof(uint16_t));
        free(e->positions);
        free(e->corrections);
        e->positions = np;
        e->corrections = nc;
        e->capacity = ncap;
    }
    e->positions[e->count++] = position;
    return true;
}

void poporon_erasure_reset(poporon_erasure_t *e)
{
    if (e)
        e->count = 0;
}

void poporon_erasure_destroy(poporon_erasure_t *e)
{
    if (!e)
        return;
    free(e->positions);
    free(e->corrections);
    free(e);
}
