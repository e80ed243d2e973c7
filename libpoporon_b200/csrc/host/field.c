/* field.c — GF(2^m) log/antilog tables, host side (public poporon_gf_* API).
 * Behaviour follows reference src/gf.c:29-86 and gf_mod (src/internal/common.h:102-110).
 * The tables are uploaded to the device when a codec handle is created; no codeword arithmetic
 * happens here. */
#include "pp_internal.h"

poporon_gf_t *poporon_gf_create(uint8_t symbol_size, uint16_t generator_polynomial)
{
    poporon_gf_t *gf;
    unsigned x = 1, i;

    /* The reference admits m <= 16 but stores the field size in 8 bits, so m > 8 can never pass its
     * primitivity check; m <= 8 is the real domain. */
    if (symbol_size < 1 || symbol_size > 8)
        return NULL;
    gf = (poporon_gf_t *)calloc(1, sizeof *gf);
    if (!gf)
        return NULL;
    gf->symbol_size = symbol_size;
    gf->nn = (uint16_t)((1u << symbol_size) - 1u);
    gf->poly = generator_polynomial;
    gf->index_of[0] = gf->nn;
    gf->alpha_to[gf->nn] = 0;
    for (i = 0; i < gf->nn; i++) {
        gf->index_of[x] = (uint16_t)i;
        gf->alpha_to[i] = (uint16_t)x;
        x <<= 1;
        if (x >> symbol_size)
            x ^= generator_polynomial;
        x &= gf->nn;
    }
    if (x != 1u) { /* alpha did not return to 1 after nn steps: not a primitive polynomial */
        free(gf);
        return NULL;
    }
    return gf;
}

void poporon_gf_destroy(poporon_gf_t *gf)
{
    free(gf);
}

uint8_t poporon_gf_mod(poporon_gf_t *gf, uint16_t value)
{
    /* the reference folds value into [0, nn); that is value mod nn for every 16-bit input */
    return (uint8_t)(value % gf->nn);
}
