/* rs_code.c — Reed-Solomon code description built on the host (poporon_rs_create).
 * Generator polynomial prod_i (x - alpha^{(fcr+i)*prim}) in index form, as reference
 * src/rs.c:59-80 stores it, and the inverse of prim modulo nn (reference src/poporon.c:84-93).
 * 16-bit wrap-around of the running root exponent is kept because it is observable for large
 * fcr/prim. */
#include "pp_internal.h"

static uint16_t gmul_log(const poporon_gf_t *gf, uint16_t poly_value, uint16_t log_b)
{
    /* poly_value * alpha^log_b, with the exponent sum truncated to 16 bits before reduction */
    return gf->alpha_to[(uint16_t)(gf->index_of[poly_value] + log_b) % gf->nn];
}

poporon_rs_t *poporon_rs_create(uint8_t symbol_size, uint16_t generator_polynomial, uint16_t first_consecutive_root,
                                uint16_t primitive_element, uint8_t num_roots)
{
    poporon_rs_t *rs;
    poporon_gf_t *gf = poporon_gf_create(symbol_size, generator_polynomial);
    uint16_t g[257];
    uint16_t root;
    unsigned i, j;

    if (!gf)
        return NULL;
    rs = (poporon_rs_t *)calloc(1, sizeof *rs);
    if (!rs) {
        poporon_gf_destroy(gf);
        return NULL;
    }
    rs->gf = gf;
    rs->fcr = first_consecutive_root;
    rs->prim = primitive_element;
    rs->nroots = num_roots;

    memset(g, 0, sizeof g);
    g[0] = 1;
    root = (uint16_t)(first_consecutive_root * primitive_element);
    for (i = 0; i < num_roots; i++) {
        /* multiply g by (x + alpha^root) */
        g[i + 1] = 1;
        for (j = i; j > 0; j--)
            g[j] = g[j] ? (uint16_t)(g[j - 1] ^ gmul_log(gf, g[j], root)) : g[j - 1];
        g[0] = gmul_log(gf, g[0], root);
        root = (uint16_t)(root + primitive_element);
    }
    for (i = 0; i <= num_roots; i++)
        rs->genpoly[i] = gf->index_of[g[i]];

    /* iprim * prim == 1 (mod nn); left 0 when prim is 0 or shares a factor with nn (handle creation
     * rejects those) */
    rs->iprim = 0;
    if (primitive_element != 0) {
        uint16_t probe = 1;
        uint32_t tries = 0;
        while ((probe % primitive_element) != 0 && ++tries <= (uint32_t)gf->nn * 2u)
            probe = (uint16_t)(probe + gf->nn);
        if ((probe % primitive_element) == 0)
            rs->iprim = (uint16_t)(probe / primitive_element);
    }
    return rs;
}

void poporon_rs_destroy(poporon_rs_t *rs)
{
    if (!rs)
        return;
    poporon_gf_destroy(rs->gf);
    free(rs);
}
