/* handle.c — opaque config / handle objects of the public API (poporon_*_config_create,
 * poporon_create, poporon_destroy, getters, version).  Mirrors the observable behaviour of
 * reference src/poporon.c:57-110,148-363; a handle additionally owns one device "lane"
 * (libpoporon_cuda context + staging buffers) per GPU it may use. */
#include "pp_internal.h"

poporon_config_t *poporon_rs_config_create(uint8_t symbol_size, uint16_t generator_polynomial,
                                           uint16_t first_consecutive_root, uint16_t primitive_element,
                                           uint8_t num_roots, poporon_erasure_t *erasure, uint16_t *syndrome)
{
    poporon_config_t *c = (poporon_config_t *)calloc(1, sizeof *c);
    if (!c)
        return NULL;
    c->fec_type = PPLN_FEC_RS;
    c->symbol_size = symbol_size;
    c->gfpoly = generator_polynomial;
    c->fcr = first_consecutive_root;
    c->prim = primitive_element;
    c->nroots = num_roots;
    c->erasure = erasure;
    c->syndrome = syndrome;
    return c;
}

poporon_config_t *poporon_bch_config_create(uint8_t symbol_size, uint16_t generator_polynomial,
                                            uint8_t correction_capability)
{
    poporon_config_t *c = (poporon_config_t *)calloc(1, sizeof *c);
    if (!c)
        return NULL;
    c->fec_type = PPLN_FEC_BCH;
    c->symbol_size = symbol_size;
    c->gfpoly = generator_polynomial;
    c->t = correction_capability;
    return c;
}

poporon_config_t *poporon_config_rs_default(void)
{
    return poporon_rs_config_create(8, 0x11D, 1, 1, 32, NULL, NULL);
}

poporon_config_t *poporon_config_bch_default(void)
{
    return poporon_bch_config_create(4, 0x13, 3);
}

void poporon_config_destroy(poporon_config_t *config)
{
    free(config);
}

static poporon_t *handle_alloc(poporon_fec_type_t type)
{
    poporon_t *h = (poporon_t *)calloc(1, sizeof *h);
    if (!h)
        return NULL;
    h->fec_type = type;
    h->nlanes = 1;
    h->lane[0].device = 0;
    h->chunk_codewords = (size_t)1 << 18; /* 256 Ki codewords (~64 MiB of RS(255,223)) per staging slot */
    return h;
}

static poporon_t *create_rs(const poporon_config_t *cfg)
{
    poporon_t *h;
    poporon_rs_t *rs = poporon_rs_create(cfg->symbol_size, cfg->gfpoly, cfg->fcr, cfg->prim, cfg->nroots);
    if (!rs)
        return NULL;
    /* prim == 0 or no inverse of prim modulo nn: the reference's constructor fails (src/poporon.c:78-93).
     * nroots == 0 or nroots >= nn leaves no message symbols and is undefined there; rejected here. */
    if (cfg->prim == 0 || rs->iprim == 0 || cfg->nroots == 0 || cfg->nroots >= rs->gf->nn) {
        poporon_rs_destroy(rs);
        return NULL;
    }
    h = handle_alloc(PPLN_FEC_RS);
    if (!h) {
        poporon_rs_destroy(rs);
        return NULL;
    }
    h->rs = rs;
    h->erasure = cfg->erasure;
    h->ext_syndrome = cfg->syndrome;
    h->parity_bytes = cfg->nroots;
    if (!pp_lane_prepare(h, &h->lane[0])) { /* no usable GPU: there is no CPU fallback */
        poporon_destroy(h);
        return NULL;
    }
    return h;
}

poporon_t *pp_handle_from_bch(poporon_bch_t *bch)
{
    poporon_t *h = handle_alloc(PPLN_FEC_BCH);
    if (!h)
        return NULL;
    h->bch = bch;
    h->parity_bytes = (size_t)((bch->parity_bits + 7u) / 8u);
    if (bch->n <= 31 && !pp_lane_prepare(h, &h->lane[0])) {
        h->bch = NULL;
        poporon_destroy(h);
        return NULL;
    }
    return h;
}

static poporon_t *create_bch(const poporon_config_t *cfg)
{
    poporon_t *h;
    poporon_bch_t *bch = poporon_bch_create(cfg->symbol_size, cfg->gfpoly, cfg->t);
    if (!bch)
        return NULL;
    h = pp_handle_from_bch(bch);
    if (!h)
        poporon_bch_destroy(bch);
    return h;
}

poporon_t *poporon_create(const poporon_config_t *config)
{
    if (!config)
        return NULL;
    switch (config->fec_type) {
    case PPLN_FEC_RS:
        return create_rs(config);
    case PPLN_FEC_BCH:
        return create_bch(config);
    default:
        return NULL;
    }
}

void poporon_destroy(poporon_t *h)
{
    if (!h)
        return;
    pp_handle_release_lanes(h);
    poporon_rs_destroy(h->rs);
    if (h->bch && h->bch->self != h) /* not the helper handle of a standalone poporon_bch_t */
        poporon_bch_destroy(h->bch);
    free(h);
}

poporon_fec_type_t poporon_get_fec_type(const poporon_t *h)
{
    return h ? h->fec_type : PPLN_FEC_UNKNOWN;
}

uint32_t poporon_get_iterations_used(const poporon_t *h)
{
    (void)h;
    return 0; /* only LDPC iterates (reference src/poporon.c:315-322) */
}

size_t poporon_get_parity_size(const poporon_t *h)
{
    if (!h)
        return 0;
    if (h->fec_type == PPLN_FEC_RS)
        return h->rs->nroots;
    if (h->fec_type == PPLN_FEC_BCH && h->bch)
        return (size_t)((uint16_t)(h->bch->n - h->bch->k) + 7u) / 8u;
    return 0;
}

size_t poporon_get_info_size(const poporon_t *h)
{
    if (!h)
        return 0;
    if (h->fec_type == PPLN_FEC_RS)
        return (size_t)(h->rs->gf->nn - h->rs->nroots);
    if (h->fec_type == PPLN_FEC_BCH && h->bch)
        return (size_t)(h->bch->k + 7u) / 8u;
    return 0;
}

uint32_t poporon_version_id(void)
{
    return PP_VERSION_ID;
}

poporon_buildtime_t poporon_buildtime(void)
{
    return (poporon_buildtime_t)POPORON_BUILDTIME;
}
