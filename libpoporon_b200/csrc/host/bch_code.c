/* bch_code.c — binary BCH code description built on the host (poporon_bch_create) and the
 * word-level convenience API the reference exposes through src/internal/common.h:116-124.
 *
 * Construction follows reference src/bch.c:168-322: the generator is the product of the distinct
 * minimal polynomials of alpha^1 .. alpha^2t, accumulated in a 32-bit word (which silently
 * truncates for long codes — kept, because the getters expose the result).  Codeword arithmetic
 * itself runs on the device (csrc/cuda/bch.cu); the word-level calls below pack one word into the
 * byte form of poporon_encode / poporon_decode and launch a one-codeword batch. */
#include "pp_internal.h"

static uint32_t min_poly_bits(const poporon_gf_t *gf, int e)
{
    uint16_t c[64];
    int deg = 0, conj = e, i, j;
    uint32_t bits = 0;

    memset(c, 0, sizeof c);
    c[0] = 1;
    do { /* c(x) *= (x + alpha^conj) */
        const uint16_t r = gf->alpha_to[conj];
        for (j = deg; j >= 0; j--) {
            if (j + 1 < 64)
                c[j + 1] ^= c[j];
            c[j] = (c[j] && r) ? gf->alpha_to[(gf->index_of[c[j]] + gf->index_of[r]) % gf->nn] : 0;
        }
        deg++;
        conj = (conj * 2) % gf->nn;
    } while (conj != e);
    for (i = 0; i <= deg; i++)
        if (c[i] == 1)
            bits |= 1u << i;
    return bits;
}

static int top_bit(uint32_t p)
{
    int d = -1;
    while (p) {
        d++;
        p >>= 1;
    }
    return d;
}

poporon_bch_t *poporon_bch_create(uint8_t symbol_size, uint16_t generator_polynomial, uint8_t t)
{
    poporon_bch_t *b;
    uint8_t seen[256];
    uint32_t gen = 1;
    int gen_deg = 0, i;

    if (symbol_size < 3 || symbol_size > 16 || t < 1 || t > 16)
        return NULL;
    b = (poporon_bch_t *)calloc(1, sizeof *b);
    if (!b)
        return NULL;
    b->gf = poporon_gf_create(symbol_size, generator_polynomial);
    if (!b->gf) {
        free(b);
        return NULL;
    }
    b->t = t;
    b->n = (uint16_t)((1u << symbol_size) - 1u);

    memset(seen, 0, sizeof seen);
    for (i = 1; i <= 2 * t; i++) {
        const int e = i % b->gf->nn;
        uint32_t mp, acc = 0;
        int c = e, s;
        if (seen[e])
            continue;
        do {
            seen[c] = 1;
            c = (c * 2) % b->gf->nn;
        } while (c != e);
        mp = min_poly_bits(b->gf, e);
        for (s = 0; s <= gen_deg; s++)
            if (gen & (1u << s))
                acc ^= mp << s;
        gen = acc;
        gen_deg = top_bit(gen);
    }
    b->gen = gen;
    b->gen_deg = (uint16_t)gen_deg;
    b->parity_bits = (uint16_t)gen_deg;
    b->k = (uint16_t)(b->n - b->parity_bits);
    return b;
}

void poporon_bch_destroy(poporon_bch_t *b)
{
    if (!b)
        return;
    if (b->self) {
        b->self->bch = NULL; /* borrowed by the helper handle: do not free twice */
        poporon_destroy(b->self);
    }
    poporon_gf_destroy(b->gf);
    free(b);
}

uint16_t poporon_bch_get_codeword_length(const poporon_bch_t *b)
{
    return b ? b->n : 0;
}

uint16_t poporon_bch_get_data_length(const poporon_bch_t *b)
{
    return b ? b->k : 0;
}

uint8_t poporon_bch_get_correction_capability(const poporon_bch_t *b)
{
    return b ? b->t : 0;
}

uint32_t poporon_bch_extract_data(const poporon_bch_t *b, uint32_t codeword)
{
    if (!b || b->k >= 32 || b->parity_bits >= 32)
        return 0;
    return (codeword >> b->parity_bits) & ((1u << b->k) - 1u);
}

static poporon_t *word_handle(poporon_bch_t *b)
{
    if (!b->self)
        b->self = pp_handle_from_bch(b);
    return b->self;
}

static void word_to_bytes(uint32_t v, uint8_t *out, unsigned nbytes)
{
    unsigned i;
    for (i = 0; i < nbytes; i++)
        out[nbytes - 1 - i] = (uint8_t)(v >> (8 * i));
}

bool poporon_bch_encode(poporon_bch_t *b, uint32_t data, uint32_t *codeword)
{
    uint8_t d[4] = {0, 0, 0, 0}, p[4] = {0, 0, 0, 0};
    unsigned db, pb, i;
    uint32_t pv = 0;
    poporon_t *h;

    if (!b || !codeword || b->n > 31)
        return false;
    if (data >= (1u << b->k)) /* reference src/bch.c:362-364 */
        return false;
    h = word_handle(b);
    if (!h)
        return false;
    db = (b->k + 7u) / 8u;
    pb = (b->parity_bits + 7u) / 8u;
    word_to_bytes(data, d, db);
    if (!poporon_encode(h, d, db, p))
        return false;
    for (i = 0; i < pb; i++)
        pv = (pv << 8) | p[i];
    *codeword = (data << b->parity_bits) | pv;
    return true;
}

bool poporon_bch_decode(poporon_bch_t *b, uint32_t received, uint32_t *corrected, int32_t *num_errors)
{
    uint8_t d[4] = {0, 0, 0, 0}, p[4] = {0, 0, 0, 0};
    unsigned db, pb;
    size_t nc = 0;
    uint32_t fixed_data, cw;
    poporon_t *h;

    if (!b || !corrected || b->n > 31)
        return false;
    received &= (1u << b->n) - 1u;
    *corrected = received;
    if (num_errors)
        *num_errors = 0;
    h = word_handle(b);
    if (!h)
        return false;
    db = (b->k + 7u) / 8u;
    pb = (b->parity_bits + 7u) / 8u;
    word_to_bytes(poporon_bch_extract_data(b, received), d, db);
    word_to_bytes(received & ((1u << b->parity_bits) - 1u), p, pb);
    if (!poporon_decode(h, d, db, p, &nc))
        return false;
    /* The byte-level decode returns the corrected message bits only; a successful decode means the
     * corrected word is a codeword (the re-syndrome check at reference src/bch.c:423-426 passed), so
     * its parity bits are the encoding of those message bits. */
    fixed_data = 0;
    for (unsigned i = 0; i < db; i++)
        fixed_data = (fixed_data << 8) | d[i];
    if (nc == 0) {
        cw = received;
    } else if (!poporon_bch_encode(b, fixed_data, &cw)) {
        return false;
    }
    *corrected = cw;
    if (num_errors)
        *num_errors = (int32_t)nc;
    return true;
}
