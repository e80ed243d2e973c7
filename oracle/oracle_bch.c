/*
 * oracle_bch.c — plain-C restatement of the reference binary BCH path.  TEST INFRASTRUCTURE ONLY
 * (see oracle.h).  Parity status: pinned for n <= 31 (m in {3,4,5}) against the compiled reference,
 * SURVEY Appendix B.8 and the exhaustive BCH(15,5) sweep in tests/golden/.
 *
 * Restated (reference file:line):
 *   generator = LCM of minimal polynomials ... src/bch.c:168-207, 209-286
 *   word encoder ............................ src/bch.c:352-380
 *   syndromes / BM / Chien / verify ......... src/bch.c:25-166, 382-435
 *   byte wrappers ........................... src/encode.c:199-234, src/decode.c:542-594
 *
 * The reference keeps the whole codeword in one uint32_t and shifts by bit positions up to n-1,
 * so encode/decode are undefined for n >= 32; construction however is well defined
 * (32-bit truncating arithmetic) and is mirrored here so that the getters agree for m = 6..8.
 * orc_bch_encode_word / orc_bch_decode_word refuse n >= 32.
 */
#include "oracle_internal.h"

#include <string.h>

#define BCH_ARR 64 /* fixed work-array length of the reference (src/bch.c:12) */

static uint32_t minimal_poly(const orc_field_t *f, int e)
{
    uint16_t c[BCH_ARR];
    int deg = 0, conj = e;
    uint32_t bits = 0;

    memset(c, 0, sizeof c);
    c[0] = 1;
    do { /* multiply c(x) by (x + alpha^conj) */
        uint16_t root = f->alpha_to[conj];
        for (int j = deg; j >= 0; j--) {
            if (j + 1 < BCH_ARR)
                c[j + 1] ^= c[j];
            if (c[j] != 0 && root != 0)
                c[j] = f->alpha_to[(f->index_of[c[j]] + f->index_of[root]) % f->nn];
            else
                c[j] = 0;
        }
        deg++;
        conj = (conj * 2) % f->nn;
    } while (conj != e);
    for (int i = 0; i <= deg; i++)
        if (c[i] == 1)
            bits |= 1u << i;
    return bits;
}

static int degree32(uint32_t p)
{
    if (p == 0)
        return -1;
    return 31 - __builtin_clz(p);
}

int orc_bch_init(orc_bch_t *b, uint8_t m, uint16_t poly, uint8_t t)
{
    uint8_t seen[256];
    uint32_t gen = 1;
    int gen_deg = 0;

    memset(b, 0, sizeof *b);
    if (m < 3 || m > 16 || t < 1 || t > 16) /* src/bch.c:293-299 */
        return -1;
    if (orc_field_init(&b->f, m, poly) != 0)
        return -1;
    b->t = t;
    b->n = (uint16_t)((1u << m) - 1u);

    memset(seen, 0, sizeof seen);
    for (int i = 1; i <= 2 * t; i++) {
        int e = i % b->f.nn, c;
        uint32_t mp, prod = 0;
        if (seen[e])
            continue;
        c = e;
        do {
            seen[c] = 1;
            c = (c * 2) % b->f.nn;
        } while (c != e);
        mp = minimal_poly(&b->f, e);
        for (int s = 0; s <= gen_deg; s++) /* 32-bit truncating carry-less product, src/bch.c:209-221 */
            if (gen & (1u << s))
                prod ^= mp << s;
        gen = prod;
        gen_deg = degree32(gen); /* -1 for the zero polynomial, src/bch.c:223-239 */
    }
    b->gen = gen;
    b->gen_deg = (uint16_t)gen_deg;
    b->parity_bits = (uint16_t)gen_deg;
    b->k = (uint16_t)(b->n - b->parity_bits);
    return 0;
}

int orc_bch_encode_word(const orc_bch_t *b, uint32_t data, uint32_t *codeword)
{
    uint32_t shifted, rem;
    if (b->n >= 32)
        return 0; /* undefined in the reference */
    if (data >= (1u << b->k))
        return 0;
    shifted = data << b->parity_bits;
    rem = shifted;
    for (int i = b->n - 1; i >= (int)b->gen_deg; i--)
        if (rem & (1u << i))
            rem ^= b->gen << (i - b->gen_deg);
    *codeword = shifted ^ rem;
    return 1;
}

static int bch_syndromes(const orc_bch_t *b, uint32_t w, uint16_t *s)
{
    int any = 0;
    for (int i = 0; i < 2 * b->t; i++) {
        s[i] = 0;
        for (int j = 0; j < b->n; j++)
            if (w & (1u << j))
                s[i] ^= b->f.alpha_to[((i + 1) * j) % b->f.nn];
        if (s[i])
            any = 1;
    }
    return any;
}

static uint16_t poly_eval(const orc_field_t *f, const uint16_t *p, int deg, uint16_t x)
{
    uint16_t acc = 0, lx;
    if (x == 0)
        return p[0];
    lx = f->index_of[x];
    for (int i = 0; i <= deg; i++)
        if (p[i] != 0)
            acc ^= f->alpha_to[(f->index_of[p[i]] + (lx * i) % f->nn) % f->nn];
    return acc;
}

/* textbook BM on polynomial-form syndromes with 64-entry arrays (src/bch.c:78-142) */
static int bch_bm(const orc_bch_t *b, const uint16_t *s, uint16_t *sigma)
{
    const orc_field_t *f = &b->f;
    uint16_t cur[BCH_ARR], prev[BCH_ARR], keep[BCH_ARR];
    uint16_t dprev = 1;
    int L = 0, shift = 1;

    memset(cur, 0, sizeof cur);
    memset(prev, 0, sizeof prev);
    cur[0] = prev[0] = 1;
    for (int it = 0; it < 2 * b->t; it++) {
        uint16_t d = s[it];
        for (int i = 1; i <= L; i++)
            if (cur[i] != 0 && s[it - i] != 0)
                d ^= f->alpha_to[(f->index_of[cur[i]] + f->index_of[s[it - i]]) % f->nn];
        if (d == 0) {
            shift++;
            continue;
        }
        {
            uint16_t mult = f->alpha_to[(f->nn - f->index_of[dprev] + f->index_of[d]) % f->nn];
            int grow = (2 * L <= it);
            if (grow)
                memcpy(keep, cur, sizeof keep);
            for (int i = 0; i < BCH_ARR - shift; i++)
                if (prev[i] != 0)
                    cur[i + shift] ^= f->alpha_to[(f->index_of[prev[i]] + f->index_of[mult]) % f->nn];
            if (grow) {
                memcpy(prev, keep, sizeof prev);
                L = it + 1 - L;
                dprev = d;
                shift = 1;
            } else {
                shift++;
            }
        }
    }
    memcpy(sigma, cur, sizeof cur);
    return L;
}

int orc_bch_decode_word(const orc_bch_t *b, uint32_t received, uint32_t *corrected, int32_t *num_errors)
{
    uint16_t s[BCH_ARR], sigma[BCH_ARR], pos[16];
    uint32_t fixed;
    int L, found = 0;

    if (b->n >= 32)
        return 0; /* undefined in the reference */
    memset(s, 0, sizeof s);
    received &= (1u << b->n) - 1u;
    *corrected = received;
    if (num_errors)
        *num_errors = 0;
    if (!bch_syndromes(b, received, s))
        return 1;
    L = bch_bm(b, s, sigma);
    if (L > b->t)
        return 0;
    for (int i = 0; i < b->n; i++) { /* Chien: position i <-> root alpha^{-i} (src/bch.c:144-166) */
        uint16_t x = b->f.alpha_to[(b->f.nn - i) % b->f.nn];
        if (poly_eval(&b->f, sigma, L, x) == 0) {
            pos[found++] = (uint16_t)i;
            if (found >= L)
                break;
        }
    }
    if (found != L)
        return 0;
    fixed = received;
    for (int i = 0; i < found; i++)
        fixed ^= 1u << pos[i];
    if (bch_syndromes(b, fixed, s))
        return 0;
    *corrected = fixed;
    if (num_errors)
        *num_errors = found;
    return 1;
}

/* poporon_encode for BCH (src/encode.c:199-234): big-endian bytes <-> one 32-bit word */
int orc_bch_encode_bytes(const orc_bch_t *b, const uint8_t *data, size_t size, uint8_t *parity)
{
    const unsigned db = (b->k + 7u) / 8u, pb = (b->parity_bits + 7u) / 8u;
    uint32_t d = 0, cw = 0, pv;
    if (size < db)
        return 0;
    for (unsigned i = 0; i < db && i < 4; i++)
        d |= (uint32_t)data[i] << (8 * (db - 1 - i));
    if (b->k < 32)
        d &= (1u << b->k) - 1u;
    if (!orc_bch_encode_word(b, d, &cw))
        return 0;
    pv = cw & ((1u << b->parity_bits) - 1u);
    memset(parity, 0, pb);
    for (unsigned i = 0; i < pb && i < 4; i++)
        parity[pb - 1 - i] = (uint8_t)(pv >> (8 * i));
    return 1;
}

/* poporon_decode for BCH (src/decode.c:542-594).  On failure neither data nor *corrected is touched. */
int orc_bch_decode_bytes(const orc_bch_t *b, uint8_t *data, size_t size, const uint8_t *parity, size_t *corrected)
{
    const unsigned db = (b->k + 7u) / 8u, pb = (b->parity_bits + 7u) / 8u;
    uint32_t d = 0, p = 0, fixed = 0, out;
    int32_t ne = 0;
    if (size < db)
        return 0;
    for (unsigned i = 0; i < db && i < 4; i++)
        d |= (uint32_t)data[i] << (8 * (db - 1 - i));
    if (b->k < 32)
        d &= (1u << b->k) - 1u;
    for (unsigned i = 0; i < pb && i < 4; i++)
        p |= (uint32_t)parity[i] << (8 * (pb - 1 - i));
    if (b->parity_bits < 32)
        p &= (1u << b->parity_bits) - 1u;
    if (!orc_bch_decode_word(b, (d << b->parity_bits) | p, &fixed, &ne))
        return 0;
    out = (fixed >> b->parity_bits) & ((1u << b->k) - 1u);
    for (unsigned i = 0; i < db && i < 4; i++)
        data[db - 1 - i] = (uint8_t)(out >> (8 * i));
    if (corrected)
        *corrected = ne > 0 ? (size_t)ne : 0;
    return 1;
}
