/*
 * oracle_rs.c — plain-C restatement of the reference Reed-Solomon path.  TEST INFRASTRUCTURE ONLY
 * (see oracle.h).  Parity status: pinned against the compiled reference and SURVEY Appendix B.
 *
 * What is restated (reference file:line):
 *   field tables ............ src/gf.c:29-86, gf_mod src/internal/common.h:102-110
 *   generator polynomial .... src/rs.c:59-80
 *   systematic encoder ...... src/encode.c:120-144 (scalar branch; the AVX2/NEON branches are
 *                             bit-identical)
 *   syndromes ............... src/decode.c:375-415
 *   locator / BM / Chien / Forney / verification / apply ... src/decode.c:17-230
 *   decode glue ............. src/decode.c:418-487, 596-612
 *
 * The restatement keeps the reference's integer widths wherever they are observable
 * (int16 location counter, uint16 truncation of every exponent sum handed to the modular
 * reduction, uint8 erasure root term) so that exotic (fcr, prim) settings wrap identically.
 */
#include "oracle_internal.h"

#include <stdlib.h>
#include <string.h>

/* v mod nn for any 16-bit v: this is what the reference's fold loop computes (common.h:102-110),
 * returned through a uint8_t like the reference does. */
static inline uint8_t redc(const orc_rs_t *rs, uint16_t v)
{
    return (uint8_t)(v % rs->nn);
}

int orc_field_init(orc_field_t *f, uint8_t m, uint16_t poly)
{
    /* src/gf.c:35 accepts 1..16, but field_size is a uint8_t (common.h:48): only m <= 8 can
     * pass the primitivity check at src/gf.c:79-83.  We reject m > 8 outright. */
    unsigned x = 1;
    if (m < 1 || m > 8)
        return -1;
    f->m = m;
    f->nn = (uint16_t)((1u << m) - 1u);
    f->poly = poly;
    memset(f->alpha_to, 0, sizeof f->alpha_to);
    memset(f->index_of, 0, sizeof f->index_of);
    f->index_of[0] = f->nn;      /* log(0) = A0 = nn  (src/gf.c:62) */
    f->alpha_to[f->nn] = 0;      /* antilog(A0) = 0   (src/gf.c:63) */
    for (unsigned i = 0; i < f->nn; i++) {
        f->index_of[x] = (uint16_t)i;
        f->alpha_to[i] = (uint16_t)x;
        x <<= 1;
        if (x & (1u << m))
            x ^= poly;
        x &= f->nn;
    }
    /* the walk must close on alpha^0 for a primitive polynomial (src/gf.c:79-83) */
    return (x == f->alpha_to[0]) ? 0 : -1;
}

int orc_rs_init(orc_rs_t *rs, const orc_rs_params_t *p)
{
    uint16_t root;
    unsigned i, j;
    uint16_t g[ORC_MAX_ROOTS + 1];

    memset(rs, 0, sizeof *rs);
    if (orc_field_init(&rs->f, p->symbol_size, p->gfpoly) != 0)
        return -1;
    if (p->prim == 0) /* src/poporon.c:78-82 */
        return -1;
    rs->nn = rs->f.nn;
    rs->fcr = p->fcr;
    rs->prim = p->prim;
    rs->nroots = p->nroots;

    /* g(x) = prod_{i<nroots} (x + alpha^{(fcr+i)*prim}), built in polynomial form with the running
     * exponent kept in 16 bits exactly like src/rs.c:59-76, then stored as logs (src/rs.c:78-80). */
    g[0] = 1;
    root = (uint16_t)(p->fcr * p->prim);
    for (i = 0; i < p->nroots; i++, root = (uint16_t)(root + p->prim)) {
        g[i + 1] = 1;
        for (j = i; j > 0; j--) {
            if (g[j] != 0)
                g[j] = g[j - 1] ^ rs->f.alpha_to[redc(rs, (uint16_t)(rs->f.index_of[g[j]] + root))];
            else
                g[j] = g[j - 1];
        }
        g[0] = rs->f.alpha_to[redc(rs, (uint16_t)(rs->f.index_of[g[0]] + root))];
    }
    for (i = 0; i <= p->nroots; i++)
        rs->genpoly[i] = rs->f.index_of[g[i]];

    /* iprim: smallest x with x*prim == 1 (mod nn), searched as 1, 1+nn, 1+2nn, ... in 16 bits
     * (src/poporon.c:84-93); more than 2*nn probes -> constructor fails. */
    {
        uint16_t probe;
        uint32_t iters = 0;
        for (probe = 1; (probe % p->prim) != 0; probe = (uint16_t)(probe + rs->nn)) {
            if (++iters > (uint32_t)rs->nn * 2u)
                return -1;
        }
        rs->iprim = (uint16_t)(probe / p->prim);
    }
    return 0;
}

/* parity <- (data(x) * x^nroots) mod g(x); parity[0] is the x^{nroots-1} coefficient. */
void orc_rs_encode(const orc_rs_t *rs, const uint8_t *data, size_t size, uint8_t *parity)
{
    const unsigned nr = rs->nroots;
    const uint16_t A0 = rs->nn;
    uint16_t i; /* the reference indexes with a uint16_t (src/encode.c:121) */

    memset(parity, 0, nr);
    for (i = 0; i < size; i++) {
        uint16_t fb = rs->f.index_of[((uint16_t)data[i] & rs->nn) ^ parity[0]];
        if (fb != A0) {
            for (unsigned j = 1; j < nr; j++)
                parity[j] ^= (uint8_t)rs->f.alpha_to[redc(rs, (uint16_t)(fb + rs->genpoly[nr - j]))];
        }
        memmove(parity, parity + 1, nr - 1);
        parity[nr - 1] = (fb != A0) ? (uint8_t)rs->f.alpha_to[redc(rs, (uint16_t)(fb + rs->genpoly[0]))] : 0;
    }
}

/* Horner evaluation of data||parity at the nroots generator roots; leaves syn[] in index form and
 * returns 1 when any syndrome is non-zero (src/decode.c:375-415). */
int orc_rs_syndromes(const orc_rs_t *rs, const uint8_t *data, size_t size, const uint8_t *parity, uint16_t *syn)
{
    const unsigned nr = rs->nroots;
    uint16_t any = 0;
    unsigned i;
    size_t j;

    for (i = 0; i < nr; i++)
        syn[i] = data[0] & rs->nn;
    for (j = 1; j < size + nr; j++) {
        const uint16_t sym = (j < size ? data[j] : parity[j - size]) & rs->nn;
        for (i = 0; i < nr; i++) {
            if (syn[i] == 0)
                syn[i] = sym;
            else
                syn[i] = sym ^ rs->f.alpha_to[redc(
                                   rs, (uint16_t)(rs->f.index_of[syn[i]] + (int)(rs->fcr + i) * (int)rs->prim))];
        }
    }
    for (i = 0; i < nr; i++) {
        any |= syn[i];
        syn[i] = rs->f.index_of[syn[i]];
    }
    return any != 0;
}

/*
 * Error/erasure correction from index-form syndromes.  Returns the reference's bool; *ncorr is
 * written exactly when the reference writes *errors_corrected (after the Chien stage).
 * erasure_apply != 0 selects the `corrections && erasure_positions` branch at src/decode.c:211-214.
 */
int orc_rs_correct(const orc_rs_t *rs, uint8_t *data, size_t size, uint8_t *parity, const uint16_t *syn,
                   uint32_t ecount, const uint32_t *epos, int erasure_apply, int16_t pad, size_t *ncorr)
{
    const unsigned nr = rs->nroots;
    const uint16_t A0 = rs->nn;
    const uint16_t *alpha_to = rs->f.alpha_to, *index_of = rs->f.index_of;
    uint16_t lambda[ORC_MAX_ROOTS + 1]; /* polynomial form during BM, index form afterwards */
    uint16_t b[ORC_MAX_ROOTS + 1];      /* index form */
    uint16_t t[ORC_MAX_ROOTS + 1];
    uint16_t omega[ORC_MAX_ROOTS + 1];
    uint16_t root[ORC_MAX_ROOTS + 1], loc[ORC_MAX_ROOTS + 1], reg[ORC_MAX_ROOTS + 1];
    uint32_t r, el;
    uint16_t deg, count;
    int16_t i, j, k;

    /* --- locator initialisation from erasures (src/decode.c:31-47) --- */
    memset(lambda, 0, sizeof lambda);
    lambda[0] = 1;
    if (ecount > 0) {
        lambda[1] = alpha_to[redc(rs, (uint16_t)(rs->prim * (rs->nn - 1 - (epos[0] + pad))))];
        for (i = 1; (uint32_t)i < ecount; i++) {
            uint8_t term = redc(rs, (uint16_t)(rs->prim * (rs->nn - 1 - (epos[i] + pad))));
            for (j = i + 1; j > 0; j--) {
                uint16_t lg = index_of[lambda[j - 1]];
                if (lg != A0)
                    lambda[j] ^= alpha_to[redc(rs, (uint16_t)(term + lg))];
            }
        }
    }
    for (i = 0; i < (int)nr + 1; i++)
        b[i] = index_of[lambda[i]];

    /* --- Berlekamp-Massey, Karn's formulation (src/decode.c:51-96) --- */
    r = ecount;
    el = ecount;
    while (++r <= nr) {
        uint16_t disc = 0;
        for (i = 0; (uint32_t)i < r; i++) {
            if (lambda[i] != 0 && syn[r - i - 1] != A0)
                disc ^= alpha_to[redc(rs, (uint16_t)(index_of[lambda[i]] + syn[r - i - 1]))];
        }
        disc = index_of[disc];
        if (disc == A0) {
            memmove(&b[1], b, nr * sizeof b[0]);
            b[0] = A0;
        } else {
            t[0] = lambda[0];
            for (i = 0; i < (int)nr; i++) {
                if (b[i] != A0)
                    t[i + 1] = lambda[i + 1] ^ alpha_to[redc(rs, (uint16_t)(disc + b[i]))];
                else
                    t[i + 1] = lambda[i + 1];
            }
            if (2 * el <= r + ecount - 1) {
                el = r + ecount - el;
                for (i = 0; i <= (int)nr; i++)
                    b[i] = (lambda[i] == 0) ? A0 : redc(rs, (uint16_t)(index_of[lambda[i]] - disc + rs->nn));
            } else {
                memmove(&b[1], b, nr * sizeof b[0]);
                b[0] = A0;
            }
            memcpy(lambda, t, (nr + 1) * sizeof t[0]);
        }
    }

    /* --- degree (src/decode.c:98-110) --- */
    deg = 0;
    for (i = 0; i < (int)nr + 1; i++) {
        lambda[i] = index_of[lambda[i]];
        if (lambda[i] != A0)
            deg = (uint16_t)i;
    }
    if (deg == 0)
        return 0;

    /* --- Chien search (src/decode.c:112-145) --- */
    memcpy(&reg[1], &lambda[1], nr * sizeof reg[0]);
    count = 0;
    for (i = 1, k = (int16_t)(rs->iprim - 1); i <= (int)rs->nn; i++, k = redc(rs, (uint16_t)(k + rs->iprim))) {
        uint16_t q = 1;
        for (j = (int16_t)deg; j > 0; j--) {
            if (reg[j] != A0) {
                reg[j] = redc(rs, (uint16_t)(reg[j] + j));
                q ^= alpha_to[reg[j]];
            }
        }
        if (q != 0)
            continue;
        if (k < pad)
            return 0;
        root[count] = (uint16_t)i;
        loc[count] = (uint16_t)k;
        if (++count == deg)
            break;
    }
    if (deg != count)
        return 0;

    /* --- evaluator omega = S*lambda mod x^deg (src/decode.c:147-158) --- */
    for (i = 0; i <= (int)deg - 1; i++) {
        uint16_t acc = 0;
        for (j = i; j >= 0; j--) {
            if (syn[i - j] != A0 && lambda[j] != A0)
                acc ^= alpha_to[redc(rs, (uint16_t)(syn[i - j] + lambda[j]))];
        }
        omega[i] = index_of[acc];
    }

    /* --- Forney magnitudes into b[] (the reference reuses `coefficients`, src/decode.c:159-191) --- */
    *ncorr = 0;
    for (j = (int16_t)(count - 1); j >= 0; j--) {
        uint16_t num1 = 0, num2, den = 0;
        for (i = (int16_t)(deg - 1); i >= 0; i--) {
            if (omega[i] != A0)
                num1 ^= alpha_to[redc(rs, (uint16_t)(omega[i] + i * root[j]))];
        }
        if (num1 == 0) {
            b[j] = 0;
            continue;
        }
        num2 = alpha_to[redc(rs, (uint16_t)(root[j] * (rs->fcr - 1) + rs->nn))];
        for (i = (int16_t)((deg < nr - 1 ? deg : nr - 1) & ~1u); i >= 0; i -= 2) {
            if (lambda[i + 1] != A0)
                den ^= alpha_to[redc(rs, (uint16_t)(lambda[i + 1] + i * root[j]))];
        }
        b[j] = alpha_to[redc(rs, (uint16_t)(index_of[num1] + index_of[num2] + rs->nn - index_of[den]))];
        (*ncorr)++;
    }

    /* --- the reference's extra re-syndrome check (src/decode.c:193-209) --- */
    for (i = 0; i < (int)nr; i++) {
        uint16_t acc = 0;
        for (j = 0; j < (int)count; j++) {
            if (b[j] == 0)
                continue;
            /* product formed in int, stored to an int16_t, added back in int, truncated to uint16 */
            k = (int16_t)((int64_t)(rs->fcr + i) * (int64_t)rs->prim * (int64_t)(rs->nn - loc[j] - 1));
            acc ^= alpha_to[redc(rs, (uint16_t)(index_of[b[j]] + k))];
        }
        if (acc != alpha_to[syn[i]])
            return 0;
    }

    /* --- apply (src/decode.c:211-227) --- */
    if (erasure_apply) {
        for (i = 0; i < (int)count; i++)
            data[epos[i]] ^= (uint8_t)b[i];
    } else {
        for (i = 0; i < (int)count; i++) {
            int32_t p = (int32_t)loc[i] - (int32_t)pad;
            if (p >= 0 && p < (int32_t)size)
                data[p] ^= (uint8_t)b[i];
            else if (p >= (int32_t)size && p < (int32_t)(size + nr))
                parity[p - size] ^= (uint8_t)b[i];
            else
                return 0;
        }
    }
    return 1;
}

/* poporon_decode for RS: src/decode.c:596-600 then 431-487.  *corrected is only touched when the
 * reference touches *corrected_num. */
int orc_rs_decode(const orc_rs_t *rs, int mode, uint8_t *data, size_t size, uint8_t *parity, uint32_t ecount,
                  const uint32_t *epos, const uint16_t *ext_syn, size_t *corrected)
{
    uint16_t syn[ORC_MAX_ROOTS + 1];
    size_t ncorr = 0;
    int ok = 0;
    int16_t pad;

    if (!data || !parity || size == 0)
        return 0; /* *corrected untouched */

    pad = (int16_t)((size_t)rs->nn - rs->nroots - size); /* src/decode.c:418-429 */
    if (pad < 0 || pad >= (int16_t)(rs->nn - rs->nroots))
        goto finish;

    if (mode == ORC_MODE_EXTSYN) {
        int dirty = 0;
        for (unsigned i = 0; i < rs->nroots; i++)
            if (ext_syn[i] != rs->nn) {
                dirty = 1;
                break;
            }
        if (dirty && !orc_rs_correct(rs, data, size, parity, ext_syn, 0, NULL, 0, pad, &ncorr))
            goto finish;
        ok = 1;
    } else if (mode == ORC_MODE_ERASURE) {
        ok = !orc_rs_syndromes(rs, data, size, parity, syn) ||
             orc_rs_correct(rs, data, size, parity, syn, ecount, epos, 1, pad, &ncorr);
    } else {
        ok = !orc_rs_syndromes(rs, data, size, parity, syn) ||
             orc_rs_correct(rs, data, size, parity, syn, 0, NULL, 0, pad, &ncorr);
    }
finish:
    if (corrected)
        *corrected = ncorr;
    return ok;
}
