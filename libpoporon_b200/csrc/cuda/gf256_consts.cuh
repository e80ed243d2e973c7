// gf256_consts.cuh — compile-time GF(2^8)/0x11D algebra for the RS(255,223) bitsliced kernels.
//
// Everything the XOR networks need is derived here with constexpr arithmetic from first principles
// (field polynomial 0x11D, alpha = 2, generator roots alpha^1..alpha^32, i.e. the reference's default
// config, src/poporon.c:281-284 and src/rs.c:59-80) — no table is copied from anywhere:
//
//   g(x)  = prod_{j=1..32} (x - alpha^j)            the RS(255,223) generator
//   gA(x) = prod_{j=1..16} (x - alpha^j),  gB(x) = prod_{j=17..32} (x - alpha^j),  g = gA * gB
//
// The kernels run two independent degree-16 LFSRs (one per warp "role") instead of one degree-32
// LFSR, because 256 bit-planes of state do not fit one thread's registers.  The Chinese remainder
// theorem recombines them for the encoder:  R = sA * FA + sB * FB  (mod g), where sX is the final
// state of LFSR X (= D(x) * x^16 mod gX) and FX = x^16 * EX mod g with EX = 1 mod gX, 0 mod gY.
// The decoder never recombines: "all syndromes zero" <=> sA == 0 and sB == 0, and the syndromes
// S_j = c(alpha^j) follow from the half that has alpha^j as a root.
#pragma once
#include <stdint.h>

namespace gf256 {

constexpr unsigned kPoly = 0x11D;

constexpr uint8_t mul(uint8_t a, uint8_t b)
{
    unsigned acc = 0, aa = a;
    for (int i = 0; i < 8; i++) {
        if (b & (1u << i))
            acc ^= aa;
        aa <<= 1;
        if (aa & 0x100)
            aa ^= kPoly;
    }
    return (uint8_t)acc;
}

struct ExpLog {
    uint8_t exp[256];
    uint8_t log[256];
};

constexpr ExpLog make_explog()
{
    ExpLog t{};
    unsigned x = 1;
    for (int i = 0; i < 255; i++) {
        t.exp[i] = (uint8_t)x;
        t.log[x] = (uint8_t)i;
        x <<= 1;
        if (x & 0x100)
            x ^= kPoly;
    }
    t.exp[255] = t.exp[0];
    t.log[0] = 0;
    return t;
}

constexpr ExpLog kExpLog = make_explog();

constexpr uint8_t pow_alpha(int e)
{
    e %= 255;
    if (e < 0)
        e += 255;
    return kExpLog.exp[e];
}

constexpr uint8_t inv(uint8_t a)
{
    return kExpLog.exp[(255 - kExpLog.log[a]) % 255];
}

// polynomials over GF(256), low degree first, fixed capacity 64
struct Poly {
    uint8_t c[64];
};

constexpr Poly zero_poly()
{
    Poly p{};
    for (int i = 0; i < 64; i++)
        p.c[i] = 0;
    return p;
}

constexpr Poly roots_poly(int first, int count) // prod (x - alpha^j), j = first .. first+count-1
{
    Poly p = zero_poly();
    p.c[0] = 1;
    int deg = 0;
    for (int j = first; j < first + count; j++) {
        const uint8_t r = pow_alpha(j);
        for (int k = deg + 1; k > 0; k--)
            p.c[k] = p.c[k - 1] ^ mul(p.c[k], r);
        p.c[0] = mul(p.c[0], r);
        deg++;
    }
    return p;
}

constexpr uint8_t eval(const Poly &p, int deg, uint8_t x)
{
    uint8_t acc = 0;
    for (int k = deg; k >= 0; k--)
        acc = mul(acc, x) ^ p.c[k];
    return acc;
}

constexpr Poly mul_poly(const Poly &a, int da, const Poly &b, int db)
{
    Poly r = zero_poly();
    for (int i = 0; i <= da; i++)
        for (int j = 0; j <= db; j++)
            r.c[i + j] ^= mul(a.c[i], b.c[j]);
    return r;
}

// a mod m (m monic of degree dm), a of degree <= da
constexpr Poly mod_poly(Poly a, int da, const Poly &m, int dm)
{
    for (int k = da; k >= dm; k--) {
        const uint8_t q = a.c[k];
        if (q)
            for (int i = 0; i <= dm; i++)
                a.c[k - dm + i] ^= mul(q, m.c[i]);
    }
    return a;
}

// Lagrange interpolation: the polynomial of degree < count taking value v[i] at alpha^(first+i)
constexpr Poly interpolate(int first, int count, const uint8_t *v)
{
    Poly r = zero_poly();
    for (int i = 0; i < count; i++) {
        // basis_i = prod_{k != i} (x - a_k) / (a_i - a_k)
        Poly b = zero_poly();
        b.c[0] = 1;
        int deg = 0;
        uint8_t denom = 1;
        const uint8_t ai = pow_alpha(first + i);
        for (int k = 0; k < count; k++) {
            if (k == i)
                continue;
            const uint8_t ak = pow_alpha(first + k);
            for (int d = deg + 1; d > 0; d--)
                b.c[d] = b.c[d - 1] ^ mul(b.c[d], ak);
            b.c[0] = mul(b.c[0], ak);
            deg++;
            denom = mul(denom, ai ^ ak);
        }
        const uint8_t s = mul(v[i], inv(denom));
        for (int d = 0; d <= deg; d++)
            r.c[d] ^= mul(b.c[d], s);
    }
    return r;
}

struct Consts {
    Poly g, gA, gB; // degrees 32, 16, 16 (monic)
    Poly FA, FB;    // degree < 32: recombination multipliers (see header comment)
};

constexpr Consts make_consts()
{
    Consts k{};
    k.gA = roots_poly(1, 16);
    k.gB = roots_poly(17, 16);
    k.g = mul_poly(k.gA, 16, k.gB, 16);
    // EA = gB * (gB^-1 mod gA): interpolate 1/gB(a) over the roots a of gA
    uint8_t va[16] = {}, vb[16] = {};
    for (int i = 0; i < 16; i++) {
        va[i] = inv(eval(k.gB, 16, pow_alpha(1 + i)));
        vb[i] = inv(eval(k.gA, 16, pow_alpha(17 + i)));
    }
    const Poly ia = interpolate(1, 16, va);  // gB^-1 mod gA
    const Poly ib = interpolate(17, 16, vb); // gA^-1 mod gB
    Poly EA = mul_poly(k.gB, 16, ia, 15);    // degree <= 31
    Poly EB = mul_poly(k.gA, 16, ib, 15);
    // FX = x^16 * EX mod g
    Poly sa = zero_poly(), sb = zero_poly();
    for (int i = 0; i < 32; i++) {
        sa.c[i + 16] = EA.c[i];
        sb.c[i + 16] = EB.c[i];
    }
    k.FA = mod_poly(sa, 47, k.g, 32);
    k.FB = mod_poly(sb, 47, k.g, 32);
    return k;
}

constexpr Consts kC = make_consts();

// sanity: EA + EB == 1 would be the textbook check; here we check the generator against the
// reference's known constant term and one CRT identity at compile time.
static_assert(kC.g.c[32] == 1, "g must be monic");
static_assert(kC.g.c[0] == pow_alpha(18), "g_0 = alpha^18 (SURVEY Appendix B.1: genpoly[0] = 18)");
static_assert(kC.g.c[1] == pow_alpha(251), "g_1 = alpha^251 (SURVEY Appendix B.1)");

// Row mask of "multiply by constant c": output bit t of c*x is the XOR of the input bits of x whose
// positions are set in rowmask(c, t).
constexpr uint8_t rowmask(uint8_t c, int t)
{
    unsigned m = 0;
    for (int a = 0; a < 8; a++)
        if (mul(c, (uint8_t)(1u << a)) & (1u << t))
            m |= 1u << a;
    return (uint8_t)m;
}

// ---- per-role LFSR update masks -------------------------------------------------------------------
// state p[0..15], p[0] = coefficient of x^15.  One step with input byte d (reference src/encode.c:120-144
// restated for a degree-16 generator): fb = d ^ p[0]; p'[j] = p[j+1] ^ fb*g_{15-j} (j < 15); p'[15] = fb*g_0.
template <int ROLE>
struct LfsrMasks {
    uint8_t m[16][8];
};

template <int ROLE>
constexpr LfsrMasks<ROLE> make_lfsr_masks()
{
    LfsrMasks<ROLE> r{};
    const Poly &gx = ROLE == 0 ? kC.gA : kC.gB;
    for (int j = 0; j < 16; j++)
        for (int t = 0; t < 8; t++)
            r.m[j][t] = rowmask(gx.c[15 - j], t);
    return r;
}

// ---- recombination map masks ------------------------------------------------------------------------
// Q = s * F mod g with s = sum_k s[k] x^(15-k).  Output coefficient of x^(31-o) (= parity byte o) receives
// s[k] * coef_{31-o}(x^(15-k) * F mod g).
template <int ROLE>
struct CrtMasks {
    uint8_t m[16][32][8]; // [input byte k][output parity byte o][bit t]
};

template <int ROLE>
constexpr CrtMasks<ROLE> make_crt_masks()
{
    CrtMasks<ROLE> r{};
    const Poly &F = ROLE == 0 ? kC.FA : kC.FB;
    for (int k = 0; k < 16; k++) {
        Poly s = zero_poly();
        for (int i = 0; i < 32; i++)
            s.c[i + (15 - k)] = F.c[i];
        const Poly col = mod_poly(s, 46, kC.g, 32);
        for (int o = 0; o < 32; o++)
            for (int t = 0; t < 8; t++)
                r.m[k][o][t] = rowmask(col.c[31 - o], t);
    }
    return r;
}

// ---- syndrome evaluation masks -----------------------------------------------------------------------
// Horner step acc = acc * alpha^j ^ s[k] for the 16 roots of the role's half.
template <int ROLE>
struct SynMasks {
    uint8_t m[16][8]; // [root index within the half][bit t]
};

template <int ROLE>
constexpr SynMasks<ROLE> make_syn_masks()
{
    SynMasks<ROLE> r{};
    for (int q = 0; q < 16; q++)
        for (int t = 0; t < 8; t++)
            r.m[q][t] = rowmask(pow_alpha(1 + 16 * ROLE + q), t);
    return r;
}

} // namespace gf256
