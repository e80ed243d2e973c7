#!/usr/bin/env python3
"""Generator of the XOR networks of the RS(255,223) syndrome transform (libpoporon_b200/csrc/cuda/rs255_fft_gen.cuh).

What is computed (replaces the arithmetic of reference src/decode.c:375-415, calculate_syndrome_u8, and - through
"parity = V^-1 * syndromes of the message" - of src/encode.c:120-144, rs_encode):

    S_j = sum_e r_e alpha^(j e),   j = 1..32,   e = exponent of the symbol (last byte of the codeword: e = 0)

Cooley-Tukey over 255 = 15 * 17.  Write e = 17 d + c (d < 15, c < 17) and beta = alpha^17 (order 15, a generator of
the subfield GF(16)):

    R_c[k] = sum_d r_(17 d + c) beta^(d k)          k = 0..14      "stage 1": a 15-point DFT with GF(16) twiddles
    S_j    = sum_c alpha^(j c) R_c[j mod 15]                        "stage 2": Horner in c with the multiplier alpha^j

Stage 1 is 15 times smaller than the 32 x 255 Horner scheme because its twiddles live in GF(16): in a basis
{b0..b3, theta b0..theta b3} of GF(256) over GF(2) (a "tower" basis) a GF(16) constant acts on the two 4-bit halves
separately with the same 4x4 matrix, and 15 = 3 * 5 splits the DFT again (Good-Thomas).  All networks are straight-line
programs of 2- and 3-input XORs on bit-planes, found here with Paar's greedy common-subexpression elimination and
checked here bit-exactly against the matrices they implement, which in turn are checked against the direct formula.

Run:  python tools/gen_rs255_fft.py [--out libpoporon_b200/csrc/cuda/rs255_fft_gen.cuh] [--search]
"""
from __future__ import annotations

import argparse
import itertools
import os
import random
import re
import sys

POLY = 0x11D

EXP = [0] * 510
LOG = [0] * 256
_x = 1
for _i in range(255):
    EXP[_i] = _x
    LOG[_x] = _i
    _x <<= 1
    if _x & 0x100:
        _x ^= POLY
for _i in range(255, 510):
    EXP[_i] = EXP[_i - 255]


def gmul(a: int, b: int) -> int:
    if a == 0 or b == 0:
        return 0
    return EXP[LOG[a] + LOG[b]]


def gpow(e: int) -> int:
    return EXP[e % 255]


def ginv(a: int) -> int:
    return EXP[(255 - LOG[a]) % 255]


# ---------------------------------------------------------------------------------------------------------
# GF(2) matrices as lists of row bitmasks (bit i of a row = coefficient of input i)
# ---------------------------------------------------------------------------------------------------------
def mat_from_cols(cols, nrows):
    rows = [0] * nrows
    for j, c in enumerate(cols):
        for i in range(nrows):
            if (c >> i) & 1:
                rows[i] |= 1 << j
    return rows


def mat_mul(a, b, ncols_b):
    """a (rows over inputs of size len(b)) * b (rows over ncols_b inputs)"""
    out = []
    for ra in a:
        acc = 0
        k = 0
        while ra:
            if ra & 1:
                acc ^= b[k]
            ra >>= 1
            k += 1
        out.append(acc)
    return out


def mat_inv(a, n):
    rows = [a[i] | (1 << (n + i)) for i in range(n)]
    for col in range(n):
        piv = next(r for r in range(col, n) if (rows[r] >> col) & 1)
        rows[col], rows[piv] = rows[piv], rows[col]
        for r in range(n):
            if r != col and (rows[r] >> col) & 1:
                rows[r] ^= rows[col]
    return [(r >> n) & ((1 << n) - 1) for r in rows]


def mat_apply(rows, x):
    y = 0
    for i, r in enumerate(rows):
        if bin(r & x).count("1") & 1:
            y |= 1 << i
    return y


def mulmat_std(c):
    """8x8 matrix of x -> c*x in the standard (polynomial) basis"""
    return mat_from_cols([gmul(c, 1 << i) for i in range(8)], 8)


# ---------------------------------------------------------------------------------------------------------
# straight-line XOR programs
# ---------------------------------------------------------------------------------------------------------
class SLP:
    """inputs are variables 0..nin-1; ops[i] = tuple of 2 or 3 source variables defining variable nin+i;
    outs[r] = variable holding output r (or -1 for constant zero)"""

    def __init__(self, nin, ops, outs):
        self.nin, self.ops, self.outs = nin, ops, outs

    def cost(self):
        return len(self.ops)

    def matrix(self):
        val = [1 << i for i in range(self.nin)]
        for srcs in self.ops:
            v = 0
            for s in srcs:
                v ^= val[s]
            val.append(v)
        return [0 if o < 0 else val[o] for o in self.outs]


def paar(rows, nin):
    """Paar's greedy CSE: returns (pair_ops, rows over the extended variable set)"""
    rows = list(rows)
    nvar = nin
    ops = []
    while True:
        # most frequent pair
        best, bestc = None, 1
        cols = [0] * nvar  # column bitmask over rows
        for r, row in enumerate(rows):
            k = 0
            rr = row
            while rr:
                if rr & 1:
                    cols[k] |= 1 << r
                rr >>= 1
                k += 1
        live = [k for k in range(nvar) if cols[k]]
        for ia in range(len(live)):
            a = live[ia]
            ca = cols[a]
            for ib in range(ia + 1, len(live)):
                b = live[ib]
                c = bin(ca & cols[b]).count("1")
                if c > bestc:
                    best, bestc = (a, b), c
        if best is None:
            break
        a, b = best
        v = nvar
        nvar += 1
        ops.append((a, b))
        both = (1 << a) | (1 << b)
        for r in range(len(rows)):
            if rows[r] & both == both:
                rows[r] = (rows[r] & ~both) | (1 << v)
    return ops, rows, nvar


def build_slp(rows, nin):
    """matrix -> SLP of 2/3-input XORs (Paar + fusion of single-use nodes + ternary row sums)"""
    pair_ops, rrows, nvar = paar(rows, nin)
    ops = [tuple(p) for p in pair_ops]  # variable nin+i
    outs = []
    for row in rrows:
        terms = [k for k in range(nvar) if (row >> k) & 1]
        if not terms:
            outs.append(-1)
            continue
        while len(terms) > 1:
            take = terms[:3]
            terms = terms[3:]
            ops.append(tuple(take))
            terms.insert(0, nin + len(ops) - 1)
        outs.append(terms[0])
    slp = SLP(nin, ops, outs)
    return fuse(slp)


def fuse(slp):
    """merge a 2-input node used exactly once (and not an output) into its consumer when that keeps <= 3 inputs"""
    nin = slp.nin
    ops = [list(o) for o in slp.ops]
    changed = True
    while changed:
        changed = False
        uses = {}
        for i, o in enumerate(ops):
            if o is None:
                continue
            for s in o:
                uses[s] = uses.get(s, 0) + 1
        outset = set(slp.outs)
        for i, o in enumerate(ops):
            if o is None:
                continue
            for s in list(o):
                if s >= nin and ops[s - nin] is not None and uses.get(s, 0) == 1 and s not in outset:
                    src = ops[s - nin]
                    if len(o) - 1 + len(src) <= 3:
                        o.remove(s)
                        o.extend(src)
                        ops[s - nin] = None
                        changed = True
                        break
    # renumber, dropping dead nodes, in dependency order (sources may now have larger indices than users)
    remap = {i: i for i in range(nin)}
    new_ops = []

    def emit(v):
        if v in remap:
            return remap[v]
        o = ops[v - nin]
        srcs = tuple(emit(s) for s in o)
        new_ops.append(srcs)
        remap[v] = nin + len(new_ops) - 1
        return remap[v]

    outs = [(-1 if o < 0 else emit(o)) for o in slp.outs]
    return SLP(nin, new_ops, outs)


def check_slp(slp, rows):
    assert slp.matrix() == list(rows), "SLP does not implement its matrix"


# ---------------------------------------------------------------------------------------------------------
# the algebra
# ---------------------------------------------------------------------------------------------------------
BETA = gpow(17)          # order 15
GF16 = [0] + [gpow(17 * i) for i in range(15)]


class Tower:
    def __init__(self, b, theta):
        self.b = list(b)
        self.theta = theta
        self.vecs = list(b) + [gmul(theta, x) for x in b]
        self.B = mat_from_cols(self.vecs, 8)       # std = B * tower
        self.C = mat_inv(self.B, 8)                # tower = C * std

    def mul_tower(self, c):
        """8x8 matrix of multiplication by c in tower coordinates"""
        return mat_mul(mat_mul(self.C, mulmat_std(c), 8), self.B, 8)

    def mul16(self, c):
        """4x4 matrix of a GF(16) constant on one half"""
        m = self.mul_tower(c)
        lo = [m[i] & 15 for i in range(4)]
        hi = [(m[4 + i] >> 4) & 15 for i in range(4)]
        assert lo == hi and all((m[i] >> 4) == 0 for i in range(4)) and all((m[4 + i] & 15) == 0 for i in range(4))
        return lo


def independent(vs):
    seen = {0}
    for v in vs:
        if v in seen:
            return False
        seen |= {s ^ v for s in seen}
    return True


def dft_block_rows(tw, npts, root):
    """npts-point DFT with twiddles root^(i*k) in GF(16), on 4-bit halves: (4*npts) x (4*npts) matrix"""
    rows = [0] * (4 * npts)
    for k in range(npts):
        for d in range(npts):
            c = 1
            for _ in range((d * k) % npts):
                c = gmul(c, root)
            m = tw.mul16(c)
            for i in range(4):
                rows[4 * k + i] |= m[i] << (4 * d)
    return rows


def horner_rows(tw, j):
    """acc' = alpha^j * acc + r : 8 x 16 matrix over [acc(8) | r(8)]"""
    m = tw.mul_tower(gpow(j))
    return [m[i] | (1 << (8 + i)) for i in range(8)]


def cost_of(tw, full=True):
    c_in = build_slp(tw.C, 8).cost()
    w3 = gpow(17 * 10)
    z5 = gpow(17 * 6)
    d3 = build_slp(dft_block_rows(tw, 3, w3), 12).cost()
    d5 = build_slp(dft_block_rows(tw, 5, z5), 20).cost()
    hs = sum(build_slp(horner_rows(tw, j), 16).cost() for j in range(1, 33)) if full else 0
    stage1 = 2 * (5 * d3 + 3 * d5)
    total = 255 * c_in + 17 * stage1 + 17 * hs
    return total, dict(conv=c_in, dft3=d3, dft5=d5, stage1_per_class=stage1, horner_all_j=hs)


def search():
    best = None
    nz16 = GF16[1:]
    bases = []
    for g in nz16:
        b = [1, g, gmul(g, g), gmul(g, gmul(g, g))]
        if independent(b):
            bases.append(b)
    for g in nz16:  # normal bases
        b = [g, gmul(g, g)]
        b.append(gmul(b[1], b[1]))
        b.append(gmul(b[2], b[2]))
        if independent(b):
            bases.append(b)
    # GF(4)-tower bases {1, w, t, w t}
    w = gpow(85)
    for t in nz16:
        b = [1, w, t, gmul(w, t)]
        if independent(b):
            bases.append(b)
    seen = set()
    cand = []
    for b in bases:
        key = tuple(b)
        if key in seen:
            continue
        seen.add(key)
        cand.append(b)
    print(f"{len(cand)} GF(16) bases", file=sys.stderr)
    results = []
    for b in cand:
        tw0 = Tower(b, 2)
        w3 = gpow(170)
        z5 = gpow(102)
        d3 = build_slp(dft_block_rows(tw0, 3, w3), 12).cost()
        d5 = build_slp(dft_block_rows(tw0, 5, z5), 20).cost()
        s1 = 2 * (5 * d3 + 3 * d5)
        # theta by conversion + Horner cost
        for theta in range(2, 256):
            if theta in GF16:
                continue
            tw = Tower(b, theta)
            c_in = build_slp(tw.C, 8).cost()
            results.append((255 * c_in + 17 * s1, b, theta, c_in, d3, d5))
    results.sort(key=lambda r: r[0])
    top = results[:40]
    scored = []
    for _, b, theta, c_in, d3, d5 in top:
        tw = Tower(b, theta)
        total, parts = cost_of(tw)
        scored.append((total, b, theta, parts))
        print(total, b, theta, parts, file=sys.stderr)
    scored.sort(key=lambda r: r[0])
    return scored[0]


# chosen by --search (kept fixed so that the generated header is reproducible)
DEFAULT_BASIS = ([1, 10, 68, 146], 72)


# ---------------------------------------------------------------------------------------------------------
# emission
# ---------------------------------------------------------------------------------------------------------
def emit_slp_body(slp, in_names, out_names, indent="    ", tmp="t"):
    """C statements computing out_names from in_names (out_names may alias in_names: all outputs are assigned
    at the end from temporaries)"""
    lines = []
    name = {i: in_names[i] for i in range(slp.nin)}
    for i, srcs in enumerate(slp.ops):
        v = slp.nin + i
        nm = f"{tmp}{i}"
        name[v] = nm
        a = [name[s] for s in srcs]
        if len(a) == 2:
            lines.append(f"{indent}const uint32_t {nm} = PPX2({a[0]}, {a[1]});")
        else:
            lines.append(f"{indent}const uint32_t {nm} = PPX3({a[0]}, {a[1]}, {a[2]});")
    # outputs: assigned after all temporaries exist (outputs may alias inputs)
    final = []
    for r, o in enumerate(slp.outs):
        src = "0u" if o < 0 else name[o]
        if src != out_names[r]:
            final.append((out_names[r], src))
    dsts = {d for d, _ in final}
    keep = {}
    for _, src in final:
        if src in dsts and src not in keep:  # an input passed through to another output while being overwritten
            keep[src] = "k_" + re.sub(r"\W", "_", src)
    for src, nm in keep.items():
        lines.append(f"{indent}const uint32_t {nm} = {src};")
    for dst, src in final:
        lines.append(f"{indent}{dst} = {keep.get(src, src)};")
    return lines


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "libpoporon_b200",
                                                  "csrc", "cuda", "rs255_fft_gen.cuh"))
    ap.add_argument("--search", action="store_true")
    ap.add_argument("--basis", default=None, help="b0,b1,b2,b3,theta")
    args = ap.parse_args()

    if args.search:
        total, b, theta, parts = search()
        print("best:", total, b, theta, parts)
        return
    if args.basis:
        v = [int(x, 0) for x in args.basis.split(",")]
        b, theta = v[:4], v[4]
    else:
        b, theta = DEFAULT_BASIS
    tw = Tower(b, theta)
    total, parts = cost_of(tw)
    print("basis", b, "theta", theta, "cost", total, parts, file=sys.stderr)
    generate(tw, args.out, total, parts)


def direct_syndromes(r):
    """r[e] for e = 0..254 -> [S_1..S_32]"""
    out = []
    for j in range(1, 33):
        s = 0
        for e in range(255):
            if r[e]:
                s ^= gmul(r[e], gpow(j * e))
        out.append(s)
    return out


# ---------------------------------------------------------------------------------------------------------
# Chien search as a transform (rs255_solve_core.h): Lambda(alpha^j) for all 255 j from the 17 locator coefficients.
#   j <-> (u = j mod 17, w = j mod 15);  alpha^(j d) = beta^(8 w d) * gamma^(8 u d),  beta = alpha^17, gamma = alpha^15
#   Lambda(alpha^j) = sum_{d<15} beta^(8 w d) Z_d(u),   Z_d(u) = gamma^(8 u d) (Lambda_d + [d < 2] gamma^u Lambda_(d+15))
# i.e. for every u one 15-point DFT (the generated dft15 has root beta: its output k is w = 2 k mod 15) of 15 values
# that advance from u to u + 1 by constant multipliers: Z_d *= gamma^(8 d) for d >= 2, and the four running terms
# Lambda_15 gamma^u, Lambda_1 gamma^(8u), Lambda_16 gamma^(9u) (Z_0 = Lambda_0 + the first, Z_1 = the other two).
# ---------------------------------------------------------------------------------------------------------
GAMMA_E = 15


def chien_exponents():
    """alpha-exponents of the per-step multipliers: [Z_2 .. Z_14], then the three running terms"""
    zs = [(GAMMA_E * ((8 * d) % 17)) % 255 for d in range(2, 15)]
    return zs, [(GAMMA_E * 1) % 255, (GAMMA_E * 8) % 255, (GAMMA_E * 9) % 255]


def chien_j_of(u, k):
    """the j of DFT output k in step u"""
    w = (2 * k) % 15
    return (120 * u + 136 * w) % 255


def check_chien_algebra(rnd):
    zs, (eb, ec, ed) = chien_exponents()
    for _ in range(20):
        lam = [rnd.randrange(256) for _ in range(17)]
        if rnd.random() < 0.5:  # plant roots
            lam = [1]
            for _r in range(rnd.randrange(1, 17)):
                x = gpow(rnd.randrange(255))
                lam = [a ^ gmul(b, x) for a, b in zip(lam + [0], [0] + lam)]
            lam = (lam + [0] * 17)[:17]
        direct = [0] * 255
        for j in range(255):
            v = 0
            for e in range(17):
                v ^= gmul(lam[e], gpow(j * e))
            direct[j] = v
        z = list(lam[:15])
        tb, tc, td = lam[15], lam[1], lam[16]
        seen = set()
        for u in range(17):
            z[0] = lam[0] ^ tb
            z[1] = tc ^ td
            for k in range(15):
                v = 0
                for d in range(15):
                    v ^= gmul(z[d], gpow(17 * ((d * k) % 15)))
                j = chien_j_of(u, k)
                assert j % 17 == u and j not in seen
                seen.add(j)
                assert v == direct[j], (u, k, j)
            for i, d in enumerate(range(2, 15)):
                z[d] = gmul(z[d], gpow(zs[i]))
            tb, tc, td = gmul(tb, gpow(eb)), gmul(tc, gpow(ec)), gmul(td, gpow(ed))
        assert len(seen) == 255


def generate(tw, path, total, parts):
    rnd = random.Random(1)
    check_chien_algebra(rnd)
    # ---- matrices and their programs -------------------------------------------------------------------
    conv_in = build_slp(tw.C, 8)
    check_slp(conv_in, tw.C)
    conv_out = build_slp(tw.B, 8)
    check_slp(conv_out, tw.B)
    w3 = gpow(17 * 10)
    z5 = gpow(17 * 6)
    r3 = dft_block_rows(tw, 3, w3)
    r5 = dft_block_rows(tw, 5, z5)
    dft3 = build_slp(r3, 12)
    check_slp(dft3, r3)
    dft5 = build_slp(r5, 20)
    check_slp(dft5, r5)
    hs = {}
    for j in range(1, 33):
        rows = horner_rows(tw, j)
        hs[j] = build_slp(rows, 16)
        check_slp(hs[j], rows)

    # ---- whole-transform check on random words, matrices only (the SLPs equal their matrices) ----------
    def to_tower(x):
        return mat_apply(tw.C, x)

    def from_tower(x):
        return mat_apply(tw.B, x)

    def m16(c, h):
        return mat_apply(tw.mul16(c), h)

    for _ in range(3):
        r = [rnd.randrange(256) for _ in range(255)]
        want = direct_syndromes(r)
        acc = [0] * 33
        for c in range(16, -1, -1):
            x = [to_tower(r[17 * d + c]) for d in range(15)]
            R = []
            for half in range(2):
                h = [(v >> (4 * half)) & 15 for v in x]
                # Good-Thomas with the in-place index map d = (10 d3 + 6 d5) mod 15
                y = list(h)
                for d5 in range(5):
                    idx = [(10 * d3 + 6 * d5) % 15 for d3 in range(3)]
                    vin = sum(y[idx[i]] << (4 * i) for i in range(3))
                    vout = mat_apply(r3, vin)
                    for i in range(3):
                        y[idx[i]] = (vout >> (4 * i)) & 15
                for k3 in range(3):
                    idx = [(10 * k3 + 6 * d5) % 15 for d5 in range(5)]
                    vin = sum(y[idx[i]] << (4 * i) for i in range(5))
                    vout = mat_apply(r5, vin)
                    for i in range(5):
                        y[idx[i]] = (vout >> (4 * i)) & 15
                R.append(y)
            Rc = [R[0][k] | (R[1][k] << 4) for k in range(15)]
            # reference for stage 1
            for k in range(15):
                s = 0
                for d in range(15):
                    s ^= gmul(r[17 * d + c], gpow(17 * d * k))
                assert from_tower(Rc[k]) == s, "stage 1 mismatch"
            for j in range(1, 33):
                acc[j] = mat_apply(horner_rows(tw, j), acc[j] | (Rc[j % 15] << 8))
        got = [from_tower(acc[j]) for j in range(1, 33)]
        assert got == want, "transform mismatch"

    # ---- encoder solve: parity symbols p_t (coefficient of x^t) from S_j = (m x^n)(alpha^j), j = 1..n ---------
    # V[j][t] = alpha^(j t), j = 1..n, t = 0..n-1;  p = V^-1 S.  n = 32 (the default code) and the reference's other
    # documented codes RS(255,239), RS(255,247): n = 16, 8 (reference README.md:417-419).
    def solve_masks(n):
        V = [[gpow(j * t) for t in range(n)] for j in range(1, n + 1)]
        A = [row[:] + [1 if i == k else 0 for k in range(n)] for i, row in enumerate(V)]
        for col in range(n):
            piv = next(r_ for r_ in range(col, n) if A[r_][col])
            A[col], A[piv] = A[piv], A[col]
            iv = ginv(A[col][col])
            A[col] = [gmul(iv, v) for v in A[col]]
            for r_ in range(n):
                if r_ != col and A[r_][col]:
                    f = A[r_][col]
                    A[r_] = [a ^ gmul(f, bb) for a, bb in zip(A[r_], A[col])]
        W = [row[n:] for row in A]  # W[t][j-1]
        # check on a random message
        msg = [0] * n + [rnd.randrange(256) for _ in range(255 - n)]
        S = direct_syndromes(msg)
        par = [0] * n
        for t in range(n):
            for jj in range(n):
                par[t] ^= gmul(W[t][jj], S[jj])
        cw = par + msg[n:]
        assert direct_syndromes(cw)[:n] == [0] * n, "encoder solve mismatch"
        # masks: output bit (parity byte index q = n - 1 - t in memory order, bit u) as a subset of the 8 tower planes
        # of S_j:  std(p_t) = sum_j Mstd(W[t][j]) * B * tower(S_j)
        mask = [[[0] * 8 for _ in range(n)] for _ in range(n)]  # [j-1][q][u]
        for jj in range(n):
            for t in range(n):
                m = mat_mul(mulmat_std(W[t][jj]), tw.B, 8)
                for u in range(8):
                    mask[jj][n - 1 - t][u] = m[u]
        return mask

    enc_mask = solve_masks(32)
    enc_mask16 = solve_masks(16)
    enc_mask8 = solve_masks(8)

    zs, run3 = chien_exponents()
    cm = {}
    for e in sorted(set(zs + run3)):
        rows = tw.mul_tower(gpow(e))
        cm[e] = build_slp(rows, 8)
        check_slp(cm[e], rows)

    # ---- emit -------------------------------------------------------------------------------------------
    L = []
    L.append("// rs255_fft_gen.cuh - GENERATED by tools/gen_rs255_fft.py; do not edit.")
    L.append("// XOR networks of the RS(255,223) syndrome transform S_j = sum_e r_e alpha^(j e), j = 1..32, as a")
    L.append("// 15 x 17 Cooley-Tukey split with GF(16) twiddles (see the generator for the algebra and the checks).")
    L.append(f"// tower basis: b = {tw.b}, theta = {tw.theta}; lane-ops per 32 codewords (model): {total}  {parts}")
    L.append("#pragma once")
    L.append("#include <stdint.h>")
    L.append("#ifndef PPX3")
    L.append("#if defined(__CUDA_ARCH__)")
    L.append("#define PPFN __device__ __forceinline__")
    L.append("static __device__ __forceinline__ uint32_t ppx3_(uint32_t a, uint32_t b, uint32_t c)")
    L.append("{")
    L.append("    uint32_t d;")
    L.append('    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));')
    L.append("    return d;")
    L.append("}")
    L.append("static __device__ __forceinline__ uint32_t ppx2_(uint32_t a, uint32_t b)")
    L.append("{")
    L.append("    uint32_t d;")
    L.append('    asm("xor.b32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));')
    L.append("    return d;")
    L.append("}")
    L.append("#define PPX3(a, b, c) ppx3_(a, b, c)")
    L.append("#define PPX2(a, b) ppx2_(a, b)")
    L.append("#else")
    L.append("#define PPFN inline")
    L.append("#define PPX3(a, b, c) ((a) ^ (b) ^ (c))")
    L.append("#define PPX2(a, b) ((a) ^ (b))")
    L.append("#endif")
    L.append("#endif")
    L.append("namespace rsfft {")
    L.append("")
    L.append(f"// standard-basis bit planes s[0..7] -> tower planes t[0..7] ({conv_in.cost()} ops)")
    L.append("PPFN void conv_in(const uint32_t (&s)[8], uint32_t (&t)[8])")
    L.append("{")
    L += emit_slp_body(conv_in, [f"s[{i}]" for i in range(8)], [f"t[{i}]" for i in range(8)], tmp="c")
    L.append("}")
    L.append(f"// tower planes -> standard-basis planes ({conv_out.cost()} ops)")
    L.append("PPFN void conv_out(const uint32_t (&t)[8], uint32_t (&s)[8])")
    L.append("{")
    L += emit_slp_body(conv_out, [f"t[{i}]" for i in range(8)], [f"s[{i}]" for i in range(8)], tmp="c")
    L.append("}")
    L.append("")
    L.append(f"// 3-point DFT (root beta^10) on one 4-plane half, in place ({dft3.cost()} ops)")
    L.append("PPFN void dft3(uint32_t (&a)[4], uint32_t (&b)[4], uint32_t (&c)[4])")
    L.append("{")
    nm3 = [f"{v}[{i}]" for v in "abc" for i in range(4)]
    L += emit_slp_body(dft3, nm3, nm3, tmp="u")
    L.append("}")
    L.append(f"// 5-point DFT (root beta^6) on one 4-plane half, in place ({dft5.cost()} ops)")
    L.append("PPFN void dft5(uint32_t (&a)[4], uint32_t (&b)[4], uint32_t (&c)[4], uint32_t (&d)[4], uint32_t (&e)[4])")
    L.append("{")
    nm5 = [f"{v}[{i}]" for v in "abcde" for i in range(4)]
    L += emit_slp_body(dft5, nm5, nm5, tmp="u")
    L.append("}")
    L.append("")
    L.append("// 15-point DFT with twiddles beta^(d k), beta = alpha^17, on one half: x[d] -> x[k], in place")
    L.append(f"// ({5 * dft3.cost() + 3 * dft5.cost()} ops)")
    L.append("PPFN void dft15(uint32_t (&x)[15][4])")
    L.append("{")
    for d5 in range(5):
        idx = [(10 * d3 + 6 * d5) % 15 for d3 in range(3)]
        L.append(f"    dft3(x[{idx[0]}], x[{idx[1]}], x[{idx[2]}]);")
    for k3 in range(3):
        idx = [(10 * k3 + 6 * d5) % 15 for d5 in range(5)]
        L.append(f"    dft5(x[{idx[0]}], x[{idx[1]}], x[{idx[2]}], x[{idx[3]}], x[{idx[4]}]);")
    L.append("}")
    L.append("")
    L.append("// Horner step of stage 2 for syndrome j: acc <- alpha^j * acc + r (tower planes)")
    L.append("template <int J>")
    L.append("PPFN void hstep(uint32_t (&acc)[8], const uint32_t (&r)[8]);")
    for j in range(1, 33):
        L.append(f"template <>  // {hs[j].cost()} ops")
        L.append(f"PPFN void hstep<{j}>(uint32_t (&acc)[8], const uint32_t (&r)[8])")
        L.append("{")
        L += emit_slp_body(hs[j], [f"acc[{i}]" for i in range(8)] + [f"r[{i}]" for i in range(8)],
                           [f"acc[{i}]" for i in range(8)], tmp="h")
        L.append("}")
    L.append("")
    L.append("// Chien search as a transform (rs255_solve_core.h; algebra and check in the generator): in-place product of")
    L.append("// a tower-basis symbol with alpha^E, for the per-step multipliers of the 15 DFT inputs")
    L.append("template <int E>")
    L.append("PPFN void cmulT(uint32_t (&x)[8]);")
    for e in sorted(cm):
        L.append(f"template <>  // {cm[e].cost()} ops")
        L.append(f"PPFN void cmulT<{e}>(uint32_t (&x)[8])")
        L.append("{")
        L += emit_slp_body(cm[e], [f"x[{i}]" for i in range(8)], [f"x[{i}]" for i in range(8)], tmp="m")
        L.append("}")
    L.append("// multiplier exponents: DFT inputs d = 2..14 (Z_d *= alpha^kChienZ[d - 2] per step), then the running terms")
    L.append("// Lambda_15 gamma^u, Lambda_1 gamma^(8u), Lambda_16 gamma^(9u)")
    L.append("constexpr int kChienZ[13] = {" + ", ".join(str(e) for e in zs) + "};")
    L.append(f"constexpr int kChienB = {run3[0]}, kChienC = {run3[1]}, kChienD = {run3[2]};")
    L.append("// DFT output k of step u is Lambda(alpha^j) with j = (120 u + kChienJ[k]) mod 255")
    L.append("constexpr int kChienJ[15] = {" + ", ".join(str((136 * ((2 * k) % 15)) % 255) for k in range(15)) + "};")
    L.append(f"// model: per step 2 x dft15 + {sum(cm[e].cost() for e in zs) + sum(cm[e].cost() for e in run3)} multiplier ops")
    L.append("")
    L.append("// encoder: parity byte q (memory order), bit u = XOR over j of the subset kEncMask[j-1][q][u] of the 8 tower")
    L.append("// planes of S_j, S_j = syndromes of message * x^32  (parity = V^-1 S, V[j][t] = alpha^(j t))")
    for n_, mk in ((32, enc_mask), (16, enc_mask16), (8, enc_mask8)):
        sfx = "" if n_ == 32 else str(n_)
        if n_ != 32:
            L.append(f"// the same for the {n_}-root code RS(255,{255 - n_}): parity = V^-1 S over S_1..S_{n_}")
        L.append(f"struct EncMasks{sfx} {{")
        L.append(f"    uint8_t m[{n_}][{n_}][8];")
        L.append("};")
        L.append(f"struct EncM{sfx} {{")
        L.append(f"    static constexpr EncMasks{sfx} v = {{{{")
        for jj in range(n_):
            L.append("        {")
            for q in range(n_):
                L.append("            {" + ", ".join(f"0x{mk[jj][q][u]:02X}" for u in range(8)) + "},")
            L.append("        },")
        L.append("    }};")
        L.append("};")
    L.append("")
    L.append("} // namespace rsfft")
    with open(path, "w") as f:
        f.write("\n".join(L) + "\n")
    print(f"wrote {path}: conv_in {conv_in.cost()}, conv_out {conv_out.cost()}, dft3 {dft3.cost()}, dft5 {dft5.cost()}, "
          f"horner sum {sum(h.cost() for h in hs.values())}", file=sys.stderr)


if __name__ == "__main__":
    main()
