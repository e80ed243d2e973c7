"""Generates the committed golden fixtures in tests/golden/*.npz by running the UNMODIFIED reference
(oracle/_ref/libpoporon_ref.so, built from /root/reference by oracle/Makefile) on seeded inputs.

Run here (the container that has /root/reference):  python tests/golden/make_golden.py
The fixtures store inputs AND outputs, so the tests that consume them need neither the reference
nor numpy RNG stability.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oraclelib import MODE_ERASURE, MODE_EXTSYN, MODE_PLAIN, Oracle, build_oracle  # noqa: E402


def rs_case(ref, name, params, size, n, seed, mode=MODE_PLAIN, emax=None):
    m, _, _, _, nr = params
    nn = (1 << m) - 1
    rng = np.random.default_rng(seed)
    data = rng.integers(0, 256, (n, size), dtype=np.uint8)
    parity = ref.rs_encode(data, params)
    cw = np.concatenate([data & nn, parity], 1)
    emax = nr // 2 + 4 if emax is None else emax
    out = dict(params=np.array(params, np.uint32), size=np.uint32(size), mode=np.uint32(mode), msg=data,
               parity=parity)
    pos = cnt = syn = None
    if mode == MODE_ERASURE:
        W = max(32, nr)
        pos = np.zeros((n, W), np.uint32)
        cnt = np.zeros(n, np.uint32)
        for i in range(n):
            c = int(rng.integers(0, min(nr, size) + 1))
            pp = rng.choice(size, c, replace=False)
            if rng.random() < 0.75:
                pp = np.sort(pp)
            pos[i, :c] = pp
            cnt[i] = c
            if rng.random() < 0.2:
                pos[i, c:] = rng.integers(0, size, W - c)
            cw[i, pp] ^= rng.integers(0, nn + 1, c, dtype=np.uint8)
            v = int(rng.choice([0, 0, 0, 0, 1, 2, 3]))
            vp = rng.choice(size + nr, v, replace=False)
            cw[i, vp] ^= rng.integers(1, nn + 1, v, dtype=np.uint8)
        out.update(pos=pos, cnt=cnt)
    else:
        for i in range(n):
            e = int(rng.integers(0, min(size + nr, emax) + 1))
            pp = rng.choice(size + nr, e, replace=False)
            cw[i, pp] ^= rng.integers(1, nn + 1, e, dtype=np.uint8)
    if mode == MODE_EXTSYN:
        syn = np.full((n, nr), nn, np.uint16)
        # half: true syndromes of the corrupted word (computed through a plain decode of a copy is not
        # exposed; use random sparse index-form syndromes instead), half: all-zero (= nn)
        for i in range(n):
            if rng.random() < 0.6:
                q = int(rng.integers(1, nr + 1))
                syn[i, rng.choice(nr, q, replace=False)] = rng.integers(0, nn, q)
        out.update(syn=syn)
    rd, rp, ok, cor = ref.rs_decode(cw[:, :size], cw[:, size:], params, mode, pos, cnt, syn)
    out.update(rx_data=cw[:, :size].copy(), rx_parity=cw[:, size:].copy(), out_data=rd, out_parity=rp, ok=ok,
               corrected=cor)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "ok-rate", ok.mean(), "max corrected", cor.max())


def bch_exhaustive(ref, name, m, poly, t, dword):
    info = ref.bch_info(m, poly, t)
    n, k, p = info["n"], info["k"], info["parity_bits"]
    db, pb = (k + 7) // 8, (p + 7) // 8
    d = np.array([[(dword >> (8 * (db - 1 - i))) & 0xFF for i in range(db)]], np.uint8)
    par, _ = ref.bch_encode(d, pb, m, poly, t)
    E = np.arange(1 << n, dtype=np.uint32)
    ed, ep = E >> p, E & ((1 << p) - 1)
    dd = np.stack([((ed >> (8 * (db - 1 - i))) & 0xFF).astype(np.uint8) for i in range(db)], 1) ^ d
    pp = np.stack([((ep >> (8 * (pb - 1 - i))) & 0xFF).astype(np.uint8) for i in range(pb)], 1) ^ par
    od, ok, cor = ref.bch_decode(dd, pp, m, poly, t)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), params=np.array([m, poly, t, n, k, p], np.uint32),
                        rx_data=dd, rx_parity=pp, out_data=od, ok=ok, corrected=cor, clean_parity=par)
    print(name, int(ok.sum()), np.bincount(cor[ok == 1]))


def bch_random(ref, name, m, poly, t, N, seed):
    info = ref.bch_info(m, poly, t)
    n, k, p = info["n"], info["k"], info["parity_bits"]
    db, pb = (k + 7) // 8, (p + 7) // 8
    rng = np.random.default_rng(seed)
    msg = rng.integers(0, 256, (N, db), dtype=np.uint8)  # garbage above k bits on purpose
    par, eok = ref.bch_encode(msg, pb, m, poly, t)
    d2, p2 = msg.copy(), par.copy()
    for i in range(N):
        for bit in rng.choice(n, int(rng.integers(0, t + 3)), replace=False):
            if bit < p:
                p2[i, pb - 1 - bit // 8] ^= 1 << (bit % 8)
            else:
                q = bit - p
                d2[i, db - 1 - q // 8] ^= 1 << (q % 8)
    od, ok, cor = ref.bch_decode(d2, p2, m, poly, t)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), params=np.array([m, poly, t, n, k, p], np.uint32), msg=msg,
                        parity=par, enc_ok=eok, rx_data=d2, rx_parity=p2, out_data=od, ok=ok, corrected=cor)
    print(name, ok.mean())


def main():
    build_oracle()
    ref = Oracle("reference")
    D = (8, 0x11D, 1, 1, 32)
    rs_case(ref, "rs255_plain", D, 223, 768, 101)
    rs_case(ref, "rs255_short64", D, 64, 384, 102)
    rs_case(ref, "rs255_size1", D, 1, 128, 103)
    rs_case(ref, "rs255_erasure", D, 223, 768, 104, MODE_ERASURE)
    rs_case(ref, "rs255_erasure_short", D, 100, 256, 105, MODE_ERASURE)
    rs_case(ref, "rs255_extsyn", D, 223, 384, 106, MODE_EXTSYN)
    rs_case(ref, "rs_ccsds_like", (8, 0x187, 112, 11, 32), 223, 256, 107)
    rs_case(ref, "rs_gf16", (4, 0x13, 1, 2, 8), 7, 256, 108)
    rs_case(ref, "rs_nroots10", (8, 0x11D, 1, 1, 10), 245, 256, 109)
    rs_case(ref, "rs_nroots127", (8, 0x11D, 1, 1, 127), 128, 96, 110)
    rs_case(ref, "rs_nroots16_prim2", (8, 0x11D, 1, 2, 16), 200, 256, 111)
    rs_case(ref, "rs_gf32_erasure", (5, 0x25, 3, 3, 6), 25, 256, 112, MODE_ERASURE)
    bch_exhaustive(ref, "bch15_5_exhaustive_d21", 4, 0x13, 3, 21)
    bch_random(ref, "bch15_5_random", 4, 0x13, 3, 2048, 201)
    bch_random(ref, "bch15_7_random", 4, 0x13, 2, 1024, 202)
    bch_random(ref, "bch31_21_random", 5, 0x25, 2, 2048, 203)
    bch_random(ref, "bch31_11_random", 5, 0x25, 5, 2048, 204)
    bch_random(ref, "bch31_6_random", 5, 0x25, 7, 1024, 205)
    bch_random(ref, "bch7_4_random", 3, 0xB, 1, 512, 206)


if __name__ == "__main__":
    main()
