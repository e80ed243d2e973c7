"""CPU tests of the oracle itself (-m "not gpu"): the plain-C restatement (oracle/liboracle.so) against
  (i)  the known-answer vectors of SURVEY.md Appendix B (extracted from the reference),
  (ii) the committed golden fixtures produced by the compiled reference (tests/golden/make_golden.py),
  (iii) the compiled reference itself, when oracle/_ref is present (this container; not the GPU box).
"""
import numpy as np
import pytest
from conftest import golden
from oraclelib import MODE_ERASURE, MODE_EXTSYN, MODE_PLAIN, RS_DEFAULT

RS_GOLDEN = ["rs255_plain", "rs255_short64", "rs255_size1", "rs255_erasure", "rs255_erasure_short", "rs255_extsyn",
             "rs_ccsds_like", "rs_gf16", "rs_nroots10", "rs_nroots127", "rs_nroots16_prim2", "rs_gf32_erasure"]
BCH_GOLDEN = ["bch15_5_exhaustive_d21", "bch15_5_random", "bch15_7_random", "bch31_21_random", "bch31_11_random",
              "bch31_6_random", "bch7_4_random"]

B1_GENPOLY = [18, 251, 215, 28, 80, 107, 248, 53, 84, 194, 91, 59, 176, 99, 203, 137, 43, 104, 137, 0, 44, 149, 148,
              218, 75, 11, 173, 254, 194, 109, 8, 11, 0]


def hello():
    d = np.zeros((1, 64), np.uint8)
    msg = b"Hello, Reed-Solomon!"
    d[0, :len(msg)] = np.frombuffer(msg, np.uint8)
    return d


def test_b1_generator_and_field(orc):
    info = orc.rs_info()
    assert list(info["genpoly"]) == B1_GENPOLY
    assert info["iprim"] == 1
    assert list(info["index_of"][:8]) == [255, 0, 1, 25, 2, 50, 26, 198]


def test_b2_hello_parity_and_two_errors(orc):
    d = hello()
    par = orc.rs_encode(d)
    assert par[0].tobytes().hex() == "31e6a711c1fc30242fd40fa62307b796856794fff72ac6a79c6fb63e70086d5b"
    bad = d.copy()
    bad[0, 0] ^= 0xFF
    bad[0, 10] ^= 0xAA
    od, op, ok, cor = orc.rs_decode(bad, par)
    assert ok[0] == 1 and cor[0] == 2 and (od == d).all() and (op == par).all()


def test_b3_b4_parities(orc):
    d3 = ((7 * np.arange(223) + 3) & 0xFF).astype(np.uint8)[None]
    assert orc.rs_encode(d3)[0].tobytes().hex() == "d412d836674222939b207358871d601b961fda8a215873da452db0d80b840f75"
    d4 = ((5 * np.arange(64) + 1) & 0xFF).astype(np.uint8)[None]
    p4 = orc.rs_encode(d4)
    assert p4[0].tobytes().hex() == "9b09ba782185d0dbf3b5af815a8e749b99932fb1a8a0b0fdd152258b1188062d"
    bad = d4.copy()
    bad[0, 0] ^= 0xFF
    bad[0, 10] ^= 0xAA
    od, op, ok, cor = orc.rs_decode(bad, p4)
    assert ok[0] == 1 and cor[0] == 2 and (od == d4).all()


def test_b5_sixteen_and_seventeen_errors(orc):
    d3 = ((7 * np.arange(223) + 3) & 0xFF).astype(np.uint8)[None]
    c = np.concatenate([d3, orc.rs_encode(d3)], 1)
    bad = c.copy()
    for i in range(15):
        bad[0, 13 * i] ^= i + 1
    bad[0, 230] ^= 0x80
    od, op, ok, cor = orc.rs_decode(bad[:, :223], bad[:, 223:])
    assert ok[0] == 1 and cor[0] == 16 and (np.concatenate([od, op], 1) == c).all()
    bad = c.copy()
    for i in range(17):
        bad[0, 13 * i] ^= i + 1
    od, op, ok, cor = orc.rs_decode(bad[:, :223], bad[:, 223:])
    assert ok[0] == 0 and cor[0] == 0 and (np.concatenate([od, op], 1) == bad).all()
    # data + parity error on the 64-byte hello message
    d = hello()
    par = orc.rs_encode(d)
    bd, bp = d.copy(), par.copy()
    bd[0, 5] ^= 1
    bp[0, 3] ^= 0x55
    od, op, ok, cor = orc.rs_decode(bd, bp)
    assert ok[0] == 1 and cor[0] == 2 and (od == d).all() and (op == par).all()


def test_b6_erasure_slot_order_quirk(orc):
    d3 = ((7 * np.arange(223) + 3) & 0xFF).astype(np.uint8)[None]
    par = orc.rs_encode(d3)
    bad = d3.copy()
    bad[0, 200] ^= 1
    bad[0, 100] ^= 2
    bad[0, 3] ^= 3
    pos = np.zeros((1, 32), np.uint32)
    pos[0, :3] = [200, 100, 3]
    od, op, ok, cor = orc.rs_decode(bad, par, RS_DEFAULT, MODE_ERASURE, pos, np.array([3], np.uint32))
    assert ok[0] == 1 and cor[0] == 3
    assert od[0, 3] ^ d3[0, 3] == 0x02 and od[0, 100] ^ d3[0, 100] == 0 and od[0, 200] ^ d3[0, 200] == 0x02
    pos[0, :3] = [3, 100, 200]
    od, op, ok, cor = orc.rs_decode(bad, par, RS_DEFAULT, MODE_ERASURE, pos, np.array([3], np.uint32))
    assert ok[0] == 1 and cor[0] == 3 and (od == d3).all()
    # 32 ascending erasures
    bad = d3.copy()
    pos = (6 * np.arange(32)).astype(np.uint32)[None]
    for i in range(32):
        bad[0, 6 * i] ^= i + 1
    od, op, ok, cor = orc.rs_decode(bad, par, RS_DEFAULT, MODE_ERASURE, pos, np.array([32], np.uint32))
    assert ok[0] == 1 and cor[0] == 32 and (od == d3).all()


def test_b7_size_edges(orc):
    d = np.arange(224, dtype=np.uint8)[None]
    par = orc.rs_encode(d)  # encode does not validate size
    assert par is not None
    od, op, ok, cor = orc.rs_decode(d, par)
    assert ok[0] == 0 and cor[0] == 0
    d1 = np.array([[0x5A]], np.uint8)
    p1 = orc.rs_encode(d1)
    bad = d1.copy()
    bad[0, 0] ^= 0x11
    od, op, ok, cor = orc.rs_decode(bad, p1)
    assert ok[0] == 1 and cor[0] == 1 and od[0, 0] == 0x5A


def test_a7_extsyn_single_error_syndromes(orc):
    # single error of magnitude 1 at data[0] of a full-length word -> syndromes 254, 253, ..., 223
    d = np.zeros((1, 223), np.uint8)
    par = orc.rs_encode(d)
    syn = np.arange(254, 222, -1).astype(np.uint16)[None]
    od, op, ok, cor = orc.rs_decode(d, par, RS_DEFAULT, MODE_EXTSYN, syn=syn)
    assert ok[0] == 1 and cor[0] == 1 and od[0, 0] == 1  # "corrects" the clean word at position 0
    od, op, ok, cor = orc.rs_decode(d, par, RS_DEFAULT, MODE_EXTSYN, syn=np.full((1, 32), 255, np.uint16))
    assert ok[0] == 1 and cor[0] == 0 and (od == d).all()


def test_b8_bch_known_answers(orc):
    for dval, hexpar in [(0, "0000"), (7, "02b2"), (14, "0053"), (21, "0247"), (28, "00a6"), (0xF5, "0247")]:
        par, ok = orc.bch_encode(np.array([[dval]], np.uint8), 2)
        assert ok[0] == 1 and par[0].tobytes().hex() == hexpar
    info = orc.bch_info()
    assert (info["n"], info["k"], info["parity_bits"], info["gen"]) == (15, 5, 10, 0x537)
    par, _ = orc.bch_encode(np.array([[21]], np.uint8), 2)
    od, ok, cor = orc.bch_decode(np.array([[21 ^ 0x0A]], np.uint8), par)
    assert ok[0] == 1 and cor[0] == 2 and od[0, 0] == 21
    # clean decode masks the pad bits of the data byte
    od, ok, cor = orc.bch_decode(np.array([[0xF5]], np.uint8), par)
    assert ok[0] == 1 and cor[0] == 0 and od[0, 0] == 0x15
    # five data-bit flips: failure, data untouched
    od, ok, cor = orc.bch_decode(np.array([[21 ^ 0x1F]], np.uint8), par)
    assert ok[0] == 0 and cor[0] == 0 and od[0, 0] == 21 ^ 0x1F


def test_bch_generators_a9(orc):
    want = {(4, 0x13, 1): (11, 0x13), (4, 0x13, 2): (7, 0x1D1), (4, 0x13, 3): (5, 0x537), (5, 0x25, 1): (26, 0x25),
            (5, 0x25, 2): (21, 0x769), (5, 0x25, 3): (16, 0x8FAF), (5, 0x25, 5): (11, 0x1626D5),
            (5, 0x25, 7): (6, 0x32DEA27)}
    for (m, poly, t), (k, gen) in want.items():
        info = orc.bch_info(m, poly, t)
        assert (info["k"], info["gen"]) == (k, gen)


@pytest.mark.parametrize("name", RS_GOLDEN)
def test_rs_golden(orc, name):
    g = golden(name)
    params = tuple(int(x) for x in g["params"])
    mode = int(g["mode"])
    assert (orc.rs_encode(g["msg"], params) == g["parity"]).all()
    kw = {}
    if mode == MODE_ERASURE:
        kw = dict(pos=g["pos"], cnt=g["cnt"])
    if mode == MODE_EXTSYN:
        kw = dict(syn=g["syn"])
    od, op, ok, cor = orc.rs_decode(g["rx_data"], g["rx_parity"], params, mode, **kw)
    assert (ok == g["ok"]).all() and (cor == g["corrected"]).all()
    assert (od == g["out_data"]).all() and (op == g["out_parity"]).all()


@pytest.mark.parametrize("name", BCH_GOLDEN)
def test_bch_golden(orc, name):
    g = golden(name)
    m, poly, t, n, k, p = (int(x) for x in g["params"])
    info = orc.bch_info(m, poly, t)
    assert (info["n"], info["k"], info["parity_bits"]) == (n, k, p)
    if "msg" in g:
        par, eok = orc.bch_encode(g["msg"], (p + 7) // 8, m, poly, t)
        assert (par == g["parity"]).all() and (eok == g["enc_ok"]).all()
    od, ok, cor = orc.bch_decode(g["rx_data"], g["rx_parity"], m, poly, t)
    assert (ok == g["ok"]).all() and (cor == g["corrected"]).all() and (od == g["out_data"]).all()


def test_b8_exhaustive_histogram():
    g = golden("bch15_5_exhaustive_d21")
    ok, cor = g["ok"], g["corrected"]
    assert int(ok.sum()) == 18432 and int((ok == 0).sum()) == 14336
    assert list(np.bincount(cor[ok == 1])) == [32, 480, 3360, 14560]


# ---------------------------------------------------------------- live differential vs the reference
def _inject(rng, cw, nn, emax):
    n, L = cw.shape
    for i in range(n):
        e = int(rng.integers(0, min(L, emax) + 1))
        pos = rng.choice(L, e, replace=False)
        cw[i, pos] ^= rng.integers(1, nn + 1, e, dtype=np.uint8)


@pytest.mark.parametrize("params", [(8, 0x11D, 1, 1, 32), (8, 0x187, 112, 11, 32), (8, 0x11D, 0, 1, 16),
                                    (4, 0x13, 1, 2, 8), (7, 0x89, 1, 1, 20), (8, 0x11D, 300, 7, 33),
                                    (8, 0x11D, 1, 1, 254), (8, 0x11D, 65535, 1, 8)])
def test_restatement_matches_reference_live(orc, ref, params):
    m, _, _, _, nr = params
    nn = (1 << m) - 1
    K = nn - nr
    ri, oi = ref.rs_info(params), orc.rs_info(params)
    for key in ri:
        assert np.all(ri[key] == oi[key])
    rng = np.random.default_rng(hash(params) & 0xFFFF)
    for size in sorted({K, max(1, K // 2), 1}):
        n = 400
        data = rng.integers(0, 256, (n, size), dtype=np.uint8)
        par = ref.rs_encode(data, params)
        assert (par == orc.rs_encode(data, params)).all()
        cw = np.concatenate([data & nn, par], 1)
        _inject(rng, cw, nn, nr // 2 + 4)
        for a, b in zip(ref.rs_decode(cw[:, :size], cw[:, size:], params),
                        orc.rs_decode(cw[:, :size], cw[:, size:], params)):
            assert (a == b).all()
        W = max(32, nr)
        pos = np.zeros((n, W), np.uint32)
        cnt = np.zeros(n, np.uint32)
        cw = np.concatenate([data & nn, par], 1)
        for i in range(n):
            c = int(rng.integers(0, min(nr, size) + 1))
            pp = rng.choice(size, c, replace=False)
            pos[i, :c] = np.sort(pp) if rng.random() < 0.6 else pp
            cnt[i] = c
            cw[i, pp] ^= rng.integers(0, nn + 1, c, dtype=np.uint8)
            v = int(rng.choice([0, 0, 1, 2]))
            vp = rng.choice(size + nr, v, replace=False)
            cw[i, vp] ^= rng.integers(1, nn + 1, v, dtype=np.uint8)
        for a, b in zip(ref.rs_decode(cw[:, :size], cw[:, size:], params, MODE_ERASURE, pos, cnt),
                        orc.rs_decode(cw[:, :size], cw[:, size:], params, MODE_ERASURE, pos, cnt)):
            assert (a == b).all()


def test_invalid_codes_agree_with_reference(orc, ref):
    for params in [(0, 0x11D, 1, 1, 16), (8, 0x11C, 1, 1, 16), (8, 0x11D, 1, 0, 16), (8, 0x11D, 1, 3, 16),
                   (8, 0x11D, 1, 5, 16), (4, 0x13, 1, 3, 4), (9, 0x211, 1, 1, 8)]:
        assert (ref.rs_info(params) is None) == (orc.rs_info(params) is None), params
    for m, poly, t in [(2, 7, 1), (17, 1, 1), (4, 0x13, 0), (4, 0x13, 17), (6, 0x43, 2), (8, 0x11D, 4)]:
        ri, oi = ref.bch_info(m, poly, t), orc.bch_info(m, poly, t)
        assert (ri is None) == (oi is None)
        if ri:
            assert (ri["n"], ri["k"], ri["parity_bits"]) == (oi["n"], oi["k"], oi["parity_bits"])


def test_gf_mod_matches_reference(orc, ref):
    for m, poly in [(8, 0x11D), (4, 0x13), (5, 0x25)]:
        for v in list(range(0, 1200)) + list(range(65000, 65536)):
            assert orc.gf_mod(m, poly, v) == ref.gf_mod(m, poly, v)
