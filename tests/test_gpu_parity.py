"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI (libpoporon.so -> libpoporon_cuda.so),
against the CPU oracle (oracle/liboracle.so, the restatement pinned in test_oracle.py) and the committed
golden fixtures.  Bar: bit-exact - corrected bytes, corrected_num and the success flag, including
uncorrectable codewords and the reference's erasure-mode quirks.

Nothing here reads /root/reference."""
import ctypes as C

import numpy as np
import pytest
from conftest import golden
from oraclelib import MODE_ERASURE, MODE_EXTSYN, MODE_PLAIN, RS_DEFAULT
from test_oracle import BCH_GOLDEN, RS_GOLDEN

import libpoporon_b200.api as api

pytestmark = pytest.mark.gpu


def _rs(params, **kw):
    m, poly, fcr, prim, nr = params
    return api.Poporon.rs(m, poly, fcr, prim, nr, **kw)


def _inject(rng, cw, nn, emax):
    n, L = cw.shape
    for i in range(n):
        e = int(rng.integers(0, min(L, emax) + 1))
        pos = rng.choice(L, e, replace=False)
        cw[i, pos] ^= rng.integers(1, nn + 1, e, dtype=np.uint8)


def _gpu_encode(h, msg, nr):
    n, size = msg.shape
    msg = np.ascontiguousarray(msg)
    par = np.full((n, nr), 0xEE, np.uint8)
    assert h.encode_batch(msg, size, size, par, nr, n)
    return par


def _gpu_decode(h, rx_d, rx_p, mode=MODE_PLAIN, pos=None, cnt=None, syn=None):
    d = np.array(rx_d, np.uint8, order="C", copy=True)
    p = np.array(rx_p, np.uint8, order="C", copy=True)
    n, size = d.shape
    ok = np.full(n, 7, np.uint8)
    cor = np.full(n, 0xFFFFFFFF, np.uint32)
    if mode == MODE_ERASURE:
        pos = np.ascontiguousarray(pos, np.uint32)
        cnt = np.ascontiguousarray(cnt, np.uint32)
        assert h.decode_batch_erasure(d, size, size, p, p.shape[1], n, pos, pos.shape[1], cnt, ok, cor)
    elif mode == MODE_EXTSYN:
        syn = np.ascontiguousarray(syn, np.uint16)
        assert h.decode_batch_syndrome(d, size, size, p, p.shape[1], n, syn, ok, cor)
    else:
        assert h.decode_batch(d, size, size, p, p.shape[1], n, ok, cor)
    if mode != MODE_EXTSYN:
        # the same batch as packed records (message immediately followed by parity): for the default code this
        # takes the bitsliced fast path, the split layout above takes the general-parameter kernels
        nr = p.shape[1]
        rec = np.ascontiguousarray(np.concatenate([np.asarray(rx_d, np.uint8), np.asarray(rx_p, np.uint8)], 1))
        ok2 = np.full(n, 7, np.uint8)
        cor2 = np.full(n, 0xFFFFFFFF, np.uint32)
        st = size + nr
        if mode == MODE_ERASURE:
            assert h.decode_batch_erasure(rec.ctypes.data, st, size, rec.ctypes.data + size, st, n, pos, pos.shape[1],
                                          cnt, ok2, cor2)
        else:
            assert h.decode_batch(rec.ctypes.data, st, size, rec.ctypes.data + size, st, n, ok2, cor2)
        assert (ok2 == ok).all() and (cor2 == cor).all(), "packed and split layouts disagree (status)"
        assert (rec[:, :size] == d).all() and (rec[:, size:] == p).all(), "packed and split layouts disagree (bytes)"
    return d, p, ok, cor


def _assert_same(got, want, tag=""):
    names = ("data", "parity", "ok", "corrected")
    for name, a, b in zip(names, got, want):
        if not (a == b).all():
            bad = np.nonzero((a != b).reshape(len(a), -1).any(1))[0]
            raise AssertionError(f"{tag}: {name} differs at codewords {bad[:8]} ({len(bad)} total)")


# ------------------------------------------------------------------------------------ golden fixtures
@pytest.mark.parametrize("name", RS_GOLDEN)
def test_rs_golden_fixture(gpu_device, name):
    g = golden(name)
    params = tuple(int(x) for x in g["params"])
    mode = int(g["mode"])
    with _rs(params) as h:
        assert (_gpu_encode(h, g["msg"], params[4]) == g["parity"]).all()
        kw = {}
        if mode == MODE_ERASURE:
            kw = dict(pos=g["pos"], cnt=g["cnt"])
        if mode == MODE_EXTSYN:
            kw = dict(syn=g["syn"])
        got = _gpu_decode(h, g["rx_data"], g["rx_parity"], mode, **kw)
        _assert_same(got, (g["out_data"], g["out_parity"], g["ok"], g["corrected"]), name)


@pytest.mark.parametrize("name", BCH_GOLDEN)
def test_bch_golden_fixture(gpu_device, name):
    g = golden(name)
    m, poly, t, n, k, p = (int(x) for x in g["params"])
    with api.Poporon.bch(m, poly, t) as h:
        assert h.info_size == (k + 7) // 8 and h.parity_size == (p + 7) // 8
        pb = (p + 7) // 8
        if "msg" in g:
            msg = np.ascontiguousarray(g["msg"])
            par = np.full((len(msg), pb), 0xEE, np.uint8)
            assert h.encode_batch(msg, msg.shape[1], msg.shape[1], par, pb, len(msg))
            assert (par == g["parity"]).all()
        d = np.array(g["rx_data"], np.uint8, order="C", copy=True)
        pp = np.ascontiguousarray(g["rx_parity"])
        N = len(d)
        ok = np.full(N, 7, np.uint8)
        cor = np.full(N, 0xFFFFFFFF, np.uint32)
        assert h.decode_batch(d, d.shape[1], d.shape[1], pp, pb, N, ok, cor)
        assert (ok == g["ok"]).all() and (cor == g["corrected"]).all() and (d == g["out_data"]).all()
        assert (pp == g["rx_parity"]).all()  # BCH decode never rewrites parity


# ------------------------------------------------------------------------------------ oracle differential
@pytest.mark.parametrize("size", [223, 222, 100, 64, 33, 2, 1])
def test_rs_default_random_vs_oracle(gpu_device, orc, size):
    rng = np.random.default_rng(1000 + size)
    n = 6000 if size == 223 else 1500
    msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        par = _gpu_encode(h, msg, 32)
        assert (par == orc.rs_encode(msg)).all()
        cw = np.concatenate([msg, par], 1)
        _inject(rng, cw, 255, 20)
        got = _gpu_decode(h, cw[:, :size], cw[:, size:])
        _assert_same(got, orc.rs_decode(cw[:, :size], cw[:, size:]), f"size {size}")
        assert h.launches > 0


def test_rs_default_error_count_sweep(gpu_device, orc):
    """exactly e errors for e = 0..20: corrected_num == e and data restored for e <= 16, failure above
    (reference tests/test_codec.c:170-233)"""
    rng = np.random.default_rng(7)
    per = 200
    msg = rng.integers(0, 256, (21 * per, 223), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        par = _gpu_encode(h, msg, 32)
        cw = np.concatenate([msg, par], 1)
        clean = cw.copy()
        for i in range(len(cw)):
            e = i // per
            pos = rng.choice(255, e, replace=False)
            cw[i, pos] ^= rng.integers(1, 256, e, dtype=np.uint8)
        d, p, ok, cor = got = _gpu_decode(h, cw[:, :223], cw[:, 223:])
        _assert_same(got, orc.rs_decode(cw[:, :223], cw[:, 223:]), "sweep")
        for e in range(21):
            sl = slice(e * per, (e + 1) * per)
            if e <= 16:
                assert ok[sl].all() and (cor[sl] == e).all()
                assert (np.concatenate([d[sl], p[sl]], 1) == clean[sl]).all()
            else:
                assert not ok[sl].any() and (cor[sl] == 0).all()
                assert (np.concatenate([d[sl], p[sl]], 1) == cw[sl]).all()


def test_rs_default_erasures_vs_oracle(gpu_device, orc):
    """sorted / unsorted rows, stale slots, zero magnitudes, extra unknown errors, 2v + c > 32"""
    rng = np.random.default_rng(11)
    n, size = 6000, 223
    msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        par = _gpu_encode(h, msg, 32)
        cw = np.concatenate([msg, par], 1)
        pos = np.zeros((n, 32), np.uint32)
        cnt = np.zeros(n, np.uint32)
        for i in range(n):
            c = int(rng.integers(0, 33))
            pp = rng.choice(size, c, replace=False)
            pos[i, :c] = np.sort(pp) if rng.random() < 0.7 else pp
            cnt[i] = c
            if rng.random() < 0.25:
                pos[i, c:] = rng.integers(0, size, 32 - c)
            cw[i, pp] ^= rng.integers(0, 256, c, dtype=np.uint8)
            v = int(rng.choice([0, 0, 0, 0, 1, 2, 3]))
            vp = rng.choice(255, v, replace=False)
            cw[i, vp] ^= rng.integers(1, 256, v, dtype=np.uint8)
        got = _gpu_decode(h, cw[:, :size], cw[:, size:], MODE_ERASURE, pos, cnt)
        want = orc.rs_decode(cw[:, :size], cw[:, size:], RS_DEFAULT, MODE_ERASURE, pos, cnt)
        _assert_same(got, want, "erasure")
        assert (want[2] == 0).any()
        # the "false with corrected_num != 0" corner of SURVEY A.5: 28 sorted erasures + 3 unknown errors
        m2 = 4000
        cw2 = np.concatenate([msg[:m2], par[:m2]], 1)
        pos2 = np.zeros((m2, 32), np.uint32)
        for i in range(m2):
            pp = np.sort(rng.choice(size, 28, replace=False))
            pos2[i, :28] = pp
            cw2[i, pp] ^= rng.integers(0, 256, 28, dtype=np.uint8)
            vp = rng.choice(255, 3, replace=False)
            cw2[i, vp] ^= rng.integers(1, 256, 3, dtype=np.uint8)
        cnt2 = np.full(m2, 28, np.uint32)
        got2 = _gpu_decode(h, cw2[:, :size], cw2[:, size:], MODE_ERASURE, pos2, cnt2)
        want2 = orc.rs_decode(cw2[:, :size], cw2[:, size:], RS_DEFAULT, MODE_ERASURE, pos2, cnt2)
        _assert_same(got2, want2, "erasure 28+3")
        assert ((want2[2] == 0) & (want2[3] != 0)).any()


def test_rs_default_extsyn_vs_oracle(gpu_device, orc):
    rng = np.random.default_rng(13)
    n, size = 9000, 223
    msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        par = _gpu_encode(h, msg, 32)
        syn = np.full((n, 32), 255, np.uint16)
        for i in range(n):
            if rng.random() < 0.7:
                q = int(rng.integers(1, 33))
                syn[i, rng.choice(32, q, replace=False)] = rng.integers(0, 255, q)
        # a block of genuine single-error syndrome sets: error magnitude alpha^a at location L gives
        # S_i = a + (1+i)*(254-L) in index form
        for i in range(500):
            L, a = int(rng.integers(0, 255)), int(rng.integers(0, 255))
            syn[i] = (a + (1 + np.arange(32)) * (254 - L)) % 255
        assert h.profile(True)
        got = _gpu_decode(h, msg, par, MODE_EXTSYN, syn=syn)
        prof = h.profile_read()
        # caller-supplied syndromes run on the bitsliced solver, not on the per-thread kernels
        assert prof.get("rs255_solve_kernel", (0, 0))[0] >= 1 and "rs_generic kernels" not in prof, prof
        _assert_same(got, orc.rs_decode(msg, par, RS_DEFAULT, MODE_EXTSYN, syn=syn), "extsyn")
        assert got[2][:500].all() and (got[3][:500] == 1).all()


def _np_syndromes(cw, m, poly, fcr, prim, nr):
    """index-form syndromes of full-length words cw [n, nn] by Horner in numpy: S_i = r(alpha^((fcr + i) * prim)),
    r_0 first = highest power (reference src/decode.c:232-416); nn = zero"""
    nn = (1 << m) - 1
    exp = np.zeros(2 * nn + 2, np.int64)
    log = np.full(nn + 1, nn, np.int64)
    x = 1
    for i in range(nn):
        exp[i] = x
        log[x] = i
        x <<= 1
        if x > nn:
            x ^= poly
    exp[nn:2 * nn] = exp[:nn]
    out = np.zeros((cw.shape[0], nr), np.uint16)
    for i in range(nr):
        root = ((fcr + i) * prim) % nn
        acc = np.zeros(cw.shape[0], np.int64)
        for j in range(cw.shape[1]):
            nz = acc != 0
            acc = np.where(nz, exp[(log[acc] + root) % nn], 0) ^ cw[:, j]
        out[:, i] = log[acc]
    return out


def test_syndromes_only_output(gpu_device, orc):
    """the 32 index-form syndromes as calculate_syndrome_u8 leaves them, handed to the caller; the
    array round-trips through the external-syndrome decode with the outcome of a plain decode."""
    rng = np.random.default_rng(31)
    n = 40000
    msg = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        par = _gpu_encode(h, msg, 32)
        cw = np.concatenate([msg, par], 1)
        # SURVEY A.7 known answer: one error of magnitude 1 on the first symbol -> 254, 253, ..., 223
        cw[0, 0] ^= 1
        _inject(rng, cw[1:], 255, 18)
        cw[5] = np.concatenate([msg[5], par[5]])  # a clean word: all syndromes 255
        syn = np.full((n, 32), 0xBEEF, np.uint16)
        d, p = np.ascontiguousarray(cw[:, :223]), np.ascontiguousarray(cw[:, 223:])
        assert h.syndromes_batch(d, 223, 223, p, 32, n, syn)
        assert (d == cw[:, :223]).all() and (p == cw[:, 223:]).all()  # nothing corrected
        assert syn[0].tolist() == list(range(254, 222, -1)) and (syn[5] == 255).all()
        assert (syn[:1500] == _np_syndromes(cw[:1500], 8, 0x11D, 1, 1, 32)).all()
        # packed records give the same array
        syn2 = np.zeros((n, 32), np.uint16)
        rec = np.ascontiguousarray(cw)
        assert h.syndromes_batch(rec.ctypes.data, 255, 223, rec.ctypes.data + 223, 255, n, syn2)
        assert (syn2 == syn).all()
        # round trip: decoding from these syndromes = plain decode (reference src/decode.c:446-464 against :475-477)
        want = orc.rs_decode(cw[:, :223], cw[:, 223:])
        _assert_same(_gpu_decode(h, cw[:, :223], cw[:, 223:], MODE_EXTSYN, syn=syn), want, "syndromes round trip")
        # shortened words
        size = 100
        par_s = _gpu_encode(h, msg[:2000, :size], 32)
        cws = np.concatenate([msg[:2000, :size], par_s], 1)
        _inject(rng, cws, 255, 10)
        syn_s = np.zeros((2000, 32), np.uint16)
        assert h.syndromes_batch(np.ascontiguousarray(cws[:, :size]), size, size, np.ascontiguousarray(cws[:, size:]), 32,
                                 2000, syn_s)
        full = np.concatenate([np.zeros((2000, 223 - size), np.uint8), cws], 1)
        assert (syn_s == _np_syndromes(full, 8, 0x11D, 1, 1, 32)).all()
        # sizes outside the padding rule are refused
        assert not h.syndromes_batch(d, 224, 224, p, 32, 4, syn) and not h.syndromes_batch(d, 0, 0, p, 32, 4, syn)


@pytest.mark.parametrize("params", [(8, 0x11D, 1, 1, 16), (8, 0x187, 112, 11, 32), (4, 0x13, 1, 2, 8)])
def test_syndromes_only_general_parameters(gpu_device, orc, params):
    m, poly, fcr, prim, nr = params
    nn = (1 << m) - 1
    K = nn - nr
    rng = np.random.default_rng(77)
    n = 600
    with _rs(params) as h:
        msg = rng.integers(0, nn + 1, (n, K), dtype=np.uint8)
        par = _gpu_encode(h, msg, nr)
        cw = np.concatenate([msg, par], 1)
        _inject(rng, cw[1:], nn, nr // 2 + 1)
        syn = np.zeros((n, nr), np.uint16)
        assert h.syndromes_batch(np.ascontiguousarray(cw[:, :K]), K, K, np.ascontiguousarray(cw[:, K:]), nr, n, syn)
        assert (syn[0] == nn).all()
        assert (syn == _np_syndromes(cw, m, poly, fcr, prim, nr)).all()
        want = orc.rs_decode(cw[:, :K], cw[:, K:], params)
        _assert_same(_gpu_decode(h, cw[:, :K], cw[:, K:], MODE_EXTSYN, syn=syn), want, f"{params} syndromes round trip")


@pytest.mark.parametrize("params", [(8, 0x187, 112, 11, 32), (8, 0x11D, 0, 1, 16), (8, 0x11D, 1, 2, 16),
                                    (8, 0x11D, 1, 1, 127), (8, 0x11D, 2, 1, 4), (8, 0x11D, 1, 1, 10),
                                    (4, 0x13, 1, 2, 8), (5, 0x25, 3, 3, 6), (7, 0x89, 1, 1, 20),
                                    (8, 0x11D, 300, 7, 33), (8, 0x11D, 1, 1, 254), (3, 0xB, 1, 1, 2),
                                    (8, 0x11D, 65535, 1, 8)])
def test_rs_general_parameters_vs_oracle(gpu_device, orc, params):
    m, _, _, _, nr = params
    nn = (1 << m) - 1
    K = nn - nr
    rng = np.random.default_rng(hash(params) & 0xFFFF)
    with _rs(params) as h:
        assert h.parity_size == nr and h.info_size == K and h.fec_type == api.FEC_RS
        for size in sorted({K, max(1, K // 2), 1}):
            n = 500
            msg = rng.integers(0, 256, (n, size), dtype=np.uint8)  # bits above m are masked by the codec
            par = _gpu_encode(h, msg, nr)
            assert (par == orc.rs_encode(msg, params)).all()
            cw = np.concatenate([msg & nn, par], 1)
            _inject(rng, cw, nn, nr // 2 + 3)
            _assert_same(_gpu_decode(h, cw[:, :size], cw[:, size:]),
                         orc.rs_decode(cw[:, :size], cw[:, size:], params), f"{params} size {size}")
            W = max(32, nr)
            pos = np.zeros((n, W), np.uint32)
            cnt = np.zeros(n, np.uint32)
            cw = np.concatenate([msg & nn, par], 1)
            for i in range(n):
                c = int(rng.integers(0, min(nr, size) + 1))
                pp = rng.choice(size, c, replace=False)
                pos[i, :c] = np.sort(pp) if rng.random() < 0.6 else pp
                cnt[i] = c
                cw[i, pp] ^= rng.integers(0, nn + 1, c, dtype=np.uint8)
                v = int(rng.choice([0, 0, 1, 2]))
                vp = rng.choice(size + nr, v, replace=False)
                cw[i, vp] ^= rng.integers(1, nn + 1, v, dtype=np.uint8)
            _assert_same(_gpu_decode(h, cw[:, :size], cw[:, size:], MODE_ERASURE, pos, cnt),
                         orc.rs_decode(cw[:, :size], cw[:, size:], params, MODE_ERASURE, pos, cnt),
                         f"{params} erasure size {size}")


def test_rs_bad_sizes(gpu_device, orc):
    with api.Poporon.rs_default() as h:
        for size in (224, 230, 255, 300):
            msg = np.arange(4 * size, dtype=np.uint32).astype(np.uint8).reshape(4, size)
            par = _gpu_encode(h, msg, 32)  # encode does not validate the size
            assert (par == orc.rs_encode(msg)).all()
            d, p, ok, cor = _gpu_decode(h, msg, par)
            assert not ok.any() and not cor.any() and (d == msg).all() and (p == par).all()


def test_packed_records_and_gaps_roundtrip(gpu_device, orc):
    """packed 255-byte records (parity = data + size) and a strided layout with gaps that must survive"""
    rng = np.random.default_rng(21)
    n = 3000
    msg = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    want_par = orc.rs_encode(msg)
    with api.Poporon.rs_default() as h:
        rec = np.zeros((n, 255), np.uint8)
        rec[:, :223] = msg
        assert h.encode_batch(rec.ctypes.data, 255, 223, rec.ctypes.data + 223, 255, n)
        assert (rec[:, 223:] == want_par).all() and (rec[:, :223] == msg).all()
        _inject(rng, rec, 255, 18)
        want = orc.rs_decode(rec[:, :223], rec[:, 223:])
        ok = np.zeros(n, np.uint8)
        cor = np.zeros(n, np.uint32)
        assert h.decode_batch(rec.ctypes.data, 255, 223, rec.ctypes.data + 223, 255, n, ok, cor)
        _assert_same((rec[:, :223], rec[:, 223:], ok, cor), want, "packed")
        # strided with gaps: data stride 300, parity stride 40, gaps hold a canary
        dbuf = np.full((n, 300), 0xC3, np.uint8)
        pbuf = np.full((n, 40), 0x3C, np.uint8)
        dbuf[:, :223] = msg
        assert h.encode_batch(dbuf, 300, 223, pbuf, 40, n)
        assert (pbuf[:, :32] == want_par).all() and (pbuf[:, 32:] == 0x3C).all() and (dbuf[:, 223:] == 0xC3).all()
        # parity rows far apart at odd addresses (the encoder merges whole sectors there): canaries in the gaps,
        # before the first and after the last row
        for pstride, off in ((100, 1), (63, 2), (97, 3), (64, 0)):
            raw = np.full(n * pstride + 128, 0x5A, np.uint8)
            base = raw.ctypes.data + 64 + off
            assert h.encode_batch(dbuf, 300, 223, base, pstride, n)
            rows = np.lib.stride_tricks.as_strided(raw[64 + off:], (n, pstride), (pstride, 1))
            assert (rows[:, :32] == want_par).all(), (pstride, off)
            assert (rows[:-1, 32:] == 0x5A).all() and (raw[:64 + off] == 0x5A).all()
            assert (raw[64 + off + (n - 1) * pstride + 32:] == 0x5A).all()


def _fast_path_counts(h):
    """launches per kernel family since profile(True): (screen, generic)"""
    prof = h.profile_read()
    return prof.get("rs255_fft_kernel<SCREEN>", (0, 0.0))[0], prof.get("rs_generic kernels", (0, 0.0))[0]


@pytest.mark.parametrize("layout", ["packed", "split", "gapped", "oddbase", "oddbase_split"])
def test_every_layout_takes_the_bitsliced_path(gpu_device, orc, layout):
    """the two-pointer shape of the reference API (include/poporon.h:90-91), packed records,
    gapped strides and misaligned bases all run on the transform kernels (no fall-back to the per-thread generic
    kernels), each bit-compared with the oracle."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(77)
    n = 70000
    msg = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    par = orc.rs_encode(msg)
    cw = np.concatenate([msg, par], 1)
    _inject(rng, cw, 255, 18)
    want = orc.rs_decode(cw[:, :223], cw[:, 223:])
    dstride, pstride, doff, poff = {"packed": (255, 255, 0, 223), "split": (223, 32, 0, None),
                                    "gapped": (260, 260, 0, 225), "oddbase": (255, 255, 3, 226),
                                    "oddbase_split": (223, 32, 5, None)}[layout]
    dev = torch.device("cuda:0")
    dbuf = torch.full((n * dstride + 64,), 0xC3, dtype=torch.uint8, device=dev)
    if poff is None:
        pbuf = torch.full((n * pstride + 64,), 0x3C, dtype=torch.uint8, device=dev)
        pbase, pview, po = pbuf.data_ptr() + 7, pbuf, 7
    else:
        pbase, pview, po = dbuf.data_ptr() + poff, dbuf, poff

    def put(buf, off, stride, rows):
        v = torch.as_strided(buf, (n, rows.shape[1]), (stride, 1), off)
        v.copy_(torch.from_numpy(np.ascontiguousarray(rows)).to(dev))
        return v

    with api.Poporon.rs_default() as h:
        # (profiling attaches to the devices the handle has used: touch device 0 first)
        assert h.encode_batch(np.zeros((1, 223), np.uint8), 223, 223, np.zeros((1, 32), np.uint8), 32, 1)
        # encode on the same layout
        dv = put(dbuf, doff, dstride, msg)
        assert h.profile(True)
        assert h.encode_batch(dbuf.data_ptr() + doff, dstride, 223, pbase, pstride, n)
        prof = h.profile_read()
        assert prof.get("rs255_fft_kernel<ENCODE>", (0, 0))[0] >= 1 and prof.get("rs_generic kernels", (0, 0))[0] == 0
        pv = torch.as_strided(pview, (n, 32), (pstride, 1), po)
        assert (pv.cpu().numpy() == par).all(), f"{layout}: encode parity"
        assert (dv.cpu().numpy() == msg).all()
        # decode
        put(dbuf, doff, dstride, cw[:, :223])
        put(pview, po, pstride, cw[:, 223:])
        ok = torch.full((n,), 9, dtype=torch.uint8, device=dev)
        cor = torch.full((n,), -1, dtype=torch.int32, device=dev)
        assert h.profile(True)
        assert h.decode_batch(dbuf.data_ptr() + doff, dstride, 223, pbase, pstride, n, ok.data_ptr(), cor.data_ptr())
        screen, generic = _fast_path_counts(h)
        assert screen >= 1 and generic == 0, f"{layout}: fell back to the generic kernels"
        got = (dv.cpu().numpy(), pv.cpu().numpy(), ok.cpu().numpy(), cor.cpu().numpy().astype(np.uint32))
        _assert_same(got, want, layout)
        # bytes between the rows are untouched
        if layout == "gapped":
            gap = torch.as_strided(dbuf, (n - 1, 3), (dstride, 1), 257)
            assert (gap == 0xC3).all()


@pytest.mark.parametrize("nr", [16, 8])
def test_rs255_239_and_247_on_the_bitsliced_path(gpu_device, orc, nr):
    """The reference's other documented codes RS(255,239) and RS(255,247) (README.md:417-419, tests/test_rs.c:36-83):
    same field, fcr and prim as the default code, so they run on the transform kernels + the bitsliced solver (with 16 /
    8 syndromes) instead of the per-thread generic kernels.  Every mode against the oracle."""
    params = (8, 0x11D, 1, 1, nr)
    K = 255 - nr
    rng = np.random.default_rng(900 + nr)
    n = 50000
    msg = rng.integers(0, 256, (n, K), dtype=np.uint8)
    want_par = orc.rs_encode(msg, params, threads=8)
    with _rs(params) as h:
        assert h.encode_batch(np.zeros((1, K), np.uint8), K, K, np.zeros((1, nr), np.uint8), nr, 1)  # touch the device
        assert h.profile(True)
        # encode: split arrays and packed records
        par = np.full((n, nr), 0xEE, np.uint8)
        assert h.encode_batch(msg, K, K, par, nr, n)
        assert (par == want_par).all()
        rec = np.zeros((n, 255), np.uint8)
        rec[:, :K] = msg
        assert h.encode_batch(rec.ctypes.data, 255, K, rec.ctypes.data + K, 255, n)
        assert (rec[:, K:] == want_par).all() and (rec[:, :K] == msg).all()
        prof = h.profile_read()
        assert prof.get("rs255_fft_kernel<ENCODE>", (0, 0))[0] >= 2 and "rs_generic kernels" not in prof, prof
        # decode, 0 .. t + 2 errors
        cw = np.concatenate([msg, want_par], 1)
        _inject(rng, cw, 255, nr // 2 + 2)
        want = orc.rs_decode(cw[:, :K], cw[:, K:], params, threads=8)
        assert 0 < want[2].mean() < 1  # decodable and undecodable words both present
        assert h.profile(True)
        got = _gpu_decode(h, cw[:, :K], cw[:, K:])  # split and packed, compared with each other inside
        prof = h.profile_read()
        assert prof.get("rs255_solve_kernel", (0, 0))[0] >= 2 and "rs_generic kernels" not in prof, prof
        _assert_same(got, want, f"RS(255,{K}) plain")
        # syndromes only, and the round trip through the external-syndrome decode
        syn = np.zeros((n, nr), np.uint16)
        assert h.syndromes_batch(np.ascontiguousarray(cw[:, :K]), K, K, np.ascontiguousarray(cw[:, K:]), nr, n, syn)
        assert (syn[:1000] == _np_syndromes(cw[:1000], 8, 0x11D, 1, 1, nr)).all()
        _assert_same(_gpu_decode(h, cw[:, :K], cw[:, K:], MODE_EXTSYN, syn=syn), want, f"RS(255,{K}) extsyn")
        # shortened words
        size = 57
        par_s = _gpu_encode(h, msg[:8000, :size], nr)
        assert (par_s == orc.rs_encode(msg[:8000, :size], params)).all()
        cws = np.concatenate([msg[:8000, :size], par_s], 1)
        _inject(rng, cws, 255, nr // 2 + 1)
        _assert_same(_gpu_decode(h, cws[:, :size], cws[:, size:]), orc.rs_decode(cws[:, :size], cws[:, size:], params),
                     f"RS(255,{K}) size {size}")
        # erasure mode (general stage behind the transform screen)
        m = 6000
        cwe = np.concatenate([msg[:m], want_par[:m]], 1)
        pos = np.zeros((m, nr), np.uint32)
        cnt = rng.integers(0, nr + 1, m).astype(np.uint32)
        for i in range(m):
            c = int(cnt[i])
            pp = np.sort(rng.choice(K, c, replace=False))
            pos[i, :c] = pp
            cwe[i, pp] ^= rng.integers(0, 256, c, dtype=np.uint8)
            if rng.random() < 0.2:
                cwe[i, rng.integers(0, 255)] ^= rng.integers(1, 256)
        _assert_same(_gpu_decode(h, cwe[:, :K], cwe[:, K:], MODE_ERASURE, pos, cnt),
                     orc.rs_decode(cwe[:, :K], cwe[:, K:], params, MODE_ERASURE, pos, cnt), f"RS(255,{K}) erasure")


def test_device_pointer_path_matches_host_path(gpu_device, orc):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(23)
    n = 5000
    msg = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        rec = torch.zeros((n, 255), dtype=torch.uint8, device="cuda:0")
        rec[:, :223] = torch.from_numpy(msg).cuda()
        stream = torch.cuda.current_stream().cuda_stream
        assert h.encode_batch_async(rec.data_ptr(), 255, 223, rec.data_ptr() + 223, 255, n, stream=stream)
        torch.cuda.synchronize()
        cw = rec.cpu().numpy()
        assert (cw[:, 223:] == orc.rs_encode(msg)).all()
        _inject(rng, cw, 255, 19)
        rec.copy_(torch.from_numpy(cw))
        ok = torch.full((n,), 9, dtype=torch.uint8, device="cuda:0")
        cor = torch.full((n,), -1, dtype=torch.int32, device="cuda:0")
        assert h.decode_batch_async(rec.data_ptr(), 255, 223, rec.data_ptr() + 223, 255, n, ok.data_ptr(),
                                    cor.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        out = rec.cpu().numpy()
        want = orc.rs_decode(cw[:, :223], cw[:, 223:])
        _assert_same((out[:, :223], out[:, 223:], ok.cpu().numpy(), cor.cpu().numpy().astype(np.uint32)), want, "dev")
        # synchronous device-pointer form
        rec.copy_(torch.from_numpy(cw))
        assert h.decode_batch(rec.data_ptr(), 255, 223, rec.data_ptr() + 223, 255, n, ok.data_ptr(), cor.data_ptr())
        assert (rec.cpu().numpy() == out).all()


# ------------------------------------------------------------------------------------ single-instance API
def test_config1_hello_reed_solomon(gpu_device):
    """BASELINE config 1 / SURVEY B.2 through poporon_encode / poporon_decode"""
    with api.Poporon.rs_default() as h:
        assert h.fec_type == api.FEC_RS and h.parity_size == 32 and h.info_size == 223
        data = np.zeros(64, np.uint8)
        msg = b"Hello, Reed-Solomon!"
        data[:len(msg)] = np.frombuffer(msg, np.uint8)
        parity = np.zeros(32, np.uint8)
        assert h.encode(data, parity)
        assert parity.tobytes().hex() == "31e6a711c1fc30242fd40fa62307b796856794fff72ac6a79c6fb63e70086d5b"
        rx = data.copy()
        rx[0] ^= 0xFF
        rx[10] ^= 0xAA
        ok, cnt = h.decode(rx, parity)
        assert ok and cnt == 2 and (rx == data).all()
        ok, cnt = h.decode(rx, parity)
        assert ok and cnt == 0
        ok, _ = h.decode(rx, parity, want_count=False)  # corrected_num may be NULL
        assert ok


def test_single_api_null_and_size_rules(gpu_device):
    """reference tests/test_codec.c:72-75,223-226 and SURVEY A.5 outcome table"""
    with api.Poporon.rs_default() as h:
        d = np.zeros(64, np.uint8)
        p = np.zeros(32, np.uint8)
        host = api.host
        assert not host.poporon_encode(None, d.ctypes.data, 64, p.ctypes.data)
        assert not host.poporon_encode(h.ptr, None, 64, p.ctypes.data)
        assert not host.poporon_encode(h.ptr, d.ctypes.data, 64, None)
        assert h.encode(d, p)
        cnt = C.c_size_t(12345)
        assert not host.poporon_decode(None, d.ctypes.data, 64, p.ctypes.data, C.byref(cnt))
        assert not host.poporon_decode(h.ptr, None, 64, p.ctypes.data, C.byref(cnt))
        assert not host.poporon_decode(h.ptr, d.ctypes.data, 64, None, C.byref(cnt))
        assert not host.poporon_decode(h.ptr, d.ctypes.data, 0, p.ctypes.data, C.byref(cnt))
        assert cnt.value == 12345  # untouched on NULL / size 0
        big = np.zeros(224, np.uint8)
        assert not host.poporon_decode(h.ptr, big.ctypes.data, 224, p.ctypes.data, C.byref(cnt))
        assert cnt.value == 0  # bad size: false, corrected_num = 0
        # 17 errors: failure, corrected_num = 0, buffers untouched
        rng = np.random.default_rng(5)
        d = rng.integers(0, 256, 223, dtype=np.uint8)
        assert h.encode(d, p)
        rx = d.copy()
        rx[rng.choice(223, 17, replace=False)] ^= 0xFF
        keep = rx.copy()
        ok, c = h.decode(rx, p)
        assert not ok and c == 0 and (rx == keep).all()


@pytest.mark.parametrize("size", [223, 100, 32, 5])
def test_erasure_kernel_edges(gpu_device, orc, size):
    """rs255_erasure_kernel (erasure-only rows of erasure mode): counts 1 and 32, first / last message position,
    shortened messages, zero magnitudes, position rows wider than 32 slots and at addresses that do not allow
    quad-word loads, a batch that is not a multiple of the 1024-codeword tile; with duplicate, unsorted, empty
    and undeclared-error rows mixed in (those go to the general stage)."""
    rng = np.random.default_rng(1000 + size)
    with api.Poporon.rs_default() as h:
        # (the host pipeline puts the position rows behind n status bytes and n counts on the device: they are
        # 16-byte aligned - the kernel's quad-word loads - for n = 3148 with strides 32 and 40, and not otherwise)
        for n, stride, shift in ((3148, 32, 0), (3148, 40, 0), (3149, 32, 0), (3148, 33, 1)):
            msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
            par = _gpu_encode(h, msg, 32)
            cw = np.concatenate([msg, par], 1)
            flat = np.zeros(n * stride + 4, np.uint32)
            pos = flat[shift:shift + n * stride].reshape(n, stride)  # (shift = 1: rows at 4 mod 16)
            cnt = np.zeros(n, np.uint32)
            for i in range(n):
                kind = i % 16
                c = (1, min(32, size), int(rng.integers(1, min(32, size) + 1)))[i % 3]
                pp = np.sort(rng.choice(size, c, replace=False))
                if kind == 1:
                    pp[0] = 0
                    pp = np.unique(pp)
                if kind == 2:
                    pp[-1] = size - 1
                    pp = np.unique(pp)
                c = len(pp)
                pos[i, :c], cnt[i] = pp, c
                # slots beyond the count: anything INSIDE the message (the reference - and so the oracle - writes through
                # them unchecked when a row has undeclared errors, src/decode.c:211-214)
                pos[i, c:] = rng.integers(0, size, stride - c)
                cw[i, pp] ^= rng.integers(0, 256, c, dtype=np.uint8)
                if kind == 3 and c >= 2:  # unsorted
                    pos[i, [0, c - 1]] = pos[i, [c - 1, 0]]
                elif kind == 4 and c >= 2:  # duplicate
                    pos[i, 1] = pos[i, 0]
                elif kind == 6:  # undeclared error in the parity
                    cw[i, size + int(rng.integers(0, 32))] ^= 0x80
                elif kind == 7:  # undeclared error in the message
                    cw[i, int(rng.integers(0, size))] ^= 0x01
                elif kind == 8:  # nothing declared
                    cnt[i] = 0
            d = np.array(cw[:, :size], order="C")
            p = np.array(cw[:, size:], order="C")
            ok = np.full(n, 7, np.uint8)
            cor = np.full(n, 0xFFFFFFFF, np.uint32)
            assert h.decode_batch_erasure(d, size, size, p, 32, n, pos.ctypes.data, stride, cnt, ok, cor)
            want = orc.rs_decode(cw[:, :size], cw[:, size:], RS_DEFAULT, MODE_ERASURE,
                                 np.ascontiguousarray(pos), cnt)
            _assert_same((d, p, ok, cor), want, f"erasure kernel size={size} n={n} stride={stride} shift={shift}")


@pytest.mark.parametrize("nr", [16, 8])
def test_erasure_mode_of_the_shorter_codes(gpu_device, orc, nr):
    """RS(255,239) / RS(255,247) in erasure mode: rs255_erasure_kernel with 16 / 8 roots (position rows of 32 slots),
    the older path for rows of nr slots only; undeclared errors, unsorted and over-long rows mixed in"""
    params = (8, 0x11D, 1, 1, nr)
    rng = np.random.default_rng(500 + nr)
    with _rs(params) as h:
        for size, stride in ((255 - nr, 32), (60, 32), (255 - nr, nr)):
            n = 4000
            msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
            par = _gpu_encode(h, msg, nr)
            cw = np.concatenate([msg, par], 1)
            pos = np.zeros((n, stride), np.uint32)
            cnt = np.zeros(n, np.uint32)
            for i in range(n):
                c = int(rng.integers(0, nr + 1))
                pp = np.sort(rng.choice(size, c, replace=False))
                pos[i, :c], cnt[i] = pp, c
                pos[i, c:] = rng.integers(0, size, stride - c)
                cw[i, pp] ^= rng.integers(0, 256, c, dtype=np.uint8)
                kind = i % 10
                if kind == 1:
                    cw[i, int(rng.integers(0, size + nr))] ^= 0x11
                elif kind == 2 and c >= 2:
                    pos[i, [0, c - 1]] = pos[i, [c - 1, 0]]
                elif kind == 3 and stride > nr:
                    cnt[i] = nr + 1
            got = _gpu_decode(h, cw[:, :size], cw[:, size:], MODE_ERASURE, pos, cnt)
            want = orc.rs_decode(cw[:, :size], cw[:, size:], params, MODE_ERASURE, pos, cnt)
            _assert_same(got, want, f"erasure nr={nr} size={size} stride={stride}")


@pytest.mark.parametrize("size,F", [(223, 1), (223, 10), (223, 23), (223, 30), (60, 7), (9, 4)])
def test_erasure_rows_with_undeclared_errors(gpu_device, orc, monkeypatch, size, F):
    """rs255_ee_kernel: every row declares the same number F of positions (so that its warps accept their tiles) and has
    errors outside them - few enough to decode, one too many, or none; the slots beyond F hold stale positions, through
    which the reference applies the magnitudes of the roots it finds beyond the F-th (src/decode.c:211-214)."""
    rng = np.random.default_rng(7000 + 31 * size + F)
    n = 3 * 1024 + 500
    monkeypatch.setenv("PP_RS255_EE_MIN_ROWS", "0")  # (read when the handle is created: small batches skip the kernel)
    with api.Poporon.rs_default() as h:
        msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
        par = _gpu_encode(h, msg, 32)
        cw = np.concatenate([msg, par], 1)
        pos = np.zeros((n, 32), np.uint32)
        cnt = np.full(n, F, np.uint32)
        emax = (32 - F) // 2
        for i in range(n):
            pp = np.sort(rng.choice(size, F, replace=False))
            pos[i, :F] = pp
            pos[i, F:] = rng.integers(0, size, 32 - F)
            cw[i, pp] ^= rng.integers(0, 256, F, dtype=np.uint8)
            e = int(rng.integers(0, emax + 2))
            others = np.setdiff1d(np.arange(size + 32), pp)
            q = rng.choice(others, min(e, len(others)), replace=False)
            cw[i, q] ^= rng.integers(1, 256, len(q), dtype=np.uint8)
        got = _gpu_decode(h, cw[:, :size], cw[:, size:], MODE_ERASURE, pos, cnt)
        want = orc.rs_decode(cw[:, :size], cw[:, size:], RS_DEFAULT, MODE_ERASURE, pos, cnt)
        _assert_same(got, want, f"erasure + errors size={size} F={F}")
        assert (want[2] == 1).sum() > n // 4


def test_modes_alternate_on_one_handle(gpu_device, orc):
    """plain -> erasure -> plain -> erasure with growing batches on ONE handle: the decode scratch is sized by mode and
    by batch (the erasure kernel's is larger) and regrown in between"""
    rng = np.random.default_rng(77)
    size = 223
    with api.Poporon.rs_default() as h:
        for step, n in enumerate((3000, 5000, 9000, 4000)):
            msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
            par = _gpu_encode(h, msg, 32)
            cw = np.concatenate([msg, par], 1)
            if step % 2 == 0:
                _inject(rng, cw, 255, 18)
                got = _gpu_decode(h, cw[:, :size], cw[:, size:])
                want = orc.rs_decode(cw[:, :size], cw[:, size:])
            else:
                pos = np.zeros((n, 32), np.uint32)
                cnt = rng.integers(1, 33, n).astype(np.uint32)
                pick = np.sort(np.argsort(rng.random((n, size)), axis=1)[:, :32], axis=1)
                act = np.arange(32)[None, :] < cnt[:, None]
                pos[act] = pick[act]
                rows = np.repeat(np.arange(n), 32).reshape(n, 32)
                cw[rows[act], pick[act]] ^= rng.integers(0, 256, int(act.sum()), dtype=np.uint8)
                cw[::9, 250] ^= 0x33  # an undeclared error on every ninth row
                got = _gpu_decode(h, cw[:, :size], cw[:, size:], MODE_ERASURE, pos, cnt)
                want = orc.rs_decode(cw[:, :size], cw[:, size:], RS_DEFAULT, MODE_ERASURE, pos, cnt)
            _assert_same(got, want, f"step {step}")


def test_single_api_erasure_object(gpu_device, orc):
    """reference tests/test_codec.c:123-168, tests/test_unified.c:82-112 and SURVEY B.6"""
    er = api.Erasure(32, 16)
    with api.Poporon.rs(erasure=er) as h:
        rng = np.random.default_rng(9)
        data = rng.integers(0, 256, 64, dtype=np.uint8)
        parity = np.zeros(32, np.uint8)
        assert h.encode(data, parity)
        for count in (16, 20):
            er.reset()
            rx = data.copy()
            for pos in rng.choice(64, count, replace=False):  # unsorted, constant magnitude like tests/util.h
                rx[pos] ^= 0xFF
                er.add(int(pos))
            ok, c = h.decode(rx, parity)
            assert ok and c == count and (rx == data).all()
        # B.6 quirk: unsorted positions with distinct magnitudes
        d3 = ((7 * np.arange(223) + 3) & 0xFF).astype(np.uint8)
        p3 = np.zeros(32, np.uint8)
        assert h.encode(d3, p3)
        rx = d3.copy()
        rx[200] ^= 1
        rx[100] ^= 2
        rx[3] ^= 3
        er.reset()
        for pos in (200, 100, 3):
            er.add(pos)
        ok, c = h.decode(rx, p3)
        assert ok and c == 3
        assert rx[3] ^ d3[3] == 2 and rx[100] ^ d3[100] == 0 and rx[200] ^ d3[200] == 2
        # empty list + one unknown error: applied at slot 0's stale position (here 200)
        rx = d3.copy()
        rx[50] ^= 0x21
        er.reset()
        ok, c = h.decode(rx, p3)
        pos = np.zeros((1, 32), np.uint32)
        pos[0, :3] = [200, 100, 3]
        bad = d3.copy()
        bad[50] ^= 0x21
        wd, wp, wok, wc = orc.rs_decode(bad[None], p3[None], RS_DEFAULT, MODE_ERASURE, pos, np.zeros(1, np.uint32))
        assert ok == bool(wok[0]) and c == wc[0] and (rx == wd[0]).all()
    er.close()


def test_single_api_ext_syndrome(gpu_device):
    """reference tests/test_codec.c:78-121: all-0xFF syndromes decode as clean"""
    syn = np.full(32, 0xFF, np.uint16)
    with api.Poporon.rs(syndrome=syn) as h:
        data = np.arange(64, dtype=np.uint8)
        parity = np.zeros(32, np.uint8)
        assert h.encode(data, parity)
        rx = data.copy()
        ok, c = h.decode(rx, parity)
        assert ok and c == 0 and (rx == data).all()
        syn[:] = np.arange(254, 222, -1)  # single error of magnitude 1 at location 0 (padding for size 64)
        ok, c = h.decode(rx, parity)
        assert not ok and c == 0 and (rx == data).all()


def test_bch_default_single_api(gpu_device):
    """reference tests/test_unified.c:366-389 and SURVEY B.8"""
    with api.Poporon.bch_default() as h:
        assert h.fec_type == api.FEC_BCH and h.info_size == 1 and h.parity_size == 2
        for dval, hexpar in [(0, "0000"), (7, "02b2"), (14, "0053"), (21, "0247"), (28, "00a6"), (0xF5, "0247")]:
            d = np.array([dval], np.uint8)
            p = np.zeros(2, np.uint8)
            assert h.encode(d, p) and p.tobytes().hex() == hexpar and d[0] == dval
        d = np.array([0x15], np.uint8)
        p = np.zeros(2, np.uint8)
        assert h.encode(d, p)
        ok, c = h.decode(d, p)
        assert ok and c == 0 and d[0] == 0x15
        d[0] = 0xF5
        ok, c = h.decode(d, p)
        assert ok and c == 0 and d[0] == 0x15  # pad bits forced to zero even on the clean path
        d[0] = 21 ^ 0x0A
        ok, c = h.decode(d, p)
        assert ok and c == 2 and d[0] == 21
        bad_p = p.copy()
        bad_p[1] ^= 0x07  # three parity-bit flips: corrected count 3, parity buffer left corrupted
        ok, c = h.decode(d, bad_p)
        assert ok and c == 3 and d[0] == 21 and (bad_p != p).any()
        d[0] = 21 ^ 0x1F
        ok, c = h.decode(d, p)
        assert not ok and c is None and d[0] == 21 ^ 0x1F  # failure: data and corrected_num untouched


@pytest.mark.parametrize("t", [1, 2, 3, 5, 7])
def test_bch31_batches_vs_oracle(gpu_device, orc, t):
    """BCH(31,k): the remainder-indexed decode map (up to the 25 parity bits of t = 7: a 128 MB map) against the oracle,
    0 .. t + 3 flipped bits anywhere in the 31, garbage above the k data bits; dense rows take the four-codewords-per-
    thread kernels, the last three codewords (n = 4 q + 3) the scalar ones."""
    info = orc.bch_info(5, 0x25, t)
    k, pbits = info["k"], info["parity_bits"]
    db, pb = (k + 7) // 8, (pbits + 7) // 8
    rng = np.random.default_rng(3100 + t)
    n = 20003
    with api.Poporon.bch(5, 0x25, t) as h:
        msg = rng.integers(0, 256, (n, db), dtype=np.uint8)
        par = np.full((n, pb), 0xEE, np.uint8)
        assert h.encode_batch(msg, db, db, par, pb, n)
        wpar, wok = orc.bch_encode(msg, pb, 5, 0x25, t)
        assert wok.all() and (par == wpar).all()
        d = msg.copy()
        pp = par.copy()
        for i in range(n):
            for bit in rng.choice(31, int(rng.integers(0, t + 4)), replace=False):
                if bit < pbits:
                    pp[i, pb - 1 - bit // 8] ^= 1 << (bit % 8)
                else:
                    b = bit - pbits
                    d[i, db - 1 - b // 8] ^= 1 << (b % 8)
        rx_d, rx_p = d.copy(), pp.copy()
        ok = np.full(n, 7, np.uint8)
        cor = np.full(n, 0xFFFFFFFF, np.uint32)
        assert h.decode_batch(d, db, db, pp, pb, n, ok, cor)
        wd, wok, wcor = orc.bch_decode(rx_d, rx_p, 5, 0x25, t)
        assert (ok == wok).all() and (cor == wcor).all() and (d == wd).all() and (pp == rx_p).all()
        assert int(wok.sum()) > 0 and (t == 1 or int(wok.sum()) < n)  # both outcomes (t = 1 is a perfect code)


def test_bch_word_level_api(gpu_device, orc):
    """poporon_bch_encode/decode/extract_data as used by reference tests/test_bch.c:61-187,251-305"""
    host = api.host
    for (m, poly, t, datas) in [(4, 0x13, 3, [0, 7, 21, 31]), (5, 0x25, 2, [0, 1, 0x12345, 0x1FFFFF])]:
        b = host.poporon_bch_create(m, poly, t)
        n = host.poporon_bch_get_codeword_length(b)
        k = host.poporon_bch_get_data_length(b)
        for data in datas:
            cw = C.c_uint32(0)
            assert host.poporon_bch_encode(b, data, C.byref(cw))
            fixed, ne = C.c_uint32(0), C.c_int32(-5)
            assert host.poporon_bch_decode(b, cw.value, C.byref(fixed), C.byref(ne))
            assert fixed.value == cw.value and ne.value == 0
            assert host.poporon_bch_extract_data(b, fixed.value) == data
            for bit in range(n):  # every single-bit error
                assert host.poporon_bch_decode(b, cw.value ^ (1 << bit), C.byref(fixed), C.byref(ne))
                assert fixed.value == cw.value and ne.value == 1
            for b1 in range(0, n, 3):  # a spread of double-bit errors
                for b2 in range(b1 + 1, n, 4):
                    assert host.poporon_bch_decode(b, cw.value ^ (1 << b1) ^ (1 << b2), C.byref(fixed), C.byref(ne))
                    assert fixed.value == cw.value and ne.value == 2
        assert not host.poporon_bch_encode(b, 1 << k, C.byref(cw))  # data >= 2^k
        host.poporon_bch_destroy(b)


def test_unsupported_long_bch_reports_failure(gpu_device):
    """n >= 63 is undefined in the reference: handles exist, codec calls report false"""
    with api.Poporon.bch(6, 0x43, 2) as h:
        d = np.zeros(8, np.uint8)
        p = np.zeros(8, np.uint8)
        assert not h.encode(d, p)
        ok, _ = h.decode(d, p)
        assert not ok


def test_use_devices_and_large_host_batch(gpu_device, orc):
    """host-pointer batch larger than one staging chunk, through the multi-lane planner"""
    rng = np.random.default_rng(31)
    n = 300_000
    msg = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        assert h.use_devices(None)  # all visible devices
        par = _gpu_encode(h, msg, 32)
        sample = rng.choice(n, 4000, replace=False)
        assert (par[sample] == orc.rs_encode(msg[sample])).all()
        cw = np.concatenate([msg, par], 1)
        hit = rng.choice(n, 30_000, replace=False)
        for i in hit[:3000]:
            e = int(rng.integers(1, 17))
            cw[i, rng.choice(255, e, replace=False)] ^= rng.integers(1, 256, e, dtype=np.uint8)
        d, p, ok, cor = _gpu_decode(h, cw[:, :223], cw[:, 223:])
        assert ok.all() and (d == msg).all() and (p == par).all()
        assert int((cor > 0).sum()) == 3000


@pytest.mark.parametrize("size,elo,ehi,n", [(223, 14, 40, 40_000), (223, 0, 16, 70_000), (120, 10, 24, 20_000)])
def test_rs_default_solver_heavy_loads(gpu_device, orc, size, elo, ehi, n):
    """the bitsliced solver (rs255_solve_core.h) and its hand-over of declined codewords (register length
    > 16, vanishing leading coefficient) to the literal stage: many uncorrectable words, several queue groups"""
    rng = np.random.default_rng(size * 1000 + ehi)
    msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
    cw = np.concatenate([msg, orc.rs_encode(msg, threads=8)], 1)
    L = size + 32
    e = rng.integers(elo, ehi + 1, n)
    order = np.argsort(rng.random((n, L)), axis=1)[:, :ehi]
    mags = rng.integers(1, 256, (n, ehi), dtype=np.uint8)
    rows = np.arange(n)
    for k in range(ehi):
        sel = e > k
        cw[rows[sel], order[sel, k]] ^= mags[sel, k]
    with api.Poporon.rs_default() as h:
        got = _gpu_decode(h, cw[:, :size], cw[:, size:])
    _assert_same(got, orc.rs_decode(cw[:, :size], cw[:, size:], threads=8), f"heavy {size} {elo}-{ehi}")


def test_host_batch_many_chunks_on_two_streams(gpu_device):
    """Packed 255-byte records on the host, 5 staging chunks, every codeword dirty (0-16 errors): consecutive
    chunks sit on different CUDA streams but share the device context's decode scratch, so their kernel sections
    must be ordered (pipeline.c).  Round-trip property at a size the oracle would not finish quickly."""
    rng = np.random.default_rng(77)
    n = (1 << 18) * 4 + 12345
    rec = np.zeros((n, 255), np.uint8)
    rec[:, :223] = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        flat = rec.reshape(-1)
        assert h.encode_batch(flat.ctypes.data, 255, 223, flat.ctypes.data + 223, 255, n)
        clean = rec.copy()
        e = rng.integers(0, 17, n)
        for k in range(16):
            rows = np.nonzero(e > k)[0]
            # distinct positions per row: position = (start + 16 * k + small offset) mod 255 never repeats for k < 16
            pos = (rows * 7 + 16 * k) % 255
            rec[rows, pos] ^= rng.integers(1, 256, len(rows), dtype=np.uint8)
        ok = np.zeros(n, np.uint8)
        cor = np.zeros(n, np.uint32)
        assert h.decode_batch(flat.ctypes.data, 255, 223, flat.ctypes.data + 223, 255, n, ok, cor)
        assert ok.all()
        assert (cor == e).all()
        assert (rec == clean).all()
