"""CPU test of the bitsliced RS(255,223) solver arithmetic (libpoporon_b200/csrc/cuda/rs255_solve_core.h).

The header is plain C++; tests/native/host_solver.cpp runs it on the host one lane (32 codewords) at a
time.  Here its outcome (status, root count, locations, magnitudes) is compared with the oracle's
poporon_decode on the same received words: for every codeword the solver accepts, applying its error list
must reproduce the oracle's corrected bytes, corrected_num and success flag; codewords it declines are
counted (they go to the literal per-codeword stage on the device) and must be rare."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
from oraclelib import Oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "native", "host_solver.cpp")
INC = os.path.join(ROOT, "libpoporon_b200", "csrc", "cuda")
OUT = os.path.join(ROOT, "build", "tests", "host_solver.so")


# both forms of Berlekamp-Massey in the header: coefficients rotating through registers / locator in memory
@pytest.fixture(scope="module", params=[0, 1], ids=["bm-registers", "bm-memory"])
def solver(request):
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    out = OUT.replace(".so", f"_bm{request.param}.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", f"-DBS_BM_MEM={request.param}", f"-I{INC}", "-o",
                    out, SRC], check=True)
    lib = C.CDLL(out)
    P = C.POINTER
    lib.host_solve.argtypes = [P(C.c_uint8), C.c_size_t, C.c_int, P(C.c_int8), P(C.c_uint8), P(C.c_uint16)]
    lib.host_solve_roots.argtypes = [P(C.c_uint8), C.c_size_t, C.c_int, C.c_int, P(C.c_int8), P(C.c_uint8),
                                     P(C.c_uint16)]
    return lib


def gf_tables():
    exp = np.zeros(512, np.int64)
    log = np.zeros(256, np.int64)
    x = 1
    for i in range(255):
        exp[i] = x
        log[x] = i
        x <<= 1
        if x & 0x100:
            x ^= 0x11D
    exp[255:510] = exp[0:255]
    return exp, log


def syndromes(cw):
    """S_j = r(alpha^j), j = 1..32, first byte = highest degree."""
    exp, log = gf_tables()
    n, length = cw.shape
    out = np.zeros((n, 32), np.uint8)
    powers = (length - 1 - np.arange(length))
    nz = cw != 0
    lg = log[cw]
    for j in range(1, 33):
        e = (lg + (j * powers)[None, :]) % 255
        v = np.where(nz, exp[e], 0)
        out[:, j - 1] = np.bitwise_xor.reduce(v, axis=1)
    return out


def run_case(solver, orc, rng, n, size, emin, emax, nr=32):
    params = (8, 0x11D, 1, 1, nr)
    msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
    par = orc.rs_encode(msg, params)
    cw = np.concatenate([msg, par], 1)
    length = size + nr
    for i in range(n):
        e = int(rng.integers(emin, emax + 1))
        pos = rng.choice(length, e, replace=False)
        cw[i, pos] ^= rng.integers(1, 256, e, dtype=np.uint8)
    wd, wp, wok, wcor = orc.rs_decode(cw[:, :size], cw[:, size:], params)
    syn = np.ascontiguousarray(syndromes(cw))
    dirty = syn[:, :nr].any(axis=1)  # (the entries beyond nr are evaluations the code does not constrain: garbage)
    status = np.zeros(n, np.int8)
    nroots = np.zeros(n, np.uint8)
    lists = np.zeros((n, 16), np.uint16)
    pad = 255 - nr - size
    P = C.POINTER
    solver.host_solve_roots(syn.ctypes.data_as(P(C.c_uint8)), n, pad, nr, status.ctypes.data_as(P(C.c_int8)),
                            nroots.ctypes.data_as(P(C.c_uint8)), lists.ctypes.data_as(P(C.c_uint16)))
    declined = 0
    for i in range(n):
        if not dirty[i]:
            assert wok[i] == 1 and wcor[i] == 0
            continue
        if status[i] == 0:
            declined += 1
            continue
        if status[i] == 1:
            assert wok[i] == 0 and wcor[i] == 0, (i, wok[i], wcor[i])
            assert (wd[i] == cw[i, :size]).all() and (wp[i] == cw[i, size:]).all()
            continue
        assert wok[i] == 1, i
        fixed = cw[i].copy()
        nnz = 0
        for e in range(nroots[i]):
            loc, mag = int(lists[i, e]) & 0xFF, int(lists[i, e]) >> 8
            assert loc >= pad
            fixed[loc - pad] ^= mag
            nnz += mag != 0
        assert nnz == wcor[i], (i, nnz, wcor[i])
        assert (fixed[:size] == wd[i]).all() and (fixed[size:] == wp[i]).all(), i
    return declined, int(dirty.sum())


def test_solver_core_correctable(solver):
    orc = Oracle("port")
    rng = np.random.default_rng(11)
    declined, dirty = run_case(solver, orc, rng, 640, 223, 0, 16)
    assert declined == 0 and dirty > 500


def test_solver_core_uncorrectable(solver):
    orc = Oracle("port")
    rng = np.random.default_rng(12)
    declined, dirty = run_case(solver, orc, rng, 960, 223, 15, 40)
    assert dirty == 960
    assert declined < 0.15 * dirty  # L > 16 happens for ~6 % of random uncorrectable words


def test_solver_core_shortened(solver):
    orc = Oracle("port")
    rng = np.random.default_rng(13)
    for size in (1, 17, 64, 200):
        run_case(solver, orc, rng, 256, size, 0, 19)


def test_solver_core_16_and_8_roots(solver):
    """RS(255,239) and RS(255,247): the same core with 16 / 8 syndromes (BM iterations, length counter, evaluator and
    Forney bounds follow the syndrome count)"""
    orc = Oracle("port")
    rng = np.random.default_rng(14)
    for nr in (16, 8):
        declined, dirty = run_case(solver, orc, rng, 640, 255 - nr, 0, nr // 2, nr)
        assert declined == 0 and dirty > 400
        run_case(solver, orc, rng, 640, 255 - nr, nr // 2, nr + 4, nr)
        run_case(solver, orc, rng, 256, 100, 0, nr // 2 + 2, nr)


def test_erasure_core(solver):
    """Erasure-only form (rs255_erasure_core.h) against the oracle's erasure mode, for the codes with 32, 16 and 8 roots:
    rows whose errors all sit on the declared, ascending positions give the reference's magnitudes and count; rows with
    an error elsewhere are flagged as not qualifying."""
    from oraclelib import MODE_ERASURE

    orc = Oracle("port")
    rng = np.random.default_rng(15)
    P = C.POINTER
    solver.host_erasure_roots.argtypes = [P(C.c_uint8), P(C.c_uint8), C.c_size_t, C.c_int, P(C.c_int8), P(C.c_uint8),
                                          P(C.c_uint8)]
    for nr, size in ((32, 223), (32, 100), (32, 33), (16, 239), (16, 50), (8, 247), (8, 9)):
        params = (8, 0x11D, 1, 1, nr)
        n = 320
        pad = 255 - nr - size
        msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
        par = orc.rs_encode(msg, params)
        cw = np.concatenate([msg, par], 1)
        pos = np.zeros((n, 32), np.uint32)
        cnt = np.zeros(n, np.uint32)
        lists = np.zeros((n, 32), np.uint8)
        stray = np.zeros(n, bool)
        for i in range(n):
            # (lanes of 32 rows with a small largest count exercise the truncated loops)
            fmax = min(nr, size) if (i // 32) % 2 == 0 else min(3, size)
            f = int(rng.integers(1, fmax + 1))
            p = np.sort(rng.choice(size, f, replace=False))
            pos[i, :f], cnt[i] = p, f
            lists[i, :f] = p + pad + 1
            cw[i, p] ^= rng.integers(0, 256, f, dtype=np.uint8)  # (zero magnitudes allowed)
            if rng.random() < 0.3:  # an error the caller did not declare, anywhere in the codeword
                others = np.setdiff1d(np.arange(size + nr), p)
                q = rng.choice(others, int(rng.integers(1, 3)), replace=False)
                cw[i, q] ^= rng.integers(1, 256, len(q), dtype=np.uint8)
                stray[i] = True
        wd, wp, wok, wcor = orc.rs_decode(cw[:, :size], cw[:, size:], params, mode=MODE_ERASURE, pos=pos, cnt=cnt)
        syn = np.ascontiguousarray(syndromes(cw))
        bad = np.zeros(n, np.int8)
        mags = np.zeros((n, 32), np.uint8)
        count = np.zeros(n, np.uint8)
        solver.host_erasure_roots(syn.ctypes.data_as(P(C.c_uint8)), lists.ctypes.data_as(P(C.c_uint8)), n, nr,
                                  bad.ctypes.data_as(P(C.c_int8)), mags.ctypes.data_as(P(C.c_uint8)),
                                  count.ctypes.data_as(P(C.c_uint8)))
        # A row qualifies exactly when the reference's Berlekamp-Massey would see only zero discrepancies, i.e. when it
        # leaves the erasure locator alone: then the reference's result IS these magnitudes at the declared positions,
        # also in the rare rows where an undeclared error goes unnoticed (as many declared positions as the code has
        # roots explain any syndromes).
        good = np.nonzero(~stray)[0]
        assert (bad[good] == 0).all(), "a row without undeclared errors did not qualify"
        assert bad[stray].mean() > 0.8
        for i in np.nonzero(bad == 0)[0]:
            f = int(cnt[i])
            assert wok[i] == 1, (nr, size, i)
            fixed = cw[i, :size].copy()
            fixed[pos[i, :f]] ^= mags[i, :f]
            assert (mags[i, f:] == 0).all()
            assert (fixed == wd[i]).all() and (cw[i, size:] == wp[i]).all(), (nr, size, i)
            assert count[i] == wcor[i] == np.count_nonzero(mags[i, :f]), (nr, size, i, count[i], wcor[i])


def test_errors_on_top_of_erasures_core(solver):
    """ee_lane / ee_products (rs255_erasure_core.h): sigma from the plain BM on the Forney syndromes, the merged rank
    list and its magnitudes, against the oracle's erasure mode - including its slot-order application: the i-th
    magnitude goes to the caller's i-th slot, whatever that slot holds beyond the declared count."""
    from oraclelib import MODE_ERASURE

    orc = Oracle("port")
    rng = np.random.default_rng(16)
    P = C.POINTER
    solver.host_ee.argtypes = [P(C.c_uint8), P(C.c_uint8), C.c_size_t, P(C.c_int), C.c_int, P(C.c_int8), P(C.c_uint8),
                               P(C.c_uint8), P(C.c_uint8)]
    for size in (223, 90):
        n = 32 * 24
        pad = 223 - size
        msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
        par = orc.rs_encode(msg)
        cw = np.concatenate([msg, par], 1)
        pos = np.zeros((n, 32), np.uint32)
        cnt = np.zeros(n, np.uint32)
        lists = np.zeros((n, 32), np.uint8)
        fcount = np.zeros(n // 32, np.int32)
        for g in range(n // 32):
            F = int(rng.integers(1, 31))
            fcount[g] = F
            for i in range(32 * g, 32 * g + 32):
                p = np.sort(rng.choice(size, F, replace=False))
                pos[i, :F], cnt[i] = p, F
                pos[i, F:] = rng.integers(0, size, 32 - F)  # stale slots: the reference applies through them
                lists[i, :F] = p + pad + 1
                cw[i, p] ^= rng.integers(0, 256, F, dtype=np.uint8)
                emax = (32 - F) // 2
                e = int(rng.integers(1, emax + 3)) if emax else 1  # sometimes more than can be corrected
                others = np.setdiff1d(np.arange(size + 32), p)
                q = rng.choice(others, min(e, len(others)), replace=False)
                cw[i, q] ^= rng.integers(1, 256, len(q), dtype=np.uint8)
        wd, wp, wok, wcor = orc.rs_decode(cw[:, :size], cw[:, size:], mode=MODE_ERASURE, pos=pos, cnt=cnt)
        syn = np.ascontiguousarray(syndromes(cw))
        status = np.zeros(n, np.int8)
        merged = np.zeros((n, 32), np.uint8)
        mags = np.zeros((n, 32), np.uint8)
        count = np.zeros(n, np.uint8)
        solver.host_ee(syn.ctypes.data_as(P(C.c_uint8)), lists.ctypes.data_as(P(C.c_uint8)), n,
                       fcount.ctypes.data_as(P(C.c_int)), pad, status.ctypes.data_as(P(C.c_int8)),
                       merged.ctypes.data_as(P(C.c_uint8)), mags.ctypes.data_as(P(C.c_uint8)),
                       count.ctypes.data_as(P(C.c_uint8)))
        assert (status == 2).sum() > n // 3
        for i in range(n):
            if status[i] == 0:
                continue
            if status[i] == 1:
                assert wok[i] == 0 and wcor[i] == 0, (i, wok[i], wcor[i])
                assert (wd[i] == cw[i, :size]).all()
                continue
            k = int(np.count_nonzero(merged[i]))
            assert wok[i] == 1, i
            fixed = cw[i, :size].copy()
            for r in range(k):  # rank r -> the caller's slot r
                if pos[i, r] < size:
                    fixed[pos[i, r]] ^= mags[i, r]
            assert (fixed == wd[i]).all() and (cw[i, size:] == wp[i]).all(), i
            assert count[i] == wcor[i], (i, count[i], wcor[i])


def test_cores_under_address_sanitizer():
    """The three bitsliced cores (plain solver, erasure-only, errors on top of erasures; 32 / 16 / 8 roots) on random
    syndromes and slot lists under -fsanitize=address,undefined (tests/native/asan_main.cpp): no access outside their
    areas whatever the data (compute-sanitizer is not available on the GPU pool)."""
    exe = os.path.join(os.path.dirname(OUT), "asan_solver")
    main = os.path.join(ROOT, "tests", "native", "asan_main.cpp")
    subprocess.run(["g++", "-O1", "-std=c++17", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", f"-I{INC}",
                    "-o", exe, main, SRC], check=True)
    run = subprocess.run([exe], capture_output=True, text=True)
    assert run.returncode == 0 and "asan run done" in run.stdout, run.stderr[-2000:]
