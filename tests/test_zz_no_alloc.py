"""GPU test of poporon_cuda_rs_encode_noalloc_dev (include/poporon_cuda.h): the encoder whose scratch is a static
device array.  Called through the C ABI of libpoporon_cuda.so, its parity rows must be bit-identical to those of
poporon_cuda_rs_encode_dev on the same device buffers - on calls in a row that share the array, with different
batch sizes and layouts, so that whatever one call leaves behind is shown not to reach the next.

Named test_zz_* so that pytest collects it after every other test file."""
import ctypes as C

import numpy as np
import pytest

import libpoporon_b200.api as api

pytestmark = pytest.mark.gpu

MAX_CODEWORDS = 65536  # POPORON_CUDA_NOALLOC_MAX_CODEWORDS
ERR_UNSUPP = -4


class _Desc(C.Structure):  # poporon_cuda_rs_desc_t
    _fields_ = [("symbol_size", C.c_uint8), ("nroots", C.c_uint8), ("nn", C.c_uint16), ("gfpoly", C.c_uint16),
                ("fcr", C.c_uint16), ("prim", C.c_uint16), ("iprim", C.c_uint16),
                ("alpha_to", C.POINTER(C.c_uint16)), ("index_of", C.POINTER(C.c_uint16)),
                ("genpoly_index", C.POINTER(C.c_uint16))]


def _bind():
    L = api.cuda
    vp, sz = C.c_void_p, C.c_size_t
    L.poporon_cuda_rs_create.restype = C.c_int
    L.poporon_cuda_rs_create.argtypes = [C.POINTER(_Desc), C.c_int, C.POINTER(vp)]
    L.poporon_cuda_rs_destroy.restype = None
    L.poporon_cuda_rs_destroy.argtypes = [vp]
    enc = [vp, vp, sz, sz, vp, sz, sz, vp, vp]
    for name in ("poporon_cuda_rs_encode_dev", "poporon_cuda_rs_encode_noalloc_dev"):
        getattr(L, name).restype = C.c_int
        getattr(L, name).argtypes = enc
    L.poporon_cuda_rs_path.restype = C.c_int
    L.poporon_cuda_rs_path.argtypes = [vp, C.c_int, vp, sz, sz, vp, sz]
    L.poporon_cuda_malloc.restype = C.c_int
    L.poporon_cuda_malloc.argtypes = [C.c_int, sz, C.POINTER(vp)]
    L.poporon_cuda_free.restype = C.c_int
    L.poporon_cuda_free.argtypes = [C.c_int, vp]
    for name in ("poporon_cuda_memcpy_h2d", "poporon_cuda_memcpy_d2h"):
        getattr(L, name).restype = C.c_int
        getattr(L, name).argtypes = [C.c_int, vp, vp, sz, vp]
    L.poporon_cuda_memset.restype = C.c_int
    L.poporon_cuda_memset.argtypes = [C.c_int, vp, C.c_int, sz, vp]
    L.poporon_cuda_stream_sync.restype = C.c_int
    L.poporon_cuda_stream_sync.argtypes = [C.c_int, vp]
    return L


def _rs255_desc(nroots):
    """GF(256) / 0x11D tables and the generator polynomial in index form, fcr = 1, prim = 1 (what the host library
    computes following reference src/gf.c:29-86 and src/rs.c:29-83)."""
    nn = 255
    alpha_to = np.zeros(nn + 1, np.uint16)
    index_of = np.zeros(nn + 1, np.uint16)
    x = 1
    for i in range(nn):
        alpha_to[i] = x
        index_of[x] = i
        x <<= 1
        if x & 0x100:
            x ^= 0x11D
    index_of[0] = nn
    alpha_to[nn] = 0

    def mul(a, b):
        return 0 if a == 0 or b == 0 else int(alpha_to[(int(index_of[a]) + int(index_of[b])) % nn])

    g = [1]
    for i in range(nroots):  # times (x + alpha^(1 + i)), coefficients lowest degree first
        root = int(alpha_to[(1 + i) % nn])
        g = [(mul(g[k], root) if k < len(g) else 0) ^ (g[k - 1] if k else 0) for k in range(len(g) + 1)]
    genpoly = np.array([index_of[c] for c in g], np.uint16)
    keep = (alpha_to, index_of, genpoly)
    u16p = C.POINTER(C.c_uint16)
    d = _Desc(8, nroots, nn, 0x11D, 1, 1, 1, alpha_to.ctypes.data_as(u16p), index_of.ctypes.data_as(u16p),
              genpoly.ctypes.data_as(u16p))
    return d, keep


class _Dev:
    """A device buffer from the library's own helpers (the entry point under test allocates nothing itself)."""

    def __init__(self, L, nbytes):
        self.L, self.nbytes = L, nbytes
        self.ptr = C.c_void_p()
        assert L.poporon_cuda_malloc(0, max(nbytes, 1), C.byref(self.ptr)) == 0, api.cuda.poporon_cuda_last_error()

    def put(self, arr):
        a = np.ascontiguousarray(arr)
        assert a.nbytes <= self.nbytes
        assert self.L.poporon_cuda_memcpy_h2d(0, self.ptr, a.ctypes.data, a.nbytes, None) == 0
        assert self.L.poporon_cuda_stream_sync(0, None) == 0

    def get(self, nbytes):
        out = np.empty(nbytes, np.uint8)
        assert self.L.poporon_cuda_memcpy_d2h(0, out.ctypes.data, self.ptr, nbytes, None) == 0
        assert self.L.poporon_cuda_stream_sync(0, None) == 0
        return out

    def fill(self, value):
        assert self.L.poporon_cuda_memset(0, self.ptr, value, self.nbytes, None) == 0

    def free(self):
        self.L.poporon_cuda_free(0, self.ptr)


def _both(L, ctx, rng, n, size, nroots, packed):
    """Encode one random batch with both entry points; returns [parity rows of the allocating call, parity rows of
    the static-scratch call, the message rows]."""
    msg = rng.integers(0, 256, (n, size), dtype=np.uint8)
    if packed:  # records of message + parity, parity = data + size
        dstride = pstride = size + nroots
        rec = np.full((n, dstride), 0xEE, np.uint8)
        rec[:, :size] = msg
        d_data = _Dev(L, rec.nbytes)
        out = []
        for fn in (L.poporon_cuda_rs_encode_dev, L.poporon_cuda_rs_encode_noalloc_dev):
            d_data.put(rec)
            par = C.c_void_p(d_data.ptr.value + size)
            assert L.poporon_cuda_rs_path(ctx, 0, d_data.ptr, dstride, size, par, pstride) == 1
            rc = fn(ctx, d_data.ptr, dstride, size, par, pstride, n, None, None)
            assert rc == 0, (rc, L.poporon_cuda_last_error())
            got = d_data.get(rec.nbytes).reshape(n, dstride)
            assert (got[:, :size] == msg).all(), "the encoder wrote into message bytes"
            out.append(got[:, size:].copy())
        d_data.free()
        return out + [msg]
    d_data, d_par, d_ok = _Dev(L, msg.nbytes), _Dev(L, n * nroots), _Dev(L, n)
    d_data.put(msg)
    out = []
    for fn in (L.poporon_cuda_rs_encode_dev, L.poporon_cuda_rs_encode_noalloc_dev):
        d_par.fill(0xEE)
        d_ok.fill(7)
        assert L.poporon_cuda_rs_path(ctx, 0, d_data.ptr, size, size, d_par.ptr, nroots) == 1
        rc = fn(ctx, d_data.ptr, size, size, d_par.ptr, nroots, n, d_ok.ptr, None)
        assert rc == 0, (rc, L.poporon_cuda_last_error())
        assert (d_ok.get(n) == 1).all()
        out.append(d_par.get(n * nroots).reshape(n, nroots))
    for b in (d_data, d_par, d_ok):
        b.free()
    return out + [msg]


def test_static_scratch_encoder_matches_the_allocating_one(gpu_device, orc):
    L = _bind()
    rng = np.random.default_rng(20261105)
    desc, keep = _rs255_desc(32)
    ctx = C.c_void_p()
    assert L.poporon_cuda_rs_create(C.byref(desc), 0, C.byref(ctx)) == 0, L.poporon_cuda_last_error()
    try:
        # calls in a row on the shared array: the largest batch first, then a ragged shortened one that reuses only
        # part of what the first left behind, then the first shape again; bit-exact each time
        for n, size, packed in ((MAX_CODEWORDS, 223, False), (5003, 100, True), (MAX_CODEWORDS, 223, True),
                                (1, 1, False), (33, 223, False)):
            want, got, msg = _both(L, ctx, rng, n, size, 32, packed)
            assert (want[:64] == orc.rs_encode(msg[:64])).all(), "the allocating entry point disagrees with the oracle"
            bad = np.nonzero((want != got).any(1))[0]
            assert bad.size == 0, f"n={n} size={size} packed={packed}: {bad.size} parity rows differ, first {bad[:8]}"
        # beyond the static scratch: refused with a reason, nothing launched
        d = _Dev(L, 64)
        rc = L.poporon_cuda_rs_encode_noalloc_dev(ctx, d.ptr, 223, 223, d.ptr, 32, MAX_CODEWORDS + 1, None, None)
        assert rc == ERR_UNSUPP
        assert b"POPORON_CUDA_NOALLOC_MAX_CODEWORDS" in L.poporon_cuda_last_error()
        d.free()
    finally:
        L.poporon_cuda_rs_destroy(ctx)
    del keep


def test_static_scratch_encoder_other_codes(gpu_device):
    """RS(255,239) and RS(255,247) share the array; a code outside the bitsliced path is refused."""
    L = _bind()
    rng = np.random.default_rng(7)
    for nroots in (16, 8):
        desc, keep = _rs255_desc(nroots)
        ctx = C.c_void_p()
        assert L.poporon_cuda_rs_create(C.byref(desc), 0, C.byref(ctx)) == 0, L.poporon_cuda_last_error()
        try:
            for n, packed in ((20001, False), (4097, True)):
                want, got, _ = _both(L, ctx, rng, n, 255 - nroots, nroots, packed)
                assert (want == got).all(), (nroots, n, packed)
        finally:
            L.poporon_cuda_rs_destroy(ctx)
    desc, keep = _rs255_desc(10)
    ctx = C.c_void_p()
    assert L.poporon_cuda_rs_create(C.byref(desc), 0, C.byref(ctx)) == 0, L.poporon_cuda_last_error()
    try:
        d = _Dev(L, 4096)
        assert L.poporon_cuda_rs_encode_noalloc_dev(ctx, d.ptr, 100, 100, C.c_void_p(d.ptr.value + 2048), 10, 8, None,
                                                    None) == ERR_UNSUPP
        d.free()
    finally:
        L.poporon_cuda_rs_destroy(ctx)
