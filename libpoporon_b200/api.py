"""ctypes mirror of the C API.  Names, argument meaning and error behaviour follow the C functions
one to one (``poporon_create`` -> :class:`Poporon`, ``poporon_encode`` -> :meth:`Poporon.encode`, ...),
so the parity tests read like the reference's own tests (reference tests/test_codec.c,
tests/test_unified.c, tests/test_bch.c).

Device buffers are passed as raw integer addresses (``tensor.data_ptr()``); PyTorch is only the
allocator / stream provider in tests and bench, never part of the computation.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .build import lib_dir

FEC_RS, FEC_LDPC, FEC_BCH, FEC_UNKNOWN = 1, 2, 3, 255
MODE_PLAIN, MODE_ERASURE, MODE_EXTSYN = 0, 1, 2

_P = C.POINTER
_u8p, _u16p, _u32p = _P(C.c_uint8), _P(C.c_uint16), _P(C.c_uint32)


class NativeLibraryMissing(ImportError):
    pass


def _load():
    d = lib_dir()
    cuda_path = os.path.join(d, "libpoporon_cuda.so")
    host_path = os.path.join(d, "libpoporon.so")
    for p in (cuda_path, host_path):
        if not os.path.exists(p):
            raise NativeLibraryMissing(
                f"{p} is missing: build it with `python -m libpoporon_b200.build` "
                "(there is no Python/CPU fallback for the codecs)")
    cuda = C.CDLL(cuda_path, mode=C.RTLD_GLOBAL)
    host = C.CDLL(host_path)
    # workload generators of the bench and the tests: a library of their own, not part of the product
    synth_path = os.path.join(d, "libpoporon_synth.so")
    synth = C.CDLL(synth_path) if os.path.exists(synth_path) else None
    return cuda, host, synth


cuda, host, synth = _load()


def _sig(lib, name, restype, *argtypes):
    if lib is None:  # (libpoporon_synth.so absent: only the harness helpers are unavailable)
        return None
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


# ---- include/poporon.h -------------------------------------------------------------------------
_sig(host, "poporon_rs_config_create", C.c_void_p, C.c_uint8, C.c_uint16, C.c_uint16, C.c_uint16, C.c_uint8,
     C.c_void_p, C.c_void_p)
_sig(host, "poporon_bch_config_create", C.c_void_p, C.c_uint8, C.c_uint16, C.c_uint8)
_sig(host, "poporon_ldpc_config_create", C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_uint32, C.c_bool, C.c_bool,
     C.c_bool, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t, C.c_uint64)
_sig(host, "poporon_config_rs_default", C.c_void_p)
_sig(host, "poporon_config_bch_default", C.c_void_p)
_sig(host, "poporon_config_ldpc_default", C.c_void_p, C.c_size_t, C.c_int)
_sig(host, "poporon_config_ldpc_burst_resistant", C.c_void_p, C.c_size_t, C.c_int)
_sig(host, "poporon_config_destroy", None, C.c_void_p)
_sig(host, "poporon_create", C.c_void_p, C.c_void_p)
_sig(host, "poporon_destroy", None, C.c_void_p)
_sig(host, "poporon_encode", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)
_sig(host, "poporon_decode", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, _P(C.c_size_t))
_sig(host, "poporon_get_fec_type", C.c_int, C.c_void_p)
_sig(host, "poporon_get_iterations_used", C.c_uint32, C.c_void_p)
_sig(host, "poporon_get_parity_size", C.c_size_t, C.c_void_p)
_sig(host, "poporon_get_info_size", C.c_size_t, C.c_void_p)
_sig(host, "poporon_version_id", C.c_uint32)
_sig(host, "poporon_buildtime", C.c_uint32)
# ---- include/poporon/erasure.h, gf.h, rng.h ------------------------------------------------------
_sig(host, "poporon_erasure_create", C.c_void_p, C.c_uint16, C.c_uint32)
_sig(host, "poporon_erasure_create_from_positions", C.c_void_p, C.c_uint16, _u32p, C.c_uint32)
_sig(host, "poporon_erasure_add_position", C.c_bool, C.c_void_p, C.c_uint32)
_sig(host, "poporon_erasure_reset", None, C.c_void_p)
_sig(host, "poporon_erasure_destroy", None, C.c_void_p)
_sig(host, "poporon_gf_create", C.c_void_p, C.c_uint8, C.c_uint16)
_sig(host, "poporon_gf_destroy", None, C.c_void_p)
_sig(host, "poporon_gf_mod", C.c_uint8, C.c_void_p, C.c_uint16)
_sig(host, "poporon_rng_create", C.c_void_p, C.c_int, C.c_void_p, C.c_size_t)
_sig(host, "poporon_rng_destroy", None, C.c_void_p)
_sig(host, "poporon_rng_next", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t)
# ---- de-facto internal API used by the reference's tests (src/internal/common.h:112-124) -----------
_sig(host, "poporon_rs_create", C.c_void_p, C.c_uint8, C.c_uint16, C.c_uint16, C.c_uint16, C.c_uint8)
_sig(host, "poporon_rs_destroy", None, C.c_void_p)
_sig(host, "poporon_bch_create", C.c_void_p, C.c_uint8, C.c_uint16, C.c_uint8)
_sig(host, "poporon_bch_destroy", None, C.c_void_p)
_sig(host, "poporon_bch_get_codeword_length", C.c_uint16, C.c_void_p)
_sig(host, "poporon_bch_get_data_length", C.c_uint16, C.c_void_p)
_sig(host, "poporon_bch_get_correction_capability", C.c_uint8, C.c_void_p)
_sig(host, "poporon_bch_encode", C.c_bool, C.c_void_p, C.c_uint32, _u32p)
_sig(host, "poporon_bch_decode", C.c_bool, C.c_void_p, C.c_uint32, _u32p, _P(C.c_int32))
_sig(host, "poporon_bch_extract_data", C.c_uint32, C.c_void_p, C.c_uint32)
# ---- include/poporon/batch.h ---------------------------------------------------------------------
_sig(host, "poporon_use_devices", C.c_bool, C.c_void_p, _P(C.c_int), C.c_size_t)
_sig(host, "poporon_device_count", C.c_size_t)
_sig(host, "poporon_host_alloc", C.c_void_p, C.c_size_t)
_sig(host, "poporon_host_free", None, C.c_void_p)
_sig(host, "poporon_encode_batch", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t,
     C.c_size_t)
_sig(host, "poporon_decode_batch", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t,
     C.c_size_t, C.c_void_p, C.c_void_p)
_sig(host, "poporon_decode_batch_erasure", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
     C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p)
_sig(host, "poporon_decode_batch_syndrome", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
     C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p)
_sig(host, "poporon_syndromes_batch", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t,
     C.c_size_t, C.c_void_p)
_sig(host, "poporon_encode_batch_async", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
     C.c_size_t, C.c_size_t, C.c_void_p)
_sig(host, "poporon_decode_batch_async", C.c_bool, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
     C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p)
_sig(host, "poporon_kernel_launches", C.c_uint64, C.c_void_p)
_sig(host, "poporon_kernel_profile", C.c_bool, C.c_void_p, C.c_bool)
_sig(host, "poporon_kernel_profile_read", C.c_char_p, C.c_void_p, C.c_int, _P(C.c_uint64), _P(C.c_double))
_sig(host, "poporon_last_error", C.c_char_p)
# ---- include/poporon_cuda.h (the part tests/bench bind directly) ----------------------------------
_sig(cuda, "poporon_cuda_device_count", C.c_int)
_sig(cuda, "poporon_cuda_last_error", C.c_char_p)
_sig(synth, "poporon_cuda_synth_fill", C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint64,
     C.c_uint64, C.c_void_p)
_sig(synth, "poporon_cuda_synth_inject", C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_uint64,
     C.c_uint64, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p)
_sig(synth, "poporon_cuda_synth_erasures", C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
     C.c_size_t, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p)


def _addr(x):
    """numpy array -> host address; int -> itself (device pointer); None -> NULL."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    return int(x)


class Erasure:
    """poporon_erasure_t (include/poporon/erasure.h)."""

    def __init__(self, num_roots: int = 32, initial_capacity: int = 0, positions=None):
        if positions is not None:
            arr = np.ascontiguousarray(positions, np.uint32)
            self.ptr = host.poporon_erasure_create_from_positions(num_roots, arr.ctypes.data_as(_u32p), len(arr))
        else:
            self.ptr = host.poporon_erasure_create(num_roots, initial_capacity)

    def add(self, position: int) -> bool:
        return bool(host.poporon_erasure_add_position(self.ptr, position))

    def reset(self) -> None:
        host.poporon_erasure_reset(self.ptr)

    def close(self) -> None:
        if self.ptr:
            host.poporon_erasure_destroy(self.ptr)
            self.ptr = None


class Poporon:
    """poporon_t (include/poporon.h).  ``Poporon.rs(...)`` / ``Poporon.bch(...)`` build the config,
    create the handle and destroy the config, exactly like the reference's tests do."""

    def __init__(self, ptr):
        if not ptr:
            raise RuntimeError("poporon_create returned NULL: " + (host.poporon_last_error() or b"").decode())
        self.ptr = ptr
        self._keep = []

    # -- constructors --------------------------------------------------------------------------------
    @classmethod
    def rs(cls, symbol_size=8, gfpoly=0x11D, fcr=1, prim=1, nroots=32, erasure: Erasure | None = None,
           syndrome: np.ndarray | None = None):
        cfg = host.poporon_rs_config_create(symbol_size, gfpoly, fcr, prim, nroots,
                                            erasure.ptr if erasure else None, _addr(syndrome))
        try:
            self = cls(host.poporon_create(cfg))
        finally:
            host.poporon_config_destroy(cfg)
        self._keep = [erasure, syndrome]
        return self

    @classmethod
    def rs_default(cls):
        cfg = host.poporon_config_rs_default()
        try:
            return cls(host.poporon_create(cfg))
        finally:
            host.poporon_config_destroy(cfg)

    @classmethod
    def bch(cls, symbol_size=4, gfpoly=0x13, t=3):
        cfg = host.poporon_bch_config_create(symbol_size, gfpoly, t)
        try:
            return cls(host.poporon_create(cfg))
        finally:
            host.poporon_config_destroy(cfg)

    @classmethod
    def bch_default(cls):
        cfg = host.poporon_config_bch_default()
        try:
            return cls(host.poporon_create(cfg))
        finally:
            host.poporon_config_destroy(cfg)

    def close(self):
        if self.ptr:
            host.poporon_destroy(self.ptr)
            self.ptr = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- getters -------------------------------------------------------------------------------------
    @property
    def fec_type(self):
        return host.poporon_get_fec_type(self.ptr)

    @property
    def parity_size(self):
        return host.poporon_get_parity_size(self.ptr)

    @property
    def info_size(self):
        return host.poporon_get_info_size(self.ptr)

    @property
    def launches(self):
        return host.poporon_kernel_launches(self.ptr)

    def profile(self, on=True) -> bool:
        """start / stop per-kernel CUDA-event timing on the devices this handle has used so far"""
        return bool(host.poporon_kernel_profile(self.ptr, bool(on)))

    def profile_read(self):
        """{kernel name: (launches, total ms)} accumulated since profile(True)"""
        out, k = {}, 0
        while True:
            n, ms = C.c_uint64(0), C.c_double(0.0)
            name = host.poporon_kernel_profile_read(self.ptr, k, C.byref(n), C.byref(ms))
            if name is None:
                return out
            if n.value:
                out[name.decode()] = (int(n.value), float(ms.value))
            k += 1

    def use_devices(self, devices=None) -> bool:
        if not devices:
            return bool(host.poporon_use_devices(self.ptr, None, 0))
        arr = (C.c_int * len(devices))(*devices)
        return bool(host.poporon_use_devices(self.ptr, arr, len(devices)))

    # -- single codeword (host numpy buffers, modified in place like the C API) -----------------------
    def encode(self, data: np.ndarray, parity: np.ndarray, size=None) -> bool:
        size = len(data) if size is None else size
        return bool(host.poporon_encode(self.ptr, _addr(data), size, _addr(parity)))

    def decode(self, data: np.ndarray, parity: np.ndarray, size=None, want_count=True):
        """Returns (ok, corrected_num or None-if-untouched)."""
        size = len(data) if size is None else size
        if not want_count:
            return bool(host.poporon_decode(self.ptr, _addr(data), size, _addr(parity), None)), None
        sentinel = 0xDEADBEEF
        cnt = C.c_size_t(sentinel)
        ok = bool(host.poporon_decode(self.ptr, _addr(data), size, _addr(parity), C.byref(cnt)))
        return ok, (None if cnt.value == sentinel else cnt.value)

    # -- batches (numpy arrays = host pointers, ints / torch tensors = device pointers) ---------------
    def encode_batch(self, data, data_stride, size, parity, parity_stride, n) -> bool:
        return bool(host.poporon_encode_batch(self.ptr, _addr(data), data_stride, size, _addr(parity),
                                              parity_stride, n))

    def decode_batch(self, data, data_stride, size, parity, parity_stride, n, ok=None, corrected=None) -> bool:
        return bool(host.poporon_decode_batch(self.ptr, _addr(data), data_stride, size, _addr(parity), parity_stride,
                                              n, _addr(ok), _addr(corrected)))

    def decode_batch_erasure(self, data, data_stride, size, parity, parity_stride, n, pos, pos_stride, cnt, ok=None,
                             corrected=None) -> bool:
        return bool(host.poporon_decode_batch_erasure(self.ptr, _addr(data), data_stride, size, _addr(parity),
                                                      parity_stride, n, _addr(pos), pos_stride, _addr(cnt),
                                                      _addr(ok), _addr(corrected)))

    def decode_batch_syndrome(self, data, data_stride, size, parity, parity_stride, n, syn, ok=None,
                              corrected=None) -> bool:
        return bool(host.poporon_decode_batch_syndrome(self.ptr, _addr(data), data_stride, size, _addr(parity),
                                                       parity_stride, n, _addr(syn), _addr(ok), _addr(corrected)))

    def syndromes_batch(self, data, data_stride, size, parity, parity_stride, n, syn) -> bool:
        """syn: [n, nroots] uint16, receives the index-form syndromes (2^m - 1 = zero); nothing is corrected"""
        return bool(host.poporon_syndromes_batch(self.ptr, _addr(data), data_stride, size, _addr(parity),
                                                 parity_stride, n, _addr(syn)))

    def encode_batch_async(self, data, data_stride, size, parity, parity_stride, n, stream=0) -> bool:
        return bool(host.poporon_encode_batch_async(self.ptr, _addr(data), data_stride, size, _addr(parity),
                                                    parity_stride, n, stream))

    def decode_batch_async(self, data, data_stride, size, parity, parity_stride, n, ok=None, corrected=None, pos=None,
                           pos_stride=0, cnt=None, stream=0) -> bool:
        return bool(host.poporon_decode_batch_async(self.ptr, _addr(data), data_stride, size, _addr(parity),
                                                    parity_stride, n, _addr(pos), pos_stride, _addr(cnt), _addr(ok),
                                                    _addr(corrected), stream))


# ---- synthetic workloads (harness) ----------------------------------------------------------------
def synth_fill(device, data, stride, size, n, seed, i0=0, stream=0):
    rc = synth.poporon_cuda_synth_fill(device, _addr(data), stride, size, n, seed, i0, stream)
    if rc != 0:
        raise RuntimeError(cuda.poporon_cuda_last_error().decode())


def synth_inject(device, cw, stride, length, n, seed, i0=0, kind=0, emin=0, emax=16, p_ppm=0, stream=0):
    rc = synth.poporon_cuda_synth_inject(device, _addr(cw), stride, length, n, seed, i0, kind, emin, emax, p_ppm,
                                        stream)
    if rc != 0:
        raise RuntimeError(cuda.poporon_cuda_last_error().decode())


def synth_erasures(device, cw, stride, size, length, n, seed, pos, pos_stride, cnt, i0=0, cmax=32, stream=0):
    rc = synth.poporon_cuda_synth_erasures(device, _addr(cw), stride, size, length, n, seed, i0, cmax, _addr(pos),
                                          pos_stride, _addr(cnt), stream)
    if rc != 0:
        raise RuntimeError(cuda.poporon_cuda_last_error().decode())
