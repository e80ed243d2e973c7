"""ctypes front-end to the CPU oracle libraries (oracle/liboracle.so = our restatement, symbols
``orc_*``; oracle/_ref/libpoporon_ref.so = the unmodified reference + ref_shim.c, symbols ``ref_*``).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

MODE_PLAIN, MODE_ERASURE, MODE_EXTSYN = 0, 1, 2


class RsParams(C.Structure):
    _fields_ = [("symbol_size", C.c_uint8), ("gfpoly", C.c_uint16), ("fcr", C.c_uint16),
                ("prim", C.c_uint16), ("nroots", C.c_uint8)]


RS_DEFAULT = (8, 0x11D, 1, 1, 32)


def build_oracle() -> None:
    """(Re)build liboracle.so and, when /root/reference is present, oracle/_ref."""
    subprocess.run(["make", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)


def _u8p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8)) if a is not None else None


def _u32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32)) if a is not None else None


def _u16p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint16)) if a is not None else None


class Oracle:
    """One of the two oracle libraries behind a numpy interface."""

    def __init__(self, kind: str = "port"):
        if kind == "port":
            path, self.pfx = os.path.join(ORACLE_DIR, "liboracle.so"), "orc"
        elif kind == "reference":
            path, self.pfx = os.path.join(ORACLE_DIR, "_ref", "libpoporon_ref.so"), "ref"
        else:
            raise ValueError(kind)
        if not os.path.exists(path) and kind == "port":
            build_oracle()
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.kind = kind
        self.lib = C.CDLL(path)
        f = self._fn
        P = C.POINTER
        f("rs_encode_batch").argtypes = [P(RsParams), P(C.c_uint8), C.c_size_t, C.c_size_t, P(C.c_uint8),
                                         C.c_size_t, C.c_size_t, P(C.c_uint8), C.c_int]
        f("rs_decode_batch").argtypes = [P(RsParams), C.c_int, P(C.c_uint8), C.c_size_t, C.c_size_t, P(C.c_uint8),
                                         C.c_size_t, C.c_size_t, P(C.c_uint32), C.c_size_t, P(C.c_uint32),
                                         P(C.c_uint16), P(C.c_uint8), P(C.c_uint32), C.c_int]
        f("bch_encode_batch").argtypes = [C.c_uint8, C.c_uint16, C.c_uint8, P(C.c_uint8), C.c_size_t, C.c_size_t,
                                          P(C.c_uint8), C.c_size_t, C.c_size_t, P(C.c_uint8), C.c_int]
        f("bch_decode_batch").argtypes = [C.c_uint8, C.c_uint16, C.c_uint8, P(C.c_uint8), C.c_size_t, C.c_size_t,
                                          P(C.c_uint8), C.c_size_t, C.c_size_t, P(C.c_uint8), P(C.c_uint32), C.c_int]
        f("bch_info").argtypes = [C.c_uint8, C.c_uint16, C.c_uint8, P(C.c_uint32)]
        f("rs_info").argtypes = [P(RsParams), P(C.c_uint16), P(C.c_uint16), P(C.c_uint16), P(C.c_uint16)]
        f("gf_mod").argtypes = [C.c_uint8, C.c_uint16, C.c_uint16]
        f("gf_mod").restype = C.c_uint8
        f("kind").restype = C.c_char_p

    def _fn(self, name):
        return getattr(self.lib, f"{self.pfx}_{name}")

    # ------------------------------------------------------------------ RS
    def rs_info(self, params=RS_DEFAULT):
        p = RsParams(*params)
        nn = (1 << params[0]) - 1
        gen = np.zeros(params[4] + 1, np.uint16)
        a2 = np.zeros(nn + 1, np.uint16)
        io = np.zeros(nn + 1, np.uint16)
        ip = C.c_uint16(0)
        rc = self._fn("rs_info")(C.byref(p), _u16p(gen), _u16p(a2), _u16p(io), C.byref(ip))
        if rc != 0:
            return None
        return dict(genpoly=gen, alpha_to=a2, index_of=io, iprim=ip.value)

    def rs_encode(self, data: np.ndarray, params=RS_DEFAULT, threads: int = 1):
        """data: [n, size] uint8 (C-contiguous).  Returns parity [n, nroots] (or None if the codec is invalid)."""
        data = np.ascontiguousarray(data, np.uint8)
        n, size = data.shape
        p = RsParams(*params)
        parity = np.zeros((n, params[4]), np.uint8)
        ok = np.zeros(n, np.uint8)
        rc = self._fn("rs_encode_batch")(C.byref(p), _u8p(data), size, size, _u8p(parity), params[4], n, _u8p(ok),
                                         threads)
        return None if rc != 0 else parity

    def rs_decode(self, data: np.ndarray, parity: np.ndarray, params=RS_DEFAULT, mode=MODE_PLAIN, pos=None,
                  cnt=None, syn=None, threads: int = 1, size=None):
        """In-place semantics on copies: returns (data, parity, ok, corrected)."""
        data = np.array(data, np.uint8, order="C", copy=True)
        parity = np.array(parity, np.uint8, order="C", copy=True)
        n = data.shape[0]
        size = data.shape[1] if size is None else size
        p = RsParams(*params)
        ok = np.zeros(n, np.uint8)
        cor = np.zeros(n, np.uint32)
        pos_stride = 0
        if mode == MODE_ERASURE:
            pos = np.ascontiguousarray(pos, np.uint32)
            cnt = np.ascontiguousarray(cnt, np.uint32)
            pos_stride = pos.shape[1]
        if mode == MODE_EXTSYN:
            syn = np.ascontiguousarray(syn, np.uint16)
        rc = self._fn("rs_decode_batch")(C.byref(p), mode, _u8p(data), data.shape[1], size, _u8p(parity),
                                         parity.shape[1], n, _u32p(pos), pos_stride, _u32p(cnt), _u16p(syn),
                                         _u8p(ok), _u32p(cor), threads)
        if rc != 0:
            return None
        return data, parity, ok, cor

    # ------------------------------------------------------------------ BCH
    def bch_info(self, m=4, poly=0x13, t=3):
        out = np.zeros(4, np.uint32)
        rc = self._fn("bch_info")(m, poly, t, _u32p(out))
        return None if rc != 0 else dict(n=int(out[0]), k=int(out[1]), parity_bits=int(out[2]), gen=int(out[3]))

    def bch_encode(self, data: np.ndarray, pbytes: int, m=4, poly=0x13, t=3, threads: int = 1):
        data = np.ascontiguousarray(data, np.uint8)
        n, size = data.shape
        parity = np.zeros((n, pbytes), np.uint8)
        ok = np.zeros(n, np.uint8)
        rc = self._fn("bch_encode_batch")(m, poly, t, _u8p(data), size, size, _u8p(parity), pbytes, n, _u8p(ok),
                                          threads)
        return None if rc != 0 else (parity, ok)

    def bch_decode(self, data: np.ndarray, parity: np.ndarray, m=4, poly=0x13, t=3, threads: int = 1):
        data = np.array(data, np.uint8, order="C", copy=True)
        parity = np.ascontiguousarray(parity, np.uint8)
        n, size = data.shape
        ok = np.zeros(n, np.uint8)
        cor = np.zeros(n, np.uint32)
        rc = self._fn("bch_decode_batch")(m, poly, t, _u8p(data), size, size, _u8p(parity), parity.shape[1], n,
                                          _u8p(ok), _u32p(cor), threads)
        return None if rc != 0 else (data, ok, cor)

    def gf_mod(self, m, poly, v):
        return int(self._fn("gf_mod")(m, poly, v))


def have_reference() -> bool:
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libpoporon_ref.so"))
