"""CPU test of the generated XOR networks of the RS(255,223) syndrome transform
(libpoporon_b200/csrc/cuda/rs255_fft_gen.cuh, written by tools/gen_rs255_fft.py).

The header is plain C++ on the host; tests/native/host_fft.cpp runs the whole transform (basis change, 15-point
DFTs, Horner stage, encoder solve) on 32 codewords in bit-plane form.  Its syndromes are compared with the direct
formula and its parity with the oracle's rs_encode (reference src/encode.c:120-144)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
from oraclelib import Oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "native", "host_fft.cpp")
INC = os.path.join(ROOT, "libpoporon_b200", "csrc", "cuda")
OUT = os.path.join(ROOT, "build", "tests", "host_fft.so")


@pytest.fixture(scope="module")
def fft():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", f"-I{INC}", "-o", OUT, SRC], check=True)
    lib = C.CDLL(OUT)
    P = C.POINTER
    lib.host_fft_syndromes.argtypes = [P(C.c_uint8), C.c_int, P(C.c_uint8)]
    lib.host_fft_encode.argtypes = [P(C.c_uint8), C.c_int, P(C.c_uint8)]
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


def u8p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


@pytest.mark.parametrize("length", [255, 254, 100, 33])
def test_syndromes_match_direct_formula(fft, length):
    rng = np.random.default_rng(length)
    cw = rng.integers(0, 256, (32, length), dtype=np.uint8)
    cw[3] = 0
    cw[5, :] = 0
    cw[5, length - 1] = 1
    got = np.zeros((32, 32), np.uint8)
    fft.host_fft_syndromes(u8p(cw), length, u8p(got))
    assert (got == syndromes(cw)).all()


@pytest.mark.parametrize("size", [223, 64, 1])
def test_encoder_solve_matches_oracle(fft, size):
    orc = Oracle("port")
    rng = np.random.default_rng(size)
    msg = rng.integers(0, 256, (32, size), dtype=np.uint8)
    got = np.zeros((32, 32), np.uint8)
    fft.host_fft_encode(u8p(msg), size, u8p(got))
    assert (got == orc.rs_encode(msg)).all()


def test_generated_header_is_current():
    """the committed header is what the committed generator writes"""
    tmp = os.path.join(ROOT, "build", "tests", "rs255_fft_gen.cuh")
    os.makedirs(os.path.dirname(tmp), exist_ok=True)
    subprocess.run(["python", os.path.join(ROOT, "tools", "gen_rs255_fft.py"), "--out", tmp], check=True,
                   stderr=subprocess.DEVNULL)
    assert open(tmp).read() == open(os.path.join(INC, "rs255_fft_gen.cuh")).read()
