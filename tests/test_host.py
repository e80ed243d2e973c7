"""CPU tests (-m "not gpu") of the C-ABI libraries' host logic: exported symbols, constructors and their
limits (mirroring reference tests/test_gf.c, test_rs.c, test_erasure.c, test_bch.c:29-59, test_invalid.c,
test_basic.c).  No codeword is computed here: that needs a GPU and there is no CPU fallback."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import libpoporon_b200.api as api
from libpoporon_b200.build import lib_dir

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
host = api.host


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(poporon_[a-z0-9_]+)\s*\(", txt)))


def _exported(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(lib_dir(), lib)], capture_output=True, text=True,
                         check=True).stdout
    return {ln.split()[-1] for ln in out.splitlines() if " T " in ln}


def test_host_library_exports_every_declared_symbol():
    exp = _exported("libpoporon.so")
    for hdr in ("poporon.h", "poporon/erasure.h", "poporon/gf.h", "poporon/rng.h", "poporon/batch.h"):
        names = _declared(hdr)
        assert names, hdr
        missing = [n for n in names if n not in exp]
        assert not missing, (hdr, missing)


def test_cuda_library_exports_every_declared_symbol():
    exp = _exported("libpoporon_cuda.so")
    names = _declared("poporon_cuda.h")
    assert len(names) > 25
    missing = [n for n in names if n not in exp]
    assert not missing, missing


def test_cuda_library_contains_sm100a_code():
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", os.path.join(lib_dir(), "libpoporon_cuda.so")],
                         capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_version_and_buildtime():
    assert host.poporon_version_id() == 20000000  # reference tests/test_basic.c
    assert host.poporon_buildtime() > 0


def test_invalid_create():
    assert host.poporon_create(None) is None  # reference tests/test_invalid.c
    host.poporon_destroy(None)
    host.poporon_config_destroy(None)
    assert host.poporon_get_fec_type(None) == api.FEC_UNKNOWN
    assert host.poporon_get_parity_size(None) == 0 and host.poporon_get_info_size(None) == 0
    assert host.poporon_get_iterations_used(None) == 0
    assert not host.poporon_encode(None, None, 0, None)
    assert not host.poporon_decode(None, None, 0, None, None)


def test_ldpc_and_rng_are_link_only():
    assert host.poporon_config_ldpc_default(1024, 1) is None
    assert host.poporon_config_ldpc_burst_resistant(1024, 1) is None
    assert host.poporon_ldpc_config_create(1024, 1, 1, 3, True, True, True, 0, 0, 0, None, 0, 0) is None
    assert host.poporon_rng_create(0, None, 0) is None
    assert not host.poporon_rng_next(None, None, 0)
    host.poporon_rng_destroy(None)


def test_gf_create_and_mod():  # reference tests/test_gf.c:31-79
    assert host.poporon_gf_create(0, 0x11D) is None
    assert host.poporon_gf_create(17, 0x11D) is None
    assert host.poporon_gf_create(8, 0x11C) is None  # not primitive
    gf = host.poporon_gf_create(8, 0x11D)
    assert gf
    assert host.poporon_gf_mod(gf, 0) == 0 and host.poporon_gf_mod(gf, 1) == 1
    assert host.poporon_gf_mod(gf, 256) == 1 and host.poporon_gf_mod(gf, 257) == 2
    for i in range(255):
        assert host.poporon_gf_mod(gf, i) == i
    host.poporon_gf_destroy(gf)
    host.poporon_gf_destroy(None)


class _Gf(C.Structure):
    _fields_ = [("symbol_size", C.c_uint8), ("nn", C.c_uint16), ("poly", C.c_uint16), ("alpha_to", C.c_uint16 * 256),
                ("index_of", C.c_uint16 * 256)]


class _Rs(C.Structure):
    _fields_ = [("gf", C.POINTER(_Gf)), ("fcr", C.c_uint16), ("prim", C.c_uint16), ("iprim", C.c_uint16),
                ("nroots", C.c_uint8), ("genpoly", C.c_uint16 * 256)]


@pytest.mark.parametrize("params", [(8, 0x11D, 1, 1, 32), (8, 0x11D, 1, 1, 16), (4, 0x13, 1, 2, 8), (8, 0x11D, 1, 1, 4),
                                    (8, 0x11D, 0, 1, 16), (8, 0x11D, 2, 1, 16), (8, 0x11D, 1, 2, 16),
                                    (8, 0x187, 112, 11, 32), (8, 0x11D, 1, 1, 127), (8, 0x11D, 300, 7, 33)])
def test_rs_create_tables_match_oracle(orc, params):  # reference tests/test_rs.c:36-83
    p = host.poporon_rs_create(*params)
    assert p
    rs = C.cast(p, C.POINTER(_Rs)).contents
    info = orc.rs_info(params)
    nn = (1 << params[0]) - 1
    assert list(rs.genpoly[:params[4] + 1]) == list(info["genpoly"])
    assert list(rs.gf.contents.alpha_to[:nn + 1]) == list(info["alpha_to"])
    assert list(rs.gf.contents.index_of[:nn + 1]) == list(info["index_of"])
    assert rs.iprim == info["iprim"]
    host.poporon_rs_destroy(p)


def test_rs_create_invalid():
    assert host.poporon_rs_create(0, 0x11D, 1, 1, 16) is None
    host.poporon_rs_destroy(None)


def test_erasure_list():  # reference tests/test_erasure.c
    e = host.poporon_erasure_create(32, 0)
    assert e
    for i in range(100):  # grows past the default capacity of num_roots
        assert host.poporon_erasure_add_position(e, i)
    host.poporon_erasure_reset(e)
    assert host.poporon_erasure_add_position(e, 7)
    host.poporon_erasure_destroy(e)
    assert host.poporon_erasure_create_from_positions(32, None, 3) is None
    arr = (C.c_uint32 * 3)(1, 2, 3)
    assert host.poporon_erasure_create_from_positions(32, arr, 0) is None
    e = host.poporon_erasure_create_from_positions(32, arr, 3)
    assert e
    host.poporon_erasure_destroy(e)
    assert not host.poporon_erasure_add_position(None, 1)
    host.poporon_erasure_reset(None)
    host.poporon_erasure_destroy(None)


def test_bch_create_limits_and_lengths(orc):  # reference tests/test_bch.c:29-59
    for (m, poly, t, n) in [(4, 0x13, 3, 15), (5, 0x25, 3, 31), (6, 0x43, 2, 63)]:
        b = host.poporon_bch_create(m, poly, t)
        assert b
        assert host.poporon_bch_get_codeword_length(b) == n
        assert host.poporon_bch_get_correction_capability(b) == t
        info = orc.bch_info(m, poly, t)
        assert host.poporon_bch_get_data_length(b) == info["k"]
        host.poporon_bch_destroy(b)
    assert host.poporon_bch_create(2, 0x07, 1) is None
    assert host.poporon_bch_create(17, 0x0001, 1) is None
    assert host.poporon_bch_create(4, 0x13, 0) is None
    assert host.poporon_bch_create(4, 0x13, 17) is None
    host.poporon_bch_destroy(None)
    assert host.poporon_bch_get_codeword_length(None) == 0


def test_config_objects_are_plain_allocations():
    for mk in (host.poporon_config_rs_default, host.poporon_config_bch_default):
        cfg = mk()
        assert cfg
        host.poporon_config_destroy(cfg)


def test_create_fails_loudly_without_gpu():
    if api.cuda.poporon_cuda_device_count() > 0:
        pytest.skip("a GPU is present")
    cfg = host.poporon_config_rs_default()
    assert host.poporon_create(cfg) is None  # no CPU fallback
    host.poporon_config_destroy(cfg)
    assert b"cuda" in host.poporon_last_error().lower()
    with pytest.raises(RuntimeError):
        api.Poporon.rs_default()
