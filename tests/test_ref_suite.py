"""The reference's OWN Unity test programs (tests/test_*.c of colopl/libpoporon, unmodified), compiled against this
repo's include/ and linked with libpoporon.so: INTEGRATION.md section 1 ("a program written against poporon.h
recompiles unchanged") as a test.  tests/ref_suite/fetch.sh copies the sources into the git-ignored tests/_ref_tests/
where /root/reference exists (build() runs it); nothing here reads /root/reference.

LDPC is outside the hot path: the LDPC cases of test_unified.c are expected to fail and are the
only failures tolerated."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "_ref_tests")
BIN = os.path.join(SRC, "bin")
LINE = re.compile(r"^[^:]+:\d+:(\w+):(PASS|FAIL|IGNORE)")

needs_suite = pytest.mark.skipif(not os.path.isdir(os.path.join(SRC, "tests")),
                                 reason="tests/_ref_tests not populated (tests/ref_suite/fetch.sh needs /root/reference)")


def _build():
    subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "ref_suite")], check=True, stdout=subprocess.DEVNULL)


def _run(name):
    _build()
    p = subprocess.run([os.path.join(BIN, name)], capture_output=True, text=True, timeout=600)
    res = {}
    for line in p.stdout.splitlines():
        m = LINE.match(line.strip())
        if m:
            res[m.group(1)] = m.group(2)
    assert res, f"{name}: no Unity results\n{p.stdout[-2000:]}\n{p.stderr[-2000:]}"
    return res, p.stdout


@needs_suite
@pytest.mark.parametrize("name", ["test_gf", "test_basic", "test_invalid"])
def test_reference_programs_without_device(name):
    """object construction, getters, GF tables, argument validation: no kernel runs"""
    res, out = _run(name)
    bad = {k: v for k, v in res.items() if v == "FAIL"}
    assert not bad, f"{name}: {bad}\n{out[-3000:]}"


@needs_suite
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["test_rs", "test_bch", "test_codec", "test_erasure", "test_unified"])
def test_reference_programs_on_gpu(gpu_device, name):
    res, out = _run(name)
    bad = {k: v for k, v in res.items() if v == "FAIL" and "ldpc" not in k.lower()}
    assert not bad, f"{name}: {bad}\n{out[-3000:]}"
    assert sum(v == "PASS" for v in res.values()) >= 1
