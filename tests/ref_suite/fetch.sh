#!/bin/bash
# Copies the reference's own Unity test programs (test vectors held by the reference: RS/BCH/GF/erasure/codec/unified/
# basic/invalid) and the Unity framework into tests/_ref_tests/ (git-ignored, travels to the GPU box with gpurun).
# Runs only where /root/reference exists; the sources are used UNMODIFIED and never enter the history.
set -e
REF=${1:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/../_ref_tests
[ -d "$REF/tests" ] || { echo "no reference tree at $REF"; exit 0; }
mkdir -p "$OUT/unity" "$OUT/src/internal" "$OUT/tests"
for t in test_basic test_bch test_codec test_erasure test_gf test_invalid test_rs test_unified; do
  cp "$REF/tests/$t.c" "$OUT/tests/"
done
cp "$REF/tests/util.h" "$OUT/tests/"
cp "$REF/third_party/unity/src/unity.c" "$REF/third_party/unity/src/unity.h" "$REF/third_party/unity/src/unity_internals.h" "$OUT/unity/"
# test_rs.c / test_bch.c include the reference's internal prototypes (poporon_rs_create, poporon_bch_encode, ...)
# (some as "internal/common.h" with -Isrc, test_basic.c as "../src/internal/common.h")
cp "$REF/src/internal/common.h" "$REF/src/internal/config.h" "$OUT/src/internal/"
echo "reference tests copied to $OUT"
