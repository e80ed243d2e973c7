"""BASELINE.json's full-size configurations through size-independent properties (no oracle at these sizes):
encode -> corrupt -> decode round trips, idempotence of a second decode, exact corrected counts.  The checks live
in bench_configs.py (each runner asserts them before it prints its line); bench.py's own gate covers configs[1]."""
import argparse
import json

import pytest

pytestmark = pytest.mark.gpu


def _run(which, tmp_path, ncw=0):
    import torch

    import bench_configs
    import libpoporon_b200.api as api

    out = tmp_path / "line.jsonl"
    args = argparse.Namespace(ncw=ncw, reps=1, out=str(out))
    torch.cuda.set_device(0)
    runner = {"erasure": bench_configs.run_erasure, "bch": bench_configs.run_bch, "load": bench_configs.run_load}[which]
    runner(args, torch, api, torch.device("cuda", 0))
    torch.cuda.empty_cache()
    return [json.loads(l) for l in open(out)]


def test_config2_erasure_16mi(gpu_device, tmp_path):
    (line,) = _run("erasure", tmp_path)
    assert line["codewords"] == 16 * 1024 * 1024
    assert line["restored_fraction"] > 0.895  # every row without unknown errors is restored (asserted inside too)
    # 256 Ki sampled rows, quirk rows included, byte-compared with the CPU reference
    assert line["verify"]["mismatches"] == 0 and line["verify"]["codewords"] >= 250000


def test_config3_bch_64mi(gpu_device, tmp_path):
    (line,) = _run("bch", tmp_path)
    assert line["codewords"] == 64 * 1024 * 1024
    assert 0.30 < line["ok_fraction_of_4_bit_errors"] < 0.45  # miscorrection rate of 4-bit errors, as in the reference


def test_config4_load_sweep_128mi(gpu_device, tmp_path):
    lines = _run("load", tmp_path)
    assert [l["codewords"] for l in lines] == [128 * 1024 * 1024] * 3
    assert lines[0]["ok_fraction"] == 1.0 and lines[0]["mean_corrected"] == 0.0
    assert lines[1]["ok_fraction"] == 1.0 and lines[1]["restored_fraction"] == 1.0
    assert 0.80 < lines[2]["ok_fraction"] < 0.90  # P(more than 16 of 255 symbols hit at 5 %) is about 14 %
    assert lines[2]["ok_fraction"] == lines[2]["restored_fraction"]
    assert all(l["verify"]["mismatches"] == 0 for l in lines)
    assert 0.80 < lines[2]["verify"]["oracle_ok_fraction"] < 0.90  # the sample holds the failures too
