"""The multi-GPU split north_star describes (one process, one host thread per device inside the library, contiguous
shards, no collective: csrc/host/pipeline.c pp_run_host_job), on real hardware.  Skips unless the box has >= 2
devices: run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`."""
import numpy as np
import pytest

import libpoporon_b200.api as api

pytestmark = pytest.mark.gpu


def _devices():
    return api.cuda.poporon_cuda_device_count()


@pytest.mark.parametrize("layout", ["split", "packed"])
def test_host_batch_over_all_devices(gpu_device, orc, layout):
    ndev = _devices()
    if ndev < 2:
        pytest.skip("needs >= 2 CUDA devices")
    rng = np.random.default_rng(5)
    n = 200000 * ndev + 12345  # ragged: the last lane's shard is shorter, and not a multiple of the tile
    msg = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        assert h.use_devices(None)  # all visible devices
        par = np.zeros((n, 32), np.uint8)
        assert h.encode_batch(msg, 223, 223, par, 32, n)
        assert (par == orc.rs_encode(msg, threads=8)).all()
        cw = np.concatenate([msg, par], 1)
        nerr = rng.integers(0, 20, n)
        order = np.argsort(rng.random((n, 255)), axis=1)[:, :19]
        hit = np.arange(19)[None, :] < nerr[:, None]
        rows = np.repeat(np.arange(n), 19).reshape(n, 19)
        cw[rows[hit], order[hit]] ^= rng.integers(1, 256, int(hit.sum()), dtype=np.uint8)
        want = orc.rs_decode(cw[:, :223], cw[:, 223:], threads=8)
        ok = np.zeros(n, np.uint8)
        cor = np.zeros(n, np.uint32)
        if layout == "split":
            d, p = np.ascontiguousarray(cw[:, :223]), np.ascontiguousarray(cw[:, 223:])
            assert h.decode_batch(d, 223, 223, p, 32, n, ok, cor)
        else:
            rec = cw.copy()
            assert h.decode_batch(rec.ctypes.data, 255, 223, rec.ctypes.data + 223, 255, n, ok, cor)
            d, p = rec[:, :223], rec[:, 223:]
        assert (ok == want[2]).all() and (cor == want[3]).all() and (d == want[0]).all() and (p == want[1]).all()
        # every device took part: the handle's launch count grows on each lane; cheap proxy: many more launches than
        # one device's chunk count
        assert h.launches >= 2 * ndev


def test_erasure_and_bch_over_all_devices(gpu_device, orc):
    ndev = _devices()
    if ndev < 2:
        pytest.skip("needs >= 2 CUDA devices")
    rng = np.random.default_rng(6)
    n = 70000 * ndev
    msg = rng.integers(0, 256, (n, 223), dtype=np.uint8)
    with api.Poporon.rs_default() as h:
        assert h.use_devices(None)
        par = np.zeros((n, 32), np.uint8)
        assert h.encode_batch(msg, 223, 223, par, 32, n)
        cw = np.concatenate([msg, par], 1)
        pos = np.zeros((n, 32), np.uint32)
        cnt = rng.integers(0, 33, n).astype(np.uint32)
        for i in range(0, n, 7):  # a seventh of the rows: erasures at sorted message positions
            c = int(cnt[i])
            pp = np.sort(rng.choice(223, c, replace=False))
            pos[i, :c] = pp
            cw[i, pp] ^= rng.integers(0, 256, c, dtype=np.uint8)
        cnt[np.arange(n) % 7 != 0] = 0
        want = orc.rs_decode(cw[:, :223], cw[:, 223:], mode=1, pos=pos, cnt=cnt, threads=8)
        d, p = np.ascontiguousarray(cw[:, :223]), np.ascontiguousarray(cw[:, 223:])
        ok = np.zeros(n, np.uint8)
        cor = np.zeros(n, np.uint32)
        assert h.decode_batch_erasure(d, 223, 223, p, 32, n, pos, 32, cnt, ok, cor)
        assert (ok == want[2]).all() and (cor == want[3]).all() and (d == want[0]).all() and (p == want[1]).all()
    with api.Poporon.bch_default() as h:
        assert h.use_devices(None)
        m = 300000 * ndev
        data = rng.integers(0, 32, (m, 1), dtype=np.uint8)
        par = np.zeros((m, 2), np.uint8)
        assert h.encode_batch(data, 1, 1, par, 2, m)
        assert (par == orc.bch_encode(data, 2, threads=8)[0]).all()
