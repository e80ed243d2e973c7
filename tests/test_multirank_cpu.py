"""The N > 1 plumbing of bench.py on CPU: two processes, torch.distributed with the gloo backend (rendezvous on
127.0.0.1).  The data path has no collective - codewords are independent, a rank takes a contiguous slice of the global
codeword index space (bench.shard_of) - so what is covered here is exactly what the ranks share: disjoint slices that
tile the index space, identical results whether a batch is processed whole or slice by slice (checked with the CPU
oracle), the barrier and the max-over-ranks reduction of the timings (bench.max_over_ranks)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ncw, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist

    import bench
    from oraclelib import Oracle

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        i0, cnt = bench.shard_of(rank, ncw)
        # the same deterministic workload on every rank; each takes its slice
        msg, e, order, mags = bench.cpu_workload(world * ncw, 16, seed=77)
        sl = slice(i0, i0 + cnt)
        orc = Oracle("port")
        par = orc.rs_encode(msg[sl])
        cw = np.concatenate([msg[sl], par], 1)
        for i in range(cnt):
            k = int(e[sl][i])
            cw[i, order[sl][i, :k]] ^= mags[sl][i, :k]
        d, p, ok, cor = orc.rs_decode(cw[:, :223], cw[:, 223:])
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), i0=i0, par=par, d=d, p=p, ok=ok, cor=cor)
        dist.barrier()
        # every rank reports a different "time": the job's time is the slowest rank's
        got = bench.max_over_ranks([10.0 + rank, 5.0 - rank, 1.0])
        assert got == [10.0 + world - 1, 5.0, 1.0], got
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo(tmp_path):
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp

    import bench
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oraclelib import Oracle

    world, ncw = 2, 1500
    mp.spawn(_worker, args=(world, _free_port(), ncw, str(tmp_path)), nprocs=world, join=True)
    # the slices tile the index space ...
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    assert [int(p_["i0"]) for p_ in parts] == [0, ncw]
    # ... and slice-by-slice equals the whole batch in one process
    msg, e, order, mags = bench.cpu_workload(world * ncw, 16, seed=77)
    orc = Oracle("port")
    par = orc.rs_encode(msg)
    cw = np.concatenate([msg, par], 1)
    for i in range(world * ncw):
        k = int(e[i])
        cw[i, order[i, :k]] ^= mags[i, :k]
    d, p, ok, cor = orc.rs_decode(cw[:, :223], cw[:, 223:])
    for name, whole in (("par", par), ("d", d), ("p", p), ("ok", ok), ("cor", cor)):
        assert (np.concatenate([p_[name] for p_ in parts]) == whole).all(), name
    assert ok.all() and (d == msg).all()


def test_max_over_ranks_single_process():
    import bench

    assert bench.max_over_ranks([3.0, 1.5]) == [3.0, 1.5]
