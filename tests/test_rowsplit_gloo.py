"""N > 1 host logic on CPU: two gloo ranks shard the rows of a matmul with the package's partition rule, compute their
shards with the ORACLE (test infrastructure — stands in for the GPU kernel here), all-gather the slices and must
reproduce the unsplit result bit for bit.  Also pins the partition rule against get_row_split's properties."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, str(ROOT))
    from oracle import refenv
    from tests.conftest import load_package
    from tests.util import oracle_mul_mat, random_blocks
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pkg = load_package()
        orc = refenv.load_oracle()
        rng = np.random.default_rng(5)                        # same matrix on both ranks
        m, k, n = 1000, 512, 3
        W = random_blocks(refenv.Q4_K, m, k, rng)
        X = rng.uniform(-1, 1, (n, k)).astype(np.float32)
        row0, rows = pkg.row_shard(m, rank, world)
        part = oracle_mul_mat(orc, refenv.Q4_K, W[row0:row0 + rows], X, threads=1)          # [n, rows]
        sizes = [pkg.row_shard(m, r, world)[1] for r in range(world)]
        pad = np.zeros((n, max(sizes)), np.float32)            # gloo gathers equal-sized tensors: pad, then trim
        pad[:, :rows] = part
        out = [torch.empty((n, max(sizes)), dtype=torch.float32) for _ in sizes]
        dist.all_gather(out, torch.from_numpy(pad))
        full = torch.cat([o[:, :s] for o, s in zip(out, sizes)], dim=1).numpy()
        want = oracle_mul_mat(orc, refenv.Q4_K, W, X, threads=1)
        q.put((rank, bool(np.array_equal(full, want)), row0, rows))
    finally:
        dist.destroy_process_group()


def test_two_rank_row_split_reproduces_unsplit_result():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29650 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=60) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _, _ in res)
    assert res[0][2] == 0 and res[0][3] == 384 and res[1][2] == 384 and res[1][3] == 616     # cut rounded down to 128 rows


def test_partition_rule():
    sys.path.insert(0, str(ROOT))
    from tests.conftest import load_package
    pkg = load_package()
    for m in (128256, 4096, 1024, 14336, 1000, 127):
        for world in (1, 2, 4, 8):
            prev = 0
            for r in range(world):
                lo, rows = pkg.row_shard(m, r, world)
                assert lo == prev and rows >= 0 and (lo % 128 == 0)
                prev = lo + rows
            assert prev == m
