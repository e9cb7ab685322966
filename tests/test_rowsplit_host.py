"""CPU test (gloo, world_size 2) of the row-split bookkeeping: every rank computes its row slice of a MUL_MAT with the
oracle, the slices are all-gathered, and the result must equal the oracle on the full matrix bit for bit (no reduction
happens across ranks, so the split may not change a single value)."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(HERE, "..", "oracle"), os.path.join(HERE, "..", "ggml-hexagon_b200")):
    sys.path.insert(0, os.path.abspath(p))


def _worker(rank, world, port, types, out_q):
    import pyoracle as O
    import rowsplit as R
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ok = True
    for t in types:
        rng = np.random.default_rng(5)                      # same seed everywhere: identical full matrix on every rank
        M, K, N = 64, 512, 3
        raw = O.random_blocks(t, M, K, rng)
        x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
        mine = O.mul_mat(t, R.slice_rows(raw, rank, world), x, 1)           # [N, rows]
        parts = [torch.zeros(mine.shape, dtype=torch.float32) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(np.ascontiguousarray(mine)))
        full = np.concatenate([p.numpy() for p in parts], axis=1)            # slices concatenate along M
        ref = O.mul_mat(t, raw, x, 1)
        ok = ok and np.array_equal(full, ref)
        lo, hi = R.row_range(M, rank, world)
        ok = ok and R.dst_offset_bytes(M, rank, world, column=2) == 4 * (2 * M + lo) and hi - lo == M // world
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        out_q.put(int(flag.item()))
    dist.destroy_process_group()


def test_rowsplit_gloo_world2():
    import pyoracle as O
    O.lib()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, [O.Q4_0, O.Q8_0, O.Q4_K, O.Q6_K], q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert q.get(timeout=5) == 1


def test_row_range_rejects_uneven():
    import pytest
    import rowsplit as R
    with pytest.raises(ValueError):
        R.row_range(4097, 0, 2)
    assert R.row_range(8192, 3, 8) == (3072, 4096)
