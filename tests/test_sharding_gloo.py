"""N > 1 host logic on CPU: world_size-2 (and 4) `gloo` process groups run the 2-D partition of SURVEY.md 8(e) with the
oracle standing in for the per-rank kernel.  Pins (i) the panel/block geometry of guda_b200/sharding.py, (ii) that a rank's
block computed from its panels with full K is BIT-IDENTICAL to the same elements of the single-process product (the
property that makes N-GPU results equal 1-GPU results), (iii) the all-gather placement, (iv) sharded DOT = all-reduce of
per-rank partials within ReduceArchTolerance."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, m, n, k, ret):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from guda_b200 import sharding
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pr, pc = sharding.grid_for(world)
        r0, rows, c0, cols = sharding.block_extents(rank, pr, pc, m, n)
        a = O.uniform_f32(m * k, 11).reshape(m, k)
        b = O.uniform_f32(k * n, 12).reshape(k, n)
        a_panel = np.ascontiguousarray(a[r0:r0 + rows, :])            # m_loc x k, lda = k
        b_panel = np.ascontiguousarray(b[:, c0:c0 + cols])            # k x n_loc, ldb = n_loc
        blk = np.zeros(rows * cols, dtype=np.float32)
        O.sgemm(0, 0, rows, cols, k, 1.5, a_panel.ravel(), k, b_panel.ravel(), cols, 0.0, blk, cols)
        gathered = [torch.empty(rows * cols, dtype=torch.float32) for _ in range(world)]
        dist.all_gather(gathered, torch.from_numpy(blk))
        full = sharding.place_blocks([g.numpy() for g in gathered], pr, pc, m, n, np.empty((m, n), dtype=np.float32))
        want = O.sgemm(0, 0, m, n, k, 1.5, a.ravel(), k, b.ravel(), n, 0.0, np.zeros(m * n, dtype=np.float32), n).reshape(m, n)
        ok_gemm = np.array_equal(full, want)
        # sharded DOT: contiguous chunks, per-rank partial, all-reduce(sum)
        nvec = 100003
        x, y = O.uniform_f32(nvec, 21), O.uniform_f32(nvec, 22)
        lo, hi = sharding.dot_chunks(nvec, world)[rank]
        part = torch.tensor([float(O.sdot(hi - lo, x[lo:hi], 1, y[lo:hi], 1))], dtype=torch.float32)
        dist.all_reduce(part, op=dist.ReduceOp.SUM)
        whole = float(O.sdot(nvec, x, 1, y, 1))
        ok_dot = abs(float(part.item()) - whole) <= max(1e-5, 1e-4 * abs(whole))
        # sharded AXPY needs no collective: chunks are disjoint and cover the vector
        chunks = sharding.dot_chunks(nvec, world)
        ok_cover = chunks[0][0] == 0 and chunks[-1][1] == nvec and all(chunks[i][1] == chunks[i + 1][0] for i in range(world - 1))
        ret[rank] = (ok_gemm, ok_dot, ok_cover)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,m,n,k", [(2, 96, 128, 70), (4, 128, 192, 33)])
def test_partitioned_sgemm_and_dot_over_gloo(world, m, n, k):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, m, n, k, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert len(ret) == world
    for r in range(world):
        assert ret[r] == (True, True, True), (r, ret[r])


def test_grid_and_extents():
    from guda_b200 import sharding as S
    assert [S.grid_for(w) for w in (1, 2, 4, 8)] == [(1, 1), (1, 2), (2, 2), (2, 4)]
    assert S.grid_for(6) == (2, 3)
    assert S.rank_coords(5, 2, 4) == (1, 1)
    seen = np.zeros((8, 16), dtype=int)
    for rank in range(8):
        r0, rows, c0, cols = S.block_extents(rank, 2, 4, 8, 16)
        seen[r0:r0 + rows, c0:c0 + cols] += 1
    assert np.all(seen == 1)                       # the blocks tile C exactly once
    with pytest.raises(ValueError):
        S.block_extents(0, 2, 4, 9, 16)
    with pytest.raises(ValueError):
        S.rank_coords(8, 2, 4)
