#!/usr/bin/env python
"""Multi-GPU parity + timing of guda_sgemm_sharded / guda_sdot_sharded (one rank per GPU, launched by torchrun):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py

Every rank regenerates the full A and B from the counter-based device RNG, computes the full product on its own GPU with
the single-GPU kernel, and checks that the gathered c_full of BOTH gather routes (P2P-fused epilogue stores, NCCL all-gather)
is bit-identical to it.  Then times both routes at --size.  Not collected by pytest (needs N > 1 GPUs)."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check-size", type=int, default=2048)
    ap.add_argument("--size", type=int, default=16384)
    ap.add_argument("--iters", type=int, default=3)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import guda_b200 as g
    from guda_b200.sharding import block_extents, grid_for
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g.Init(local)
    L = g.load()
    ident = (C.c_uint8 * 128)()
    if rank == 0:
        g.capi.check(L.guda_comm_unique_id(ident), "CommUniqueId")
    obj = [bytes(ident)]
    dist.broadcast_object_list(obj, src=0)
    ident = (C.c_uint8 * 128).from_buffer_copy(obj[0])
    comm = C.c_void_p()
    g.capi.check(L.guda_comm_init_rank(C.byref(comm), world, rank, ident), "CommInitRank")
    pr, pc = grid_for(world)
    out = {"world": world, "grid": [pr, pc]}

    def run(n, mode, iters, check):
        g.capi.check(L.guda_comm_set_gather_mode(comm, mode), "SetGatherMode")
        r0, m_loc, c0, n_loc = block_extents(rank, pr, pc, n, n)
        a_panel, b_panel = g.Malloc(m_loc * n * 4), g.Malloc(n * n_loc * 4)
        c_block, c_full = g.Malloc(m_loc * n_loc * 4), g.Malloc(n * n * 4)
        g.FillUniform(a_panel, m_loc, n, n, seed=11, row0=r0, col0=0, full_cols=n)
        g.FillUniform(b_panel, n, n_loc, n_loc, seed=12, row0=0, col0=c0, full_cols=n)
        g.Memset(c_full, 0xFF, n * n * 4)

        def step():
            g.capi.check(L.guda_sgemm_sharded(comm, pr, pc, n, n, n, 1.0, a_panel.ptr, n, b_panel.ptr, n_loc, c_block.ptr, n_loc,
                                              c_full.ptr, n, None), "SgemmSharded")
        step()
        g.Synchronize()
        eff = C.c_int(-1)
        L.guda_comm_get_gather_mode(comm, C.byref(eff))
        res = {"mode_requested": mode, "mode_in_effect": eff.value}
        if check:
            a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
            g.FillUniform(a, n, n, n, seed=11); g.FillUniform(b, n, n, n, seed=12)
            g.GEMM(False, False, n, n, n, 1.0, a, n, b, n, 0.0, c, n)
            g.Synchronize()
            res["bit_identical_to_one_gpu"] = bool(np.array_equal(c_full.Float32(), c.Float32()))
            for p in (a, b, c):
                g.Free(p)
        else:
            dist.barrier(); g.Synchronize()
            e0, e1 = g.Event(), g.Event()
            e0.Record()
            for _ in range(iters):
                step()
            e1.Record(); e1.Synchronize()
            ms = torch.tensor([g.ElapsedTime(e0, e1) / iters], dtype=torch.float64, device="cuda")
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            res["ms"] = float(ms.item()); res["tflops"] = 2.0 * n ** 3 / res["ms"] / 1e9
        for p in (a_panel, b_panel, c_block, c_full):
            g.Free(p)
        return res

    def run_host(n, k, mode):
        """guda_sgemm_sharded_host: pinned host panels -> host C block + gathered c_full; k spans several 8192-deep chunks."""
        g.capi.check(L.guda_comm_set_gather_mode(comm, mode), "SetGatherMode")
        r0, m_loc, c0, n_loc = block_extents(rank, pr, pc, n, n)
        a_panel, b_panel = g.Malloc(m_loc * k * 4), g.Malloc(k * n_loc * 4)
        c_block, c_full = g.Malloc(m_loc * n_loc * 4), g.Malloc(n * n * 4)
        g.FillUniform(a_panel, m_loc, k, k, seed=21, row0=r0, col0=0, full_cols=k)
        g.FillUniform(b_panel, k, n_loc, n_loc, seed=22, row0=0, col0=c0, full_cols=n)
        ha, hb, hc = C.c_void_p(), C.c_void_p(), C.c_void_p()
        for h, nb in ((ha, m_loc * k * 4), (hb, k * n_loc * 4), (hc, m_loc * n_loc * 4)):
            g.capi.check(L.guda_malloc_host(C.byref(h), nb), "MallocHost")
        L.guda_memcpy(ha, a_panel.ptr, m_loc * k * 4, 2, None); L.guda_memcpy(hb, b_panel.ptr, k * n_loc * 4, 2, None)
        g.Memset(a_panel, 0xFF, m_loc * k * 4); g.Memset(b_panel, 0xFF, k * n_loc * 4); g.Memset(c_full, 0xFF, n * n * 4)
        g.capi.check(L.guda_sgemm_sharded_host(comm, pr, pc, n, n, k, 1.0, ha, k, hb, n_loc, hc, n_loc, a_panel.ptr, b_panel.ptr,
                                               c_block.ptr, c_full.ptr, n, None), "SgemmShardedHost")
        a, b, c = g.Malloc(n * k * 4), g.Malloc(k * n * 4), g.Malloc(n * n * 4)
        g.FillUniform(a, n, k, k, seed=21); g.FillUniform(b, k, n, n, seed=22)
        g.GEMM(False, False, n, n, k, 1.0, a, k, b, n, 0.0, c, n)
        g.Synchronize()
        want = c.Float32().reshape(n, n)
        host_block = np.ctypeslib.as_array(C.cast(hc, C.POINTER(C.c_float)), shape=(m_loc, n_loc))
        res = {"gathered_bit_identical": bool(np.array_equal(c_full.Float32().reshape(n, n), want)),
               "host_block_bit_identical": bool(np.array_equal(host_block, want[r0:r0 + m_loc, c0:c0 + n_loc]))}
        for p in (a, b, c, a_panel, b_panel, c_block, c_full):
            g.Free(p)
        for h in (ha, hb, hc):
            L.guda_free_host(h)
        return res

    def run_dgemm(n, mode):
        """guda_dgemm_sharded: gathered c_full on every rank bit-identical to the one-GPU DGEMM (DMMA route) of the same operands."""
        g.capi.check(L.guda_comm_set_gather_mode(comm, mode), "SetGatherMode")
        r0, m_loc, c0, n_loc = block_extents(rank, pr, pc, n, n)
        a_panel, b_panel = g.Malloc(m_loc * n * 8), g.Malloc(n * n_loc * 8)
        c_block, c_full = g.Malloc(m_loc * n_loc * 8), g.Malloc(n * n * 8)
        g.FillUniform(a_panel, m_loc, n, n, seed=41, row0=r0, col0=0, full_cols=n, dtype=np.float64)
        g.FillUniform(b_panel, n, n_loc, n_loc, seed=42, row0=0, col0=c0, full_cols=n, dtype=np.float64)
        g.Memset(c_full, 0xFF, n * n * 8)
        for _ in range(2):                                   # second call takes the cached peer mapping
            g.capi.check(L.guda_dgemm_sharded(comm, pr, pc, n, n, n, 1.0, a_panel.ptr, n, b_panel.ptr, n_loc, c_block.ptr, n_loc,
                                              c_full.ptr, n, None), "DgemmSharded")
        g.Synchronize()
        a, b, c = g.Malloc(n * n * 8), g.Malloc(n * n * 8), g.Malloc(n * n * 8)
        g.FillUniform(a, n, n, n, seed=41, dtype=np.float64); g.FillUniform(b, n, n, n, seed=42, dtype=np.float64)
        g.Dgemm(False, False, n, n, n, 1.0, a, n, b, n, 0.0, c, n)
        g.Synchronize()
        res = {"bit_identical_to_one_gpu": bool(np.array_equal(c_full.Float64(), c.Float64()))}
        for p in (a, b, c, a_panel, b_panel, c_block, c_full):
            g.Free(p)
        return res

    out["check_dgemm_p2p"] = run_dgemm(1024, 0)
    out["check_dgemm_nccl"] = run_dgemm(1024, 1)
    out["check_host_p2p"] = run_host(1024, 2 * 8192 + 300, 0)
    out["check_host_nccl"] = run_host(1024, 2 * 8192 + 300, 1)
    out["check_p2p"] = run(args.check_size, 0, 1, True)
    out["check_nccl"] = run(args.check_size, 1, 1, True)
    out["time_p2p"] = run(args.size, 0, args.iters, False)
    out["time_nccl"] = run(args.size, 1, args.iters, False)
    # sharded DOT vs the one-GPU DOT of the same vector
    nvec = 1 << 24
    lo = rank * (nvec // world)
    x, y = g.Malloc(nvec * 4), g.Malloc(nvec * 4)
    g.FillUniform(x, 1, nvec, nvec, 31); g.FillUniform(y, 1, nvec, nvec, 32)
    whole = g.DOT(x, y, nvec)
    r = C.c_float(0)
    g.capi.check(L.guda_sdot_sharded(comm, nvec // world, x.ptr + lo * 4, y.ptr + lo * 4, C.byref(r), None), "SdotSharded")
    out["sdot"] = {"one_gpu": float(whole), "sharded": float(r.value), "within_reduce_tolerance": bool(abs(r.value - whole) <= max(1e-5, 1e-4 * abs(whole)))}
    # sharded AXPY: this rank's chunk, bit-identical to the same chunk of the one-GPU AXPY
    yy = g.Malloc(nvec * 4); g.FillUniform(yy, 1, nvec, nvec, 32)
    g.AXPY(2.5, x, yy, nvec)
    g.capi.check(L.guda_saxpy_sharded(comm, nvec // world, 2.5, x.ptr + lo * 4, y.ptr + lo * 4, None), "SaxpySharded")
    g.Synchronize()
    out["saxpy"] = {"chunk_bit_identical": bool(np.array_equal(g.DevicePtr(y.ptr + lo * 4, (nvec // world) * 4).Float32(),
                                                                g.DevicePtr(yy.ptr + lo * 4, (nvec // world) * 4).Float32()))}
    ok = torch.tensor([1.0 if (out["check_p2p"]["bit_identical_to_one_gpu"] and out["check_nccl"]["bit_identical_to_one_gpu"]
                               and out["check_dgemm_p2p"]["bit_identical_to_one_gpu"] and out["check_dgemm_nccl"]["bit_identical_to_one_gpu"]
                               and out["saxpy"]["chunk_bit_identical"]
                               and all(out[k][f] for k in ("check_host_p2p", "check_host_nccl")
                                       for f in ("gathered_bit_identical", "host_block_bit_identical"))
                               and out["sdot"]["within_reduce_tolerance"]) else 0.0], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    out["all_ranks_ok"] = bool(ok.item() > 0.5)
    dist.barrier(); g.Synchronize()
    L.guda_comm_destroy(comm)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()
    sys.exit(0 if out["all_ranks_ok"] else 1)


if __name__ == "__main__":
    main()
