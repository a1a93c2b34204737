"""Multi-GPU host logic (m4ri-rust_b200/sharded.py) on CPU: row-block partition, k-chunk
plan, and the chunked NCCL-shaped broadcast of B run under gloo with world_size 2 (and
3, ragged), the per-chunk product being the oracle (test infrastructure standing in for
the CUDA product). The gathered blocks must equal the oracle's full product bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_row_block_partition(m4):
    from m4ri_rust_b200 import sharded
    for m in (1, 255, 256, 257, 1000, 8192, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            blocks = [sharded.row_block(m, world, r) for r in range(world)]
            assert sum(b[1] for b in blocks) == m
            pos = 0
            for r0, rows in blocks:
                assert rows >= 0 and (rows == 0 or r0 == pos)
                pos += rows
            for r0, rows in blocks[:-1]:
                assert rows % 256 == 0 or r0 + rows == m
    assert sharded.row_block(65536, 8, 3) == (3 * 8192, 8192)


def test_k_chunks(m4):
    from m4ri_rust_b200 import sharded
    for k in (1, 1000, 1024, 5000, 65536):
        for c in (1, 2, 4, 7):
            plan = sharded.k_chunks(k, c)
            assert plan[0][0] == 0 and plan[-1][1] == k and len(plan) <= c
            for (a, b), (a2, _) in zip(plan, plan[1:]):
                assert b == a2 and a % 1024 == 0 and b > a
    assert sharded.k_chunks(65536, 4) == [(i * 16384, (i + 1) * 16384) for i in range(4)]


def _worker(rank, world, port, m, k, n, chunks, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import oracle
    from m4ri_rust_b200 import sharded
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        w = oracle.width
        r0, rows = sharded.row_block(m, world, rank)
        A = oracle.random(m, k, 11)[:, : w(k)]
        Ablk = np.ascontiguousarray(A[r0:r0 + rows])
        # B is valid on rank 0 only; the others start from garbage
        B = oracle.random(k, n, 12)[:, : w(n)].copy() if rank == 0 else np.full((k, w(n)), 0x5555555555555555, np.uint64)
        tB = torch.from_numpy(B.view(np.int64))
        Cblk = np.zeros((rows, w(n)), np.uint64)

        def chunk_product(i, k0, k1, clear):
            Bc = np.ascontiguousarray(tB.numpy().view(np.uint64)[k0:k1])
            Ac = np.zeros((rows, w(k1 - k0)), np.uint64)
            Ac[:, :] = Ablk[:, k0 // 64: k0 // 64 + w(k1 - k0)]
            if (k1 - k0) % 64:
                Ac[:, -1] &= np.uint64((1 << ((k1 - k0) % 64)) - 1)
            part = oracle.mul_m4rm(Ac, Bc, k1 - k0, n)[:, : w(n)]
            if clear:
                Cblk[:] = part
            else:
                Cblk[:] ^= part

        sm = sharded.ShardedMul(m, k, n, dist=dist, src=0, chunks=chunks, chunk_product=chunk_product, tB=tB)
        assert (sm.r0, sm.rows) == (r0, rows)
        for _ in range(2):  # two steps: the plan is reusable, the result idempotent
            sm.step()
        np.save(os.path.join(out_dir, f"c{rank}.npy"), Cblk)
        np.save(os.path.join(out_dir, f"b{rank}.npy"), tB.numpy().view(np.uint64))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,m,k,n,chunks", [(2, 600, 2048, 300, 2), (2, 512, 4096, 512, 4), (3, 700, 1500, 130, 4)])
def test_sharded_product_gloo(tmp_path, orc, world, m, k, n, chunks):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, m, k, n, chunks, str(tmp_path)), nprocs=world, join=True)
    w = orc.width
    A, B = orc.random(m, k, 11), orc.random(k, n, 12)
    want = orc.mul_m4rm(A, B, k, n)[:, : w(n)]
    got = np.concatenate([np.load(tmp_path / f"c{r}.npy") for r in range(world)], axis=0)
    assert got.shape == want.shape and (got == want).all()
    for r in range(world):  # every rank ended up with rank 0's B
        assert (np.load(tmp_path / f"b{r}.npy") == B[:, : w(n)]).all()


def test_k_chunks_ramp_and_sm_reservation(m4):
    from m4ri_rust_b200 import sharded
    assert sharded.k_chunks_ramp(65536) == [(0, 8192), (8192, 32768), (32768, 65536)]
    assert sharded.k_chunks_ramp(4096) == [(0, 4096)] and sharded.k_chunks_ramp(8192) == [(0, 8192)]
    for k in (16384, 20000, 70000):
        plan = sharded.k_chunks_ramp(k)
        assert plan[0][0] == 0 and plan[-1][1] == k
        assert all(a % 1024 == 0 and b > a for a, b in plan) and all(p[1] == q[0] for p, q in zip(plan, plan[1:]))

    class FakeWork:
        def wait(self):
            pass

    class FakeDist:  # world of 2 without a process group: records the order of the calls
        def __init__(self, log):
            self.log = log
        def get_world_size(self):
            return 2
        def get_rank(self):
            return 0
        def broadcast(self, t, src, async_op):
            self.log.append(("bcast", t.shape[0]))
            return FakeWork()

    import torch
    log = []
    sm = sharded.ShardedMul(512, 65536, 64, dist=FakeDist(log), chunks="ramp", tB=torch.zeros((65536, 1), dtype=torch.int64),
                            chunk_product=lambda i, k0, k1, clear: log.append(("mul", i, clear)),
                            reserve_sms=lambda r: log.append(("reserve", r)))
    sm.step()
    assert log == [("bcast", 8192), ("bcast", 24576), ("bcast", 32768),
                   ("reserve", 8), ("mul", 0, True), ("reserve", 8), ("mul", 1, False), ("reserve", 0), ("mul", 2, False)]


def test_exchange_slabs_cover_k(m4):
    """Slab plan of the fused exchange: disjoint, ordered, aligned, covering [0, k); from 3 ranks up
    the root only serves reads."""
    from m4ri_rust_b200 import sharded
    for world in (1, 2, 3, 4, 8):
        for src, serves in ((0, False), (0, True), (world - 1, True)):
            for k in (1024, 65536, 70000, 5000, 100):
                slabs = [sharded.exchange_slab(k, world, r, src, root_serves=serves) for r in range(world)]
                if world >= 3 and serves:
                    assert slabs[src] == (0, 0)
                live = sorted(s for s in slabs if s[1] > 0)
                assert live[0][0] == 0 and sum(s[1] for s in live) == k
                assert all(a[0] + a[1] == b[0] for a, b in zip(live, live[1:]))
                assert all(s[0] % 1024 == 0 for s in live)
