"""Row-block sharded C = A*B over the GPUs of one node (BASELINE configs[3]).

One process per GPU (torchrun). Rank g owns rows [r0, r0+rows) of A and of C and
needs all of B: B lives on `src` and is broadcast by NCCL over NVLink — the path's
one exchange step. There is no reduction: the result blocks are disjoint.

The broadcast is cut into k-chunks (row slabs of B). Chunk i+1 travels on NCCL's
stream while the tensor cores multiply chunk i,
    C_block (^)= A_block[:, k_i] * B[k_i, :],
so only the first chunk's transfer is exposed. The per-chunk product is a
callable: on the GPU it is libm4ri_b200's device product on windows of the
resident operands; the CPU tests (gloo, world_size 2) pass the oracle instead, so
the partition, the chunk plan and the broadcast plumbing are covered without a GPU.
"""
import torch


def row_block(m, world, rank, align=256):
    """Contiguous row-block of rank `rank`: (first row, number of rows). Blocks are
    multiples of `align` rows (the 256-row tile of the tensor kernel) except the last
    non-empty one; trailing ranks may own nothing when m is small."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    per = -(-m // world)
    per = -(-per // align) * align
    r0 = min(m, rank * per)
    return r0, min(m, r0 + per) - r0


def k_chunks(k, chunks, align=1024):
    """Cut [0, k) into at most `chunks` slabs whose boundaries are multiples of `align`
    bits (128-bit aligned column windows of A, whole TMA boxes)."""
    chunks = max(1, int(chunks))
    per = -(-k // chunks)
    per = max(align, -(-per // align) * align)
    out, k0 = [], 0
    while k0 < k:
        out.append((k0, min(k, k0 + per)))
        k0 += per
    return out or [(0, 0)]


class ShardedMul:
    """Holds the plan of one sharded product and runs steps of it.

    tA, tC: this rank's row-blocks as int64 tensors [rows, stride_words];
    tB: the full k x n operand [k, stride_words] (valid on `src`, overwritten elsewhere).
    chunk_product(i, k0, k1, clear) computes C_block (^)= A_block[:, k0:k1] * B[k0:k1, :]
    on the current stream/device."""

    def __init__(self, m, k, n, dist=None, src=0, chunks=4, chunk_product=None, tB=None):
        self.m, self.k, self.n = m, k, n
        self.dist = dist
        self.src = src
        self.world = dist.get_world_size() if dist is not None else 1
        self.rank = dist.get_rank() if dist is not None else 0
        self.r0, self.rows = row_block(m, self.world, self.rank)
        self.plan = k_chunks(k, chunks if self.world > 1 else 1)
        self.chunk_product = chunk_product
        self.tB = tB

    def step(self):
        works = []
        if self.dist is not None and self.world > 1:
            for (k0, k1) in self.plan:  # all chunks queued at once: they serialise on NCCL's stream
                works.append(self.dist.broadcast(self.tB[k0:k1], src=self.src, async_op=True))
        for i, (k0, k1) in enumerate(self.plan):
            if works:
                works[i].wait()  # the compute stream waits for chunk i only
            if self.rows > 0:
                self.chunk_product(i, k0, k1, i == 0)


def gpu_chunk_product(m4, tA, tB, tC, rows, k, n, algo=0):
    """chunk_product for resident torch tensors on the current CUDA device: windows of
    A_block (columns k0:k1) and of B (rows k0:k1) wrapped as device matrices, product by
    libm4ri_b200 on torch's current stream."""
    sa, sb, sc = tA.stride(0), tB.stride(0), tC.stride(0)
    dC = m4.DMat.wrap(tC.data_ptr(), rows, n, sc, keepalive=tC)
    cache = {}

    def product(i, k0, k1, clear):
        if i not in cache:
            if k0 % 128:
                raise ValueError("k-chunk boundaries must be multiples of 128 bits")
            dA = m4.DMat.wrap(tA.data_ptr() + (k0 // 64) * 8, rows, k1 - k0, sa, keepalive=tA)
            dB = m4.DMat.wrap(tB.data_ptr() + k0 * sb * 8, k1 - k0, n, sb, keepalive=tB)
            cache[i] = (dA, dB)
        dA, dB = cache[i]
        m4.dmul(dC, dA, dB, clear=clear, algo=algo)

    return product
