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

RESERVED_SMS = 8  # SMs left to NCCL while a broadcast overlaps a product (bench: NCCL_MAX_NCHANNELS=8)


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


def k_chunks_ramp(k, align=1024, parts=3):
    """Slabs ending at 1/8, 1/2 and all of k (parts = 3; 1/16, 1/4, 1/2, 1 for parts = 4): the first
    transfer is the only exposed one, so it is the smallest; the later, larger products hide the
    later, larger transfers."""
    if k < 16 * align:
        return k_chunks(k, 1, align)
    cuts = [k // 8, k // 2] if parts == 3 else [k // 16, k // 4, k // 2]
    cuts = [c // align * align for c in cuts]
    return list(zip([0] + cuts, cuts + [k]))


class ShardedMul:
    """Holds the plan of one sharded product and runs steps of it.

    tA, tC: this rank's row-blocks as int64 tensors [rows, stride_words];
    tB: the full k x n operand [k, stride_words] (valid on `src`, overwritten elsewhere).
    chunk_product(i, k0, k1, clear) computes C_block (^)= A_block[:, k0:k1] * B[k0:k1, :]
    on the current stream/device."""

    def __init__(self, m, k, n, dist=None, src=0, chunks=4, chunk_product=None, tB=None,
                 reserve_sms=None):
        self.m, self.k, self.n = m, k, n
        self.dist = dist
        self.src = src
        self.world = dist.get_world_size() if dist is not None else 1
        self.rank = dist.get_rank() if dist is not None else 0
        self.r0, self.rows = row_block(m, self.world, self.rank)
        if self.world == 1:
            self.plan = k_chunks(k, 1)
        elif chunks == "auto":
            # a block tall enough for a Strassen level (leaves of 16384) gains more from it than
            # from hiding the transfer; shorter blocks hide the transfer behind ramped chunks
            self.plan = k_chunks(k, 1) if self.rows >= 32768 and k >= 32768 and n >= 32768 else k_chunks_ramp(k)
        elif chunks == "ramp":
            self.plan = k_chunks_ramp(k)
        elif chunks == "ramp4":
            self.plan = k_chunks_ramp(k, parts=4)
        else:
            self.plan = k_chunks(k, chunks)
        self.chunk_product = chunk_product
        self.tB = tB
        # reserve_sms(n): called with n > 0 before a product that must leave n SMs to the
        # collective still in flight (the persistent product kernels otherwise hold every SM and
        # NCCL's kernel could only start when they end), and with 0 before the last product
        self.reserve_sms = reserve_sms

    def step(self):
        works = []
        if self.dist is not None and self.world > 1:
            for (k0, k1) in self.plan:  # all chunks queued at once: they serialise on NCCL's stream
                works.append(self.dist.broadcast(self.tB[k0:k1], src=self.src, async_op=True))
        for i, (k0, k1) in enumerate(self.plan):
            if works:
                works[i].wait()  # the compute stream waits for chunk i only
            if self.reserve_sms is not None and works:
                self.reserve_sms(0 if i == len(self.plan) - 1 else RESERVED_SMS)
            if self.rows > 0:
                self.chunk_product(i, k0, k1, i == 0)


def release_stream(m4):
    """Hand the calling thread's library work back to the library's own stream.

    exchange() / step() and the product of gpu_chunk_product() bind the thread's library stream to
    torch's current stream (m4ri_b200_set_stream) and leave it bound: the next step needs the same
    binding, and the caller's own device-API calls between steps (checksums, Freivalds) must stay
    ordered behind the product. The binding outlives the step, so a caller that is about to destroy
    that torch stream, or that goes on to use the host ABI (mzd_mul ...) expecting the library's own
    stream, calls this first."""
    m4.lib().m4ri_b200_use_own_stream()


def gpu_chunk_product(m4, tA, tB, tC, rows, k, n, algo=0):
    """chunk_product for resident torch tensors on the current CUDA device: windows of
    A_block (columns k0:k1) and of B (rows k0:k1) wrapped as device matrices, product by
    libm4ri_b200 on torch's current stream. Side effect: the thread's library stream stays bound to
    that torch stream afterwards (release_stream undoes it)."""
    sa, sb, sc = tA.stride(0), tB.stride(0), tC.stride(0)
    dC = m4.DMat.wrap(tC.data_ptr(), rows, n, sc, keepalive=tC)
    cache = {}

    def product(i, k0, k1, clear):
        # works[i].wait() orders NCCL's stream before torch's current stream: issue the product there
        import ctypes
        m4.lib().m4ri_b200_set_stream(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        if i not in cache:
            if k0 % 128:
                raise ValueError("k-chunk boundaries must be multiples of 128 bits")
            dA = m4.DMat.wrap(tA.data_ptr() + (k0 // 64) * 8, rows, k1 - k0, sa, keepalive=tA)
            dB = m4.DMat.wrap(tB.data_ptr() + k0 * sb * 8, k1 - k0, n, sb, keepalive=tB)
            cache[i] = (dA, dB)
        dA, dB = cache[i]
        m4.dmul(dC, dA, dB, clear=clear, algo=algo)

    return product


def exchange_slab(k, world, rank, src=0, align=1024, root_serves=False):
    """k-slab of B that `rank` transposes and scatters in the fused exchange: (first row, rows).
    Slab boundaries are multiples of `align` rows (whole 1024-row tiles of the wide transpose,
    128-byte aligned column offsets in Bt).

    Default: every rank takes one slab. root_serves=True (3 or more GPUs) makes the root only
    serve its B while the other world-1 ranks share the work: the root's NVLink egress drops from
    2(G-1)/G |B| to |B|, but the workers' own read + store traffic grows by G/(G-1) and they
    become the bottleneck — measured equal at N = 4 (1.16 vs 1.21 ms,
    profiles/r1k_exchange_probe_n4*.json), so the simpler plan is the default."""
    if not root_serves or world < 3:
        return row_block(k, world, rank, align)
    if rank == src:
        return 0, 0
    return row_block(k, world - 1, rank - 1 if rank > src else rank, align)


class FusedBroadcastMul:
    """The same sharded product with the exchange fused into the transpose (GPU only).

    The tensor path consumes B transposed (Bt, n x k, k-contiguous). Instead of "NCCL
    broadcast of B, then every rank transposes all of it", rank g transposes ONE k-slab of
    B — read straight from the root's memory over NVLink — and the same kernel stores the
    slab of Bt into every rank's Bt buffer with 128-byte peer stores (transpose_wide_kernel). Root
    egress is |B| instead of (G-1)|B|, every link carries (G-1)/G |B| in parallel, no SM is
    held by a collective while the product runs, and the product is one launch over the full
    k. Buffers are torch symmetric memory (NVLink-mapped into every process of the node);
    two device-side barriers per step order the peer stores against the products.
    """

    def __init__(self, m4, m, k, n, dist, src=0, root_serves=None, chunks=None, plan=None):
        import ctypes
        import os
        import torch.distributed._symmetric_memory as symm_mem
        self.m4, self.L, self.dist, self.src = m4, m4.lib(), dist, src
        self.m, self.k, self.n = m, k, n
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        if self.world > 8:
            raise ValueError("one NVLink domain of at most 8 GPUs")
        if k % 128 or n % 128:
            raise ValueError("k and n must be multiples of 128 bits")
        self.r0, self.rows = row_block(m, self.world, self.rank)
        # Plan of the exchange (tools/gpu_exchange.sh; defaults measured at N = 4 and 8):
        #   root_serves  the root only serves its B, the other ranks share the slabs: the root's
        #                NVLink egress is the bottleneck otherwise (reads served + its own stores)
        #   chunks       a non-root rank pulls its slab with the copy engine in `chunks` pieces
        #                into a local staging buffer while the scatter kernel works on the previous
        #                piece: remote reads (ingress) overlap the peer stores (egress) instead of
        #                alternating inside one kernel
        #   plan         "peer": G peer stores per output row (above). "mc_slabs": the same slabs, but
        #                ONE store per row to the NVSwitch multicast address of the symmetric Bt
        #                buffer — the switch replicates it into every GPU's copy (NVLS), so a
        #                rank's egress is its slab, not (G-1) x its slab. "mc_root": the root alone
        #                transposes all of its (local) B and multicasts Bt: no remote reads, root
        #                egress |B| (the floor for a B that starts on one GPU). "auto": mc_root
        #                when the handle has a multicast address, else peer.
        #                (multimem.st and st.global compile to the same STG on sm_100a: a plain
        #                store to the multicast mapping is the multicast store)
        if plan is None:
            plan = os.environ.get("M4RI_B200_XCHG_PLAN", "auto")
        if root_serves is None:
            root_serves = os.environ.get("M4RI_B200_XCHG_ROOT_SERVES", "auto")
            root_serves = (self.world >= 6) if root_serves == "auto" else root_serves == "1"
        if chunks is None:
            chunks = int(os.environ.get("M4RI_B200_XCHG_CHUNKS", "4"))
        self.root_serves, self.chunks = bool(root_serves), max(1, chunks)
        self.plan = plan  # resolved below, once the buffers have a handle
        self.s0, self.srows = exchange_slab(k, self.world, self.rank, src, root_serves=self.root_serves)
        ws, kws = n // 64, k // 64
        dev = torch.device("cuda", torch.cuda.current_device())
        self.tB = symm_mem.empty((k, ws), dtype=torch.int64, device=dev)    # valid on src only
        self.tBt = symm_mem.empty((n, kws), dtype=torch.int64, device=dev)  # filled by all ranks
        self.hB = symm_mem.rendezvous(self.tB, dist.group.WORLD)
        self.hBt = symm_mem.rendezvous(self.tBt, dist.group.WORLD)
        self.tA = torch.empty((self.rows, kws), dtype=torch.int64, device=dev)
        self.tC = torch.empty((self.rows, ws), dtype=torch.int64, device=dev)
        D = m4.DMat
        self.dA = D.wrap(self.tA.data_ptr(), self.rows, k, kws, keepalive=self.tA)
        self.dC = D.wrap(self.tC.data_ptr(), self.rows, n, ws, keepalive=self.tC)
        self.dB = D.wrap(self.tB.data_ptr(), k, n, ws, keepalive=self.tB)
        self.dBt = D.wrap(self.tBt.data_ptr(), n, k, kws, keepalive=self.tBt)
        mc = int(getattr(self.hBt, "multicast_ptr", 0) or 0)
        if self.plan == "auto":
            self.plan = "mc_root" if mc else "peer"
        if self.plan != "peer" and not mc:
            raise RuntimeError("exchange plan %r needs a multicast mapping of the symmetric buffer" % self.plan)
        if self.plan == "mc_root":  # one slab: everything, on the root
            self.s0, self.srows = (0, k) if self.rank == src else (0, 0)
        self.pieces = []  # (source matrix, destination array, staging copy or None)
        if self.srows > 0:
            remote = self.rank != src and self.chunks > 1
            tiles = self.srows // 1024 if self.srows % 1024 == 0 else 0
            nch = min(self.chunks, tiles) if (remote and tiles) else 1
            if nch > 1:
                self.stage = torch.empty((self.srows, ws), dtype=torch.int64, device=dev)
                self.rootB = self.hB.get_buffer(src, (k, ws), torch.int64)
                self.copy_stream = torch.cuda.Stream(device=dev)
                self.ev_go = torch.cuda.Event()
            self.keep = []
            for c in range(nch):
                c0, c1 = (tiles * c // nch) * 1024 if nch > 1 else 0, (tiles * (c + 1) // nch) * 1024 if nch > 1 else self.srows
                if nch > 1:
                    dSrc = D.wrap(self.stage.data_ptr() + c0 * ws * 8, c1 - c0, n, ws, keepalive=self.stage)
                    copy = (self.stage[c0:c1], self.rootB[self.s0 + c0: self.s0 + c1], torch.cuda.Event())
                else:
                    dSrc = D.wrap(int(self.hB.buffer_ptrs[src]) + self.s0 * ws * 8, self.srows, n, ws, keepalive=self.tB)
                    copy = None
                ptrs = self.hBt.buffer_ptrs if self.plan == "peer" else [mc]
                dDst = [D.wrap(int(p) + ((self.s0 + c0) // 64) * 8, n, c1 - c0, kws, keepalive=self.tBt) for p in ptrs]
                self.keep.append(dDst)
                self.pieces.append((dSrc, (ctypes.c_void_p * len(dDst))(*[d.h for d in dDst]), len(dDst), copy))

    def exchange(self):
        """Distribute the root's B: afterwards every rank's Bt holds all of B transposed.
        Side effect: the thread's library stream stays bound to torch's current stream
        (release_stream undoes it)."""
        # the symmetric-memory barriers run on torch's current stream: the library must issue the
        # scatter and the product on that same stream or the barriers order nothing
        import ctypes
        self.L.m4ri_b200_set_stream(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        self.hBt.barrier(channel=0)  # every rank is done reading its Bt of the previous step
        main = torch.cuda.current_stream()
        if len(self.pieces) > 1:
            # the staging buffer is free again: the previous step's scatters precede this point
            self.ev_go.record(main)
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(self.ev_go)
                for (_, _, _, (dst, srcv, ev)) in self.pieces:
                    dst.copy_(srcv, non_blocking=True)
                    ev.record(self.copy_stream)
        for (dSrc, dst_arr, ndst, copy) in self.pieces:
            if copy is not None:
                main.wait_event(copy[2])
            self.L.m4ri_b200_dtranspose_scatter(dst_arr, ndst, dSrc.h)
        self.hBt.barrier(channel=1)  # every slab of Bt has landed everywhere

    def step(self, algo=0):
        self.exchange()
        if self.rows > 0:
            self.m4.dmul_abt(self.dC, self.dA, self.dBt, clear=True, algo=algo)
