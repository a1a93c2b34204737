"""Plans of the fused multi-GPU exchange, timed and checked (run under torchrun, N ranks):
for every (root_serves, chunks) plan: the exchange alone, the whole step, and a check that every
rank's Bt equals the transpose of the root's B (checksums against a locally transposed copy).
   python -m torch.distributed.run --nproc-per-node N tools/exchange_probe.py [n]
(the round-1 version of this probe — barriers / scatter to own / peers / all, narrow vs wide peer
stores — produced profiles/r1k_exchange_probe_*.json)"""
import os, sys, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import m4ri_rust_b200 as m4
from m4ri_rust_b200 import sharded

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
L = m4.lib()
stream = torch.cuda.current_stream()
L.m4ri_b200_set_stream(ctypes.c_void_p(stream.cuda_stream))


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps):
        fn()
    b.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / reps], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t), 3)


out = {"n": n, "world": world, "plans": {}}
plans = [("peer", False, 1), ("peer", world >= 6, 4), ("mc_slabs", False, 1), ("mc_slabs", False, 4), ("mc_slabs", world >= 3, 4), ("mc_root", False, 1)]
want = None
def note(msg):
    if os.environ.get("PROBE_VERBOSE"):
        torch.cuda.synchronize()
        print(f"[rank {rank}] {msg}", file=sys.stderr, flush=True)


for (plan, serves, chunks) in plans:
    note(f"plan {plan} {serves} {chunks}")
    try:
        fm = sharded.FusedBroadcastMul(m4, n, n, n, dist, src=0, root_serves=serves, chunks=chunks, plan=plan)
    except RuntimeError as e:
        out["plans"]["%s,root_serves=%d,chunks=%d" % (plan, serves, chunks)] = {"unavailable": str(e)}
        continue
    L.m4ri_b200_dfill_random_at(fm.dA.h, 1, fm.r0 * (n // 64))
    if rank == 0:
        fm.dB.fill_random(2)
    else:
        fm.tB.zero_()
    fm.tBt.zero_()
    torch.cuda.synchronize(); dist.barrier()
    if want is None:  # transpose of the root's B, computed locally from a broadcast copy
        tmp = fm.tB.clone()
        dist.broadcast(tmp, src=0)
        dT = m4.DMat.init(n, n)
        dTmp = m4.DMat.wrap(tmp.data_ptr(), n, n, n // 64, keepalive=tmp)
        L.m4ri_b200_dtranspose(dT.h, dTmp.h)
        want = dT.checksum()
        del dTmp, tmp, dT
    note("want ready")
    fm.exchange()
    note("first exchange done")
    rec = {"exchange_ms": timed(fm.exchange), "step_ms": timed(lambda: fm.step(0))}
    note("timed")
    fm.tBt.zero_(); torch.cuda.synchronize(); dist.barrier()
    fm.exchange(); torch.cuda.synchronize(); dist.barrier()
    ok = torch.tensor([1 if fm.dBt.checksum() == want else 0], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    rec["bt_correct_on_every_rank"] = bool(int(ok))
    rec["pieces"] = len(fm.pieces)
    out["plans"]["%s,root_serves=%d,chunks=%d" % (plan, serves, chunks)] = rec
    del fm
    torch.cuda.synchronize(); dist.barrier()
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
