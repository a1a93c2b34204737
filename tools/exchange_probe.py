"""Where the fused exchange's time goes (run under torchrun, N ranks):
   barriers alone / transpose-scatter to own buffer only / to every rank / whole exchange / product.
   python -m torch.distributed.run --nproc-per-node N tools/exchange_probe.py [n]"""
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
fm = sharded.FusedBroadcastMul(m4, n, n, n, dist, src=0)
L.m4ri_b200_dfill_random_at(fm.dA.h, 1, fm.r0 * (n // 64))
if rank == 0:
    fm.dB.fill_random(2)
else:
    fm.tB.zero_()
torch.cuda.synchronize(); dist.barrier()

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
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    return [round(float(x), 3) for x in allt]

has = fm.srows > 0  # from 3 ranks up the root only serves reads
own = (ctypes.c_void_p * 1)(fm.dDst[rank].h) if has else None
others = [d.h for i, d in enumerate(fm.dDst) if i != rank] if has else []
oth = (ctypes.c_void_p * len(others))(*others) if has else None
def scat(arr, cnt, src):
    if has:
        L.m4ri_b200_dtranspose_scatter(arr, cnt, src)
def barriers():
    fm.hBt.barrier(channel=0); fm.hBt.barrier(channel=1)
def scatter_own():
    fm.hBt.barrier(channel=0)
    scat(own, 1, (fm.dSlab.h if has else None))
def scatter_peers():
    fm.hBt.barrier(channel=0)
    scat(oth, len(others), (fm.dSlab.h if has else None))
def scatter_all():
    fm.hBt.barrier(channel=0)
    scat(fm.dst_arr if has else None, world, (fm.dSlab.h if has else None))
def exchange():
    fm.hBt.barrier(channel=0)
    scat(fm.dst_arr if has else None, world, (fm.dSlab.h if has else None))
    fm.hBt.barrier(channel=1)
def product():
    m4.dmul_abt(fm.dC, fm.dA, fm.dBt, clear=True, algo=0)
# local-source transpose of the same slab shape (rank-local B rows) for reference
loc = m4.DMat.wrap(fm.tB.data_ptr() + fm.s0 * (n // 64) * 8, fm.srows, n, n // 64, keepalive=fm.tB) if has else None
def scatter_own_localsrc():
    scat(own, 1, (loc.h if has else None))
def scatter_all_localsrc():
    fm.hBt.barrier(channel=0)
    scat(fm.dst_arr if has else None, world, (loc.h if has else None))
out = {"n": n, "world": world}
out["barriers"] = timed(barriers)
for var in ("narrow", "wide"):
    os.environ["M4RI_B200_TRANSPOSE"] = var
    for name, fn in [("scatter_own_localsrc", scatter_own_localsrc), ("scatter_own", scatter_own),
                     ("scatter_peers", scatter_peers), ("scatter_all_localsrc", scatter_all_localsrc),
                     ("scatter_all", scatter_all), ("exchange", exchange)]:
        out[var + ":" + name] = timed(fn)
# destinations rotated by rank (every rank starts with a different peer), wide kernel
if has:
    rot = [fm.dDst[(rank + 1 + j) % world].h for j in range(world)]
    rot_arr = (ctypes.c_void_p * world)(*rot)
def exchange_rot():
    fm.hBt.barrier(channel=0)
    scat(rot_arr if has else None, world, (fm.dSlab.h if has else None))
    fm.hBt.barrier(channel=1)
out["wide:exchange_rotated"] = timed(exchange_rot)
os.environ.pop("M4RI_B200_TRANSPOSE")
out["product"] = timed(product)
out["step"] = timed(lambda: fm.step(0))
L.m4ri_b200_set_strassen_levels(0)
out["product_L0"] = timed(product)
L.m4ri_b200_set_strassen_levels(-1)
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
