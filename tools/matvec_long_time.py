"""Long-row mat-vec (mul_v_long_kernel) on a few shapes, L2 flushed before every launch, median of 10.
(r2: a variant with 32-byte loads was measured with this script and was SLOWER — 35.8 vs 28.7 us at
32768 x 32768 — and was not kept.) usage: python tools/matvec_long_time.py"""
import sys,time,os,torch,ctypes
sys.path.insert(0,"/root/repo")
import m4ri_rust_b200 as m4
L=m4.lib()
stream=torch.cuda.current_stream(); L.m4ri_b200_set_stream(ctypes.c_void_p(stream.cuda_stream))
flush=torch.empty(256<<20,dtype=torch.uint8,device="cuda")
for (m,n) in [(32768,32768),(65536,16384),(16384,65536),(32768,8192)]:
    dA=m4.DMat.init(m,n).fill_random(1); dx=m4.DMat.init(1,n).fill_random(2); dy=m4.DMat.init(1,m)
    ts=[]
    for _ in range(12):
        flush.fill_(1)
        a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        a.record(stream); L.m4ri_b200_dmul_v(dy.h,dA.h,dx.h,1); b.record(stream); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    ts=sorted(ts[2:]); ms=ts[len(ts)//2]
    print(m,n,round(ms*1e3,1),"us",round(m*n/8/ms/1e6),"GB/s",flush=True)
