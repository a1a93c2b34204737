import sys,time
sys.path.insert(0,"/root/repo")
import m4ri_rust_b200 as m4
L=m4.lib()
for n in (1024,2048,4096,8192):
    dA,dB=m4.DMat.init(n,n).fill_random(1),m4.DMat.init(n,n).fill_random(2)
    hA,hB=L.mzd_init(n,n),L.mzd_init(n,n)
    L.m4ri_b200_download(hA,dA.h); L.m4ri_b200_download(hB,dB.h); L.m4ri_b200_synchronize()
    for name,f in (("mzd_mul",lambda: L.mzd_mul(None,hA,hB,0)),("mzd_mul_naive",lambda: L.mzd_mul_naive(None,hA,hB)),("mzd_mul_m4rm",lambda: L.mzd_mul_m4rm(None,hA,hB,0))):
        L.mzd_free(f())
        t0=time.perf_counter()
        for _ in range(3): L.mzd_free(f())
        print(n,name,round((time.perf_counter()-t0)/3*1e3,3),"ms",flush=True)
