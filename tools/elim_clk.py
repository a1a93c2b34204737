import ctypes, os, sys
sys.path.insert(0, "/root/repo")
import m4ri_rust_b200 as m4
L = m4.lib()
n = 8192
W = m4.DMat.init(n, n); W.fill_random(7)
L.m4ri_b200_synchronize()
r = L.m4ri_b200_drref(W.h, 0, None, 0)
L.m4ri_b200_synchronize()
out = (ctypes.c_longlong * 16)()
L.m4ri_b200_elim_clocks(out)
v = list(out)
print("rank", r, "deltas", [v[i+1]-v[i] for i in range(6)], "n,groups,t_fwd,t_ext,kp", v[8:13])
it = (ctypes.c_longlong * 64)()
L.m4ri_b200_elim_iters(it)
it = list(it)
print("per-iteration stamps (ballot, shuffles, rest):", [(it[4*k+1]-it[4*k], it[4*k+2]-it[4*k+1], it[4*k+4]-it[4*k+2]) for k in range(7)])
