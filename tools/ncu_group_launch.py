"""Three device-resident products C = A*Bt^T of one rectangular shape with the library's own
Strassen levels — the command tools/gpu_ncu_group.sh captures one grouped leaf launch of.
usage: python tools/ncu_group_launch.py m:k:n"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4
L = m4.lib()
m, k, n = (int(v) for v in sys.argv[1].split(":"))
dA, dBt, dC = m4.DMat.init(m, k).fill_random(1), m4.DMat.init(n, k).fill_random(2), m4.DMat.init(m, n)
for _ in range(3):
    m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
L.m4ri_b200_synchronize()
print("launches", m4.launch_count())
