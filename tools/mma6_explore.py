"""Build A/B variants of kernels_mma6.cu (pipeline-shape -D overrides) into m4ri-rust_b200/lib/exp/.
usage (CPU container): python tools/mma6_explore.py name:KS=512,G=1,NEPI=8,PF=2,HINT=0[,DBG] ...
then on the GPU box:   bash tools/mma6_explore.sh   (times every lib/exp/*.so)"""
import importlib.util, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b", os.path.join(ROOT, "m4ri-rust_b200", "build.py"))
b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
from concurrent.futures import ThreadPoolExecutor
def one(arg):
    name, cfg = arg.split(":")
    defs = []
    for kv in cfg.split(","):
        if kv == "DBG":
            defs.append("M4RI_B200_PERF_HOOKS")
        else:
            k, v = kv.split("=")
            defs.append(f"MMA6_{k}={v}")
    return b.build_variant(name, defs)
b.build()
with ThreadPoolExecutor(8) as ex:
    for out in ex.map(one, sys.argv[1:]):
        print(out)
