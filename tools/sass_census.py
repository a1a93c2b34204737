"""SASS census of the in-tree library: per kernel, the counts of the Blackwell-native instructions
(tcgen05 MMA forms, tensor-memory loads/stores, commit barriers, TMA loads) and of the fallbacks.

    python tools/sass_census.py > profiles/<round>_sass_census.txt        (CPU only: cuobjdump)
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "m4ri-rust_b200", "lib", "libm4ri_b200.so")
COLS = [("UTCOMMA", r"\bUTCOMMA"), ("UTCIMMA", r"\bUTCIMMA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"),
        ("UTCBAR", r"\bUTCBAR"), ("UTMALDG", r"\bUTMALDG"), ("LDGSTS", r"\bLDGSTS"), ("SYNCS", r"\bSYNCS"),
        ("IMMA", r"\bIMMA")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.splitlines()
    bodies = re.split(r"\n\s*Function : \S+\n", sass)[1:]
    print("# SASS census of the in-tree library (cuobjdump -sass m4ri-rust_b200/lib/libm4ri_b200.so, tools/sass_census.py)")
    print("# columns: kernel, " + ", ".join(c for c, _ in COLS) + "; kernels with none of them are left out")
    for name, body in sorted(zip(names, bodies)):
        counts = [len(re.findall(rx, body)) for _, rx in COLS]
        if any(counts):
            print(name, *counts)


if __name__ == "__main__":
    main()
