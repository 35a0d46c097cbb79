#!/usr/bin/env python
"""Per-kernel SASS instruction counts of libfdlp_b200.so (cuobjdump -sass): the evidence of which hardware paths the
kernels use.  DMMA = FP64 tensor path (mma.sync.m8n8k4.f64; tcgen05 has no FP64 kind), HMMA = TF32 / BF16 mma.sync,
UTC*MMA / LDTM = tcgen05 + TMEM, UTMALDG / UBLKCP = TMA bulk copies, LDG.E.128 = 16-byte global loads, STL / LDL = register
spills.    python tools/sass_counts.py > profiles/sass_r02.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "speech_recognition_tools_b200", "libfdlp_b200.so")
PATTERNS = [("DMMA", r"\bDMMA\."), ("HMMA", r"\bHMMA\."), ("IMMA", r"\bIMMA\."), ("UTCMMA", r"\bUTC\w*MMA"),
            ("LDTM", r"\bLDTM"), ("TMA", r"\bUTMALDG|\bUBLKCP|\bUTMASTG"), ("SYNCS", r"\bSYNCS\."),
            ("DFMA", r"\bDFMA\b"), ("DADD", r"\bDADD\b"), ("DMUL", r"\bDMUL\b"), ("FFMA", r"\bFFMA\b"),
            ("LDG.128", r"\bLDG\.E\.128"), ("LDG.64", r"\bLDG\.E\.64"), ("LDG", r"\bLDG\."), ("STG", r"\bSTG\."),
            ("LDS", r"\bLDS\b|\bLDS\."), ("STS", r"\bSTS\b|\bSTS\."), ("SHFL", r"\bSHFL\."), ("CCTL.PF", r"\bCCTL\.[\w.]*PF"),
            ("ATOM/RED", r"\bATOMG?\.|\bRED\."), ("STL", r"\bSTL\b|\bSTL\."), ("LDL", r"\bLDL\b|\bLDL\."), ("BAR", r"\bBAR\.")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = re.findall(r"Function : (\S+)", sass)
    dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.strip().split("\n")
    parts = re.split(r"Function : \S+", sass)[1:]
    print("# %s: %d kernels; arch %s" % (os.path.basename(LIB), len(names), ",".join(sorted(set(re.findall(r"arch = (sm_\w+)", sass))))))
    print("# %-66s %7s  %s" % ("kernel", "instr", "  ".join("%s" % n for n, _ in PATTERNS)))
    tot = collections.Counter()
    for name, body in sorted(zip(dem, parts)):
        lines = [l for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        c = {n: sum(1 for l in lines if re.search(p, l)) for n, p in PATTERNS}
        for n, v in c.items():
            tot[n] += v
        short = re.sub(r"^void fdlp::|\(.*$", "", name)
        print("%-68s %7d  %s" % (short[:68], len(lines), "  ".join("%s=%d" % (n, v) for n, v in c.items() if v)))
    print("# total: " + "  ".join("%s=%d" % (n, tot[n]) for n, _ in PATTERNS))


if __name__ == "__main__":
    sys.exit(main())
