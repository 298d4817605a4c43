#!/usr/bin/env python
"""SASS evidence for profiles/: per kernel of libsailrecon_b200.so the mnemonic histogram (tcgen05 = UTCHMMA / UTCBAR /
LDTM / STTM, TMA = UTMALDG, MUFU.EX2, FFMA2 / FADD2, F2FP ...) and the registers / stack reported by cuobjdump -res-usage.

    python scripts/sass_summary.py > profiles/rNN_sass_summary.txt
"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
objs = sorted(glob.glob(os.path.join(ROOT, "*_b200", "build", "*.o")))
KEY = ("UTCHMMA", "UTCBAR", "UTMALDG", "LDTM", "STTM", "MUFU.EX2", "FFMA2", "FADD2", "FMNMX3", "F2FP", "SYNCS", "BRA.U.ANY",
       "R2UR", "ELECT", "LDL", "STL", "LDG", "STG", "LDS", "STS", "SHFL", "USETMAXREG")
for obj in objs:
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", obj], capture_output=True, text=True).stdout
    usage = dict(re.findall(r"Function (\S+):\n\s*(REG:\d+[^\n]*)", res))
    for m in re.finditer(r"Function : (\S+)\n(.*?)(?=\n\s*Function : |\Z)", sass, re.S):
        name, body = m.group(1), m.group(2)
        ops = collections.Counter()
        total = 0
        for ln in body.split("\n"):
            mm = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if mm:
                total += 1
                op = mm.group(1)
                for k in KEY:
                    if op.startswith(k):
                        ops[k] += 1
        demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        print(f"## {os.path.basename(obj)}: {demangled[:150]}")
        print(f"   {usage.get(name, '')}  instructions {total}")
        print("   " + "  ".join(f"{k} {ops[k]}" for k in KEY if ops[k]))
