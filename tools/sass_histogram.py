"""Opcode histogram of the Blackwell-specific SASS in the built objects (evidence that the kernels are tcgen05 / TMEM /
TMA code, not recompiled legacy paths).  usage: python tools/sass_histogram.py > profiles/rNN_sass_histogram.txt
SASS mnemonics (B200_PROFILING.md): tcgen05.mma -> UTC*MMA, tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG/UBLKCP,
tcgen05.commit / mbarrier -> UTCBAR / SYNCS, legacy mma.sync -> HMMA, cp.async -> LDGSTS."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "vision-transformer_b200", "build")
KEYS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "UTCCP", "SYNCS", "HMMA",
        "LDGSTS", "UCGABAR", "REDG", "ATOMG", "MUFU.EX2"]
print("# cuobjdump -sass of vision-transformer_b200/build/*.o (nvcc -gencode arch=compute_100a,code=sm_100a)")
print(f"{'object':26s} " + " ".join(f"{k:>8s}" for k in KEYS) + "   kernels")
tot = collections.Counter()
for f in sorted(os.listdir(OBJ)):
    if not f.endswith(".o"):
        continue
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, f)], capture_output=True, text=True).stdout
    c = collections.Counter()
    nk = len(re.findall(r"^\s*Function : ", out, flags=re.M))
    for line in out.splitlines():
        m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        for k in KEYS:
            if op == k or op.startswith(k + ".") or (k == "MUFU.EX2" and op.startswith("MUFU.EX2")):
                c[k] += 1
    tot.update(c)
    print(f"{f:26s} " + " ".join(f"{c[k]:8d}" for k in KEYS) + f"   {nk}")
print(f"{'TOTAL':26s} " + " ".join(f"{tot[k]:8d}" for k in KEYS))
