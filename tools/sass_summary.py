#!/usr/bin/env python
"""Per-kernel SASS evidence of the in-tree library: counts of the mnemonics that prove tcgen05 / TMEM / TMA
(`cuobjdump -sass libqi_b200.so`; no GPU needed).  Output is committed as profiles/sass_summary.txt."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vllm_omni_draft_b200", "libqi_b200.so")
MNEMONICS = ["UTCHMMA", "UTCHMMA.2CTA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "SYNCS",
             "HMMA", "MUFU.EX2", "MUFU.TANH", "FFMA2", "ELECT", "ACQBULK", "USETMAXREG", "RED", "ST.E", "LD.E"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", cur).replace("void ", "").replace("qi::", "")
            per[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        per[cur]["_total"] += 1
        for mn in MNEMONICS:
            if mn == "UTCHMMA.2CTA":
                if op.startswith("UTCHMMA") and ".2CTA" in op:
                    per[cur][mn] += 1
            elif op == mn or op.startswith(mn + "."):
                per[cur][mn] += 1
    cols = [m for m in MNEMONICS if any(c[m] for c in per.values())]
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} : per-kernel instruction counts (static SASS)")
    print(f"# {'kernel':58s} " + " ".join(f"{c:>12s}" for c in ["instr"] + cols))
    tot = collections.Counter()
    for k, c in per.items():
        print(f"{k[:60]:60s} " + " ".join(f"{c[m]:12d}" for m in ["_total"] + cols))
        tot.update(c)
    print(f"{'TOTAL':60s} " + " ".join(f"{tot[m]:12d}" for m in ["_total"] + cols))


if __name__ == "__main__":
    main()
