"""Counts of the Blackwell-specific SASS mnemonics per kernel of libvqacx.so (evidence that the hot kernels are
tcgen05 / TMEM / TMA code): python tools/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "vqa_counterexamples_b200", "libvqacx.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
pats = ["UTCHMMA", "UTCQMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "UTMAPF"]
counts, order, cur = collections.defaultdict(collections.Counter), [], None
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        order.append(cur)
        continue
    if cur is None:
        continue
    for p in pats:
        for hit in re.findall(r"\b(" + p + r"[\w.]*)", ln):
            counts[cur][hit] += 1
print(f"# cuobjdump -sass {os.path.relpath(so, ROOT)} (sm_100a): Blackwell tensor-core / TMEM / TMA mnemonics per kernel")
for k in order:
    if counts[k]:
        print(f"{k}\n    " + ", ".join(f"{n} x{c}" for n, c in sorted(counts[k].items())))
print("# kernels without any of these mnemonics: " + ", ".join(k.split("::")[-1] for k in order if not counts[k]))
