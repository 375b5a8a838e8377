#!/usr/bin/env python
"""Static SASS instruction counts per kernel of a built .so (cuobjdump -sass); optional opcode histogram.
usage: sass_count.py lib.so [kernel-substring]"""
import collections, re, subprocess, sys
so = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else None
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
name = None
cnt = collections.Counter(); ops = collections.defaultdict(collections.Counter)
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        cnt[name] += 1
        ops[name][m.group(2).split(".")[0]] += 1
for n, c in cnt.most_common():
    if pat and pat not in n: continue
    print(f"{c:6d} {n}")
    if pat:
        print("       " + " ".join(f"{o}:{k}" for o, k in ops[n].most_common(25)))
