#!/usr/bin/env python
"""Static SASS instruction count per source line of one kernel (nvdisasm -g -c output of a cubin).
  cuobjdump -xelf all lib.so; nvdisasm -g -c kernels.sm_100a.cubin > all.sass; python scripts/sass_lines.py all.sass <kernel substring>"""
import re
import sys
from collections import Counter

path, key = sys.argv[1], sys.argv[2]
inside, cur = False, None
per_line, per_fn = Counter(), Counter()
for ln in open(path, errors="ignore"):
    if ln.startswith(".text."):
        inside = key in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        per_line[cur] += 1
        per_fn[cur[0] if cur else None] += 1
tot = sum(per_line.values())
print(f"# {tot} SASS instructions in *{key}*")
for f, n in per_fn.most_common():
    print(f"{n:6d}  {f}")
print()
for (f, l), n in sorted(per_line.items(), key=lambda kv: (kv[0][0], kv[0][1])):
    if n >= int(sys.argv[3]) if len(sys.argv) > 3 else 1:
        print(f"{n:5d}  {f}:{l}")
