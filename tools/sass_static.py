#!/usr/bin/env python
"""Static SASS instruction count per source line of one kernel (code-size / instruction-cache work).
usage: cuobjdump -xelf all lib.so; nvdisasm -g -c X.cubin > X.dis; sass_static.py X.dis <kernel substring> [file]"""
import collections
import re
import sys

path, want = sys.argv[1], sys.argv[2]
only = sys.argv[3] if len(sys.argv) > 3 else None
cur, inside = None, False
cnt = collections.Counter()
for line in open(path):
    if line.startswith(".text."):
        inside = want in line
        continue
    if line.startswith(".section"):
        inside = False
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", line):
        cnt[cur] += 1
print("total", sum(cnt.values()))
byf = collections.Counter()
for (f, l), c in cnt.items():
    byf[f] += c
print(dict(byf))
bins = collections.Counter()
for (f, l), c in cnt.items():
    if only is None or f == only:
        bins[(f, l // 10 * 10)] += c
for k, c in sorted(bins.items()):
    if c >= 80:
        print("%s:%d-%d  %d" % (k[0], k[1], k[1] + 9, c))
