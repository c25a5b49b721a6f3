#!/usr/bin/env python3
"""Instruction mix of one kernel in an object file: python tools/sass_mix.py build/eoc/select_tile.o <substring of mangled name>"""
import collections
import re
import subprocess
import sys

obj, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
cur = None
mix = collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    if cur and pat in cur:
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            mix[m.group(1)] += 1
tot = sum(mix.values())
print(pat, "total", tot)
for k, v in mix.most_common(25):
    print(f"  {k:12s} {v}")
print("  -- evidence opcodes:", {k: mix.get(k, 0) for k in ("UTMALDG", "UTMAPF", "SYNCS", "FMNMX", "FMNMX3", "IMAD", "LDS", "STS", "BAR", "LDGSTS")})
