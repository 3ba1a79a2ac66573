#!/usr/bin/env python3
"""Prints the SASS of one kernel of libwc.so between two addresses (or the whole kernel), compactly.
usage: tools/sass_loop.py <mangled-name-substring> [lo_hex hi_hex]"""
import re, subprocess, sys, os
lib = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cuda-mapreduce_b200", "libwc.so")
pat = sys.argv[1]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", out)
for b in blocks[1:]:
    name = b.split("\n", 1)[0].strip()
    if pat in name:
        lo = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0
        hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 30
        n = 0
        for l in b.split("\n"):
            m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
            if not m: continue
            a = int(m.group(1), 16)
            if lo <= a <= hi:
                print("%05x  %s" % (a, m.group(2).strip())); n += 1
        print("# %s: %d instructions shown" % (name[:60], n))
        break
