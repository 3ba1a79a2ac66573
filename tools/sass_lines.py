#!/usr/bin/env python3
"""Static view of a kernel's SASS by source line: tools/sass_lines.py lib.so kernel_substr [lo hi]
prints, per source line in [lo, hi], the number of SASS instructions, and lists local-memory ops."""
import glob, os, re, subprocess, sys, tempfile, collections
so, kern = sys.argv[1:3]
lo = int(sys.argv[3]) if len(sys.argv) > 3 else 0
hi = int(sys.argv[4]) if len(sys.argv) > 4 else 10**9
td = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=td, check=True, capture_output=True)
cnt = collections.Counter(); loc = collections.Counter(); total = 0
for cubin in glob.glob(os.path.join(td, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    fn = cur = None
    for ln in dis.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m: fn = m.group(1); continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m and fn and kern in fn and cur:
            total += 1
            cnt[cur] += 1
            if re.search(r"\b(STL|LDL)", m.group(2)): loc[cur] += 1
print("total SASS instructions in kernel:", total, " local-memory ops:", sum(loc.values()))
for (f, l), c in sorted(cnt.items()):
    if f == "wc_device.cuh" and lo <= l <= hi:
        print("%s:%d  %d%s" % (f, l, c, ("   LOCAL x%d" % loc[(f, l)]) if loc[(f, l)] else ""))
