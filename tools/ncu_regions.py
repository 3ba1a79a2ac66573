#!/usr/bin/env python3
"""Groups tools/ncu_by_line.py output into named source regions.
usage: tools/ncu_regions.py prof.ncu-rep lib.so kernel tiles "name:lo-hi,name:lo-hi,..." """
import re, subprocess, sys, os
rep, so, kern, tiles, spec = sys.argv[1:6]
tiles = float(tiles)
groups = []
for item in spec.split(","):
    n, r = item.split(":"); a, b = r.split("-"); groups.append((n, int(a), int(b)))
here = os.path.dirname(os.path.abspath(__file__))
out = subprocess.run([sys.executable, os.path.join(here, "ncu_by_line.py"), rep, so, kern, "100000"], capture_output=True, text=True).stdout
tot = 0; agg = {}
for ln in out.splitlines()[2:]:
    p = ln.split()
    name = p[0]; inst = int(p[1]); samp = int(p[3]); wf = int(p[5]); lsb = int(p[6])
    tot += inst
    m = re.match(r"wc_device.cuh:(\d+)", name)
    g = None
    if m:
        L = int(m.group(1))
        for gn, a, b in groups:
            if a <= L <= b: g = gn; break
    if g is None: g = name
    x = agg.setdefault(g, [0, 0, 0, 0]); x[0] += inst; x[1] += samp; x[2] += wf; x[3] += lsb
ts = sum(v[1] for v in agg.values())
print("total warp-instr %d = %.1f per tile; samples %d" % (tot, tot / tiles, ts))
for g, (i, s, w, l) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
    print("%-34s inst/tile %7.1f  inst%% %5.1f  samples%% %5.1f  long_sb%% %5.1f  smem_wf/tile %6.1f" % (g, i / tiles, 100 * i / tot, 100 * s / ts, 100 * l / ts, w / tiles))
