#!/usr/bin/env python3
"""Joins an ncu SASS source page (ncu -i X.ncu-rep --page source --csv) with nvdisasm line info of
the same cubin, and prints per-source-line totals: instructions executed, stall samples, shared
wavefronts.  Usage: tools/ncu_by_line.py prof.ncu-rep libwc.so kernel_substring [top_n]"""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, so, kern = sys.argv[1:4]
    topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    td = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=td, check=True, capture_output=True)
    lines_of = {}
    for cubin in glob.glob(os.path.join(td, "*.cubin")):
        dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
        cur_fn, cur_line, inl = None, None, None
        for ln in dis.splitlines():
            m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
            if m:
                cur_fn = m.group(1)
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
            if m:
                cur_line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m and cur_fn and kern in cur_fn:
                lines_of[int(m.group(1), 16)] = (cur_line, m.group(2))
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    base = None
    agg = {}
    tot_inst = tot_samp = 0
    dump = open(os.environ["NCU_DUMP"], "w") if os.environ.get("NCU_DUMP") else None  # per-instruction listing
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        addr = int(r[col["Address"]], 16)
        if base is None:
            base = addr
        off = addr - base
        line, sass = lines_of.get(off, (None, r[col["Source"]]))
        inst = int(r[col["Instructions Executed"]] or 0)
        samp = int(r[col["# Samples"]] or 0)
        wf = int(r[col["L1 Wavefronts Shared"]] or 0)
        long_sb = int(r[col["stall_long_sb"]] or 0)
        if dump:
            dump.write("%05x %-22s %10d %6d %8d %6d  %s\n" % (off, "%s:%d" % line if line else "?", inst, samp, wf, long_sb, sass))
        a = agg.setdefault(line, [0, 0, 0, 0])
        a[0] += inst; a[1] += samp; a[2] += wf; a[3] += long_sb
        tot_inst += inst; tot_samp += samp
    print("total warp-instructions %d, samples %d" % (tot_inst, tot_samp))
    print("%-28s %14s %7s %9s %7s %12s %9s" % ("line", "inst", "inst%", "samples", "samp%", "smem_wavefr", "long_sb"))
    for line, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
        name = "%s:%d" % line if line else "?"
        print("%-28s %14d %6.2f%% %9d %6.2f%% %12d %9d" % (name, a[0], 100.0 * a[0] / max(tot_inst, 1), a[1],
                                                            100.0 * a[1] / max(tot_samp, 1), a[2], a[3]))


if __name__ == "__main__":
    main()
