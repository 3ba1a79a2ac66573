#!/usr/bin/env python3
"""Small word-count jobs for compute-sanitizer (one --tool per run; see profiles/README.md):
the smoke inputs (quirk lines, CRLF, NUL, unterminated tail), an 8 MB config-2 sample and a 3 MB config-5
sample through wc_feed on the default kernel, the same through the lane-serial kernel, the sort path and
strict-ref mode.  Results are checked against the oracle, so a run that passes also computed the right rows."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import __graft_entry__ as ge
    from oracle import orc
    from tools.corpus import Corpus
    pkg = ge._load_pkg()
    c2 = Corpus(2, total_bytes=8 << 20).bytes().tobytes()
    c5 = Corpus(5, total_bytes=3_000_000).bytes().tobytes()
    quirks = b"Hello He He\nHe Hello He\na  b a\n" + c2[: 1 << 18] + b"a b\r\nc d\r\nx\0y z\n" + b"w" * 300 + b" tail"
    jobs = [("quirks", quirks), ("config2_8MB", c2), ("config5_3MB", c5)]
    variants = [("default", {}), ("lane_serial", dict(lane_serial=True)), ("sort", dict(path=pkg.PATH_SORT)),
                ("strict_ref", dict(strict_ref=True))]
    for vname, kw in variants:
        for jname, data in jobs:
            if vname == "strict_ref" and jname == "config5_3MB":
                continue  # (lines longer than 99 bytes: compared with orc_fast_cap below instead)
            with pkg.Engine(short_table_log2=18, long_table_log2=16, arena_bytes=32 << 20, chunk_bytes=1 << 20, **kw) as e:
                e.feed(data)
                got = e.finish()
            exp = orc.fast_cap(data) if vname == "strict_ref" else orc.fast(data)
            assert got.pairs() == exp.pairs() and got.total == exp.total, (vname, jname)
            print("ok", vname, jname, got.n_rows, got.total, flush=True)


if __name__ == "__main__":
    main()
