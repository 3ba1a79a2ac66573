#!/usr/bin/env python3
"""Hash path vs sort path over the share of distinct words (what WC_PATH_AUTO decides by).

    python tools/path_probe.py [--mb 1024] > profiles/r2_path_crossover.txt

For each corpus: device time of one whole job (wc_reset + count kernel(s) [+ wc_aggregate on the sort path]),
best of 3, text resident in HBM, on both forced paths; rows checked against the oracle once per corpus."""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=1024)
    ap.add_argument("--only", default="", help="comma-separated substrings of the corpus names to run")
    args = ap.parse_args()
    import torch
    import __graft_entry__ as ge
    from oracle import orc
    from tools.corpus import Corpus
    pkg = ge._load_pkg()
    n = args.mb << 20
    specs = [
        ("zipf1.1 V=50k (config 2)", dict(config=2)),
        ("zipf1.1 V=1M (config 3)", dict(config=3)),
        ("zipf0.8 V=8M", dict(config=3, vocab=8_000_000, s=0.8)),
        ("zipf0.5 V=32M", dict(config=3, vocab=32_000_000, s=0.5)),
        ("zipf0.2 V=64M", dict(config=3, vocab=64_000_000, s=0.2)),
        ("adversarial (config 5)", dict(config=5)),
    ]
    print("# %d MiB of each corpus, one B200; ms = wall time of reset + count (+ aggregate) with the device synchronised, best of 3" % args.mb)
    print("%-28s %10s %12s %9s %10s %10s %8s" % ("corpus", "words", "distinct", "ratio", "hash ms", "sort ms", "oracle"))
    for name, kw in specs:
        if args.only and not any(o in name for o in args.only.split(",")):
            continue
        c = Corpus(total_bytes=n, **kw)
        host = c.bytes()
        dev = torch.from_numpy(host).cuda()
        exp = orc.fast_mt(host, keep_rows=False)
        res = {}
        for path, pid in (("hash", pkg.PATH_HASH), ("sort", pkg.PATH_SORT)):
            distinct_guess = max(1 << 18, exp.n_rows * 4)
            s_log2 = 12 if path == "sort" else max(18, (distinct_guess - 1).bit_length())
            with pkg.Engine(path=pid, short_table_log2=min(s_log2, 31), long_table_log2=22, arena_bytes=1 << 30) as e:
                best = 1e9
                for _ in range(3):
                    e.reset(); e.sync()
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    e.count_device(dev.data_ptr(), dev.numel(), last_piece=True)
                    e.aggregate()
                    e.sync()
                    best = min(best, (time.perf_counter() - t0) * 1e3)
                got = e.finish(keep_rows=False)
                ok = (got.n_rows, got.total, got.stats["n_flush"]) == (exp.n_rows, exp.total, exp.n_flush)
                res[path] = (best, ok, got.stats["n_distinct"], got.stats["n_flush"] - got.stats["n_empty"])
        words, distinct = res["hash"][3], res["hash"][2]
        print("%-28s %10d %12d %9.4f %10.2f %10.2f %8s" % (name, words, distinct, distinct / max(words, 1), res["hash"][0],
                                                        res["sort"][0], "ok" if res["hash"][1] and res["sort"][1] else "MISMATCH"), flush=True)
        del dev


if __name__ == "__main__":
    main()
