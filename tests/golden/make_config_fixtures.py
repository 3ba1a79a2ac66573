#!/usr/bin/env python3
"""Oracle results for whole BASELINE.json configs that are too slow to count on the GPU box inside bench.py
(config 5: 4e9 bytes of mostly distinct words take the oracle about ten minutes on eight cores):

    python tests/golden/make_config_fixtures.py [5]   ->   tests/golden/config5_full_oracle.json

The text comes from tools/corpus (the seeds of SURVEY.md 8(d) G7), the counts from oracle/ (orc_fast_mt, all
host threads): rows, total, words, stop and the two-number digest of the rows (orc_result_checksum, the same
formula as libwc.so's wc_result_checksum).  bench.py compares a config-5 job with this file."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    from oracle import orc
    from tools.corpus import CONFIGS, Corpus
    for cfg in [int(a) for a in sys.argv[1:]] or [5]:
        total = CONFIGS[cfg]["bytes"]
        c = Corpus(cfg, total_bytes=total)
        t0 = time.time()
        data = c.bytes()
        t1 = time.time()
        r = orc.fast_mt(data, keep_rows=False)
        out = {"config": cfg, "bytes": total, "vocab_seed": c.vocab_seed, "text_seed": c.text_seed, "mode": "ref",
               "rows": r.n_rows, "total": r.total, "n_flush": r.n_flush, "n_empty": r.n_empty, "n_lost_empty": r.n_lost_empty,
               "stop": r.stop, "max_line": r.max_line, "checksum": r.checksum,
               "made_by": "tests/golden/make_config_fixtures.py (oracle/wc_oracle.c orc_fast_mt on %d threads; generate %.0f s, count %.0f s)"
                          % (os.cpu_count(), t1 - t0, time.time() - t1)}
        path = os.path.join(ROOT, "tests", "golden", "config%d_full_oracle.json" % cfg)
        with open(path, "w") as f:
            json.dump(out, f, indent=1)
            f.write("\n")
        print(path, out)


if __name__ == "__main__":
    main()
