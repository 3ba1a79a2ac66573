#!/usr/bin/env python3
"""Times libwc.so's host reduce (csrc/wc_finalize.cpp: finalize_rows -- what the sort path's wc_finish and
WC_HOST_REDUCE=1 run) on a synthetic table of N distinct words, on this machine's cores; no GPU needed.

    WC_TIMING=1 python tools/host_reduce_bench.py [N]        prints seconds, rows, total and wc_result_checksum
"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
    lib = C.CDLL(os.environ.get("WC_LIB") or os.path.join(ROOT, "cuda-mapreduce_b200", "libwc.so"))
    rng = np.random.default_rng(7)
    lens = rng.integers(6, 15, size=n, dtype=np.uint32)
    lens[rng.random(n) < 0.01] = 24                       # a few words past 16 bytes
    offs = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    blob = rng.integers(97, 101, size=int(offs[-1]), dtype=np.uint8)   # 4-letter alphabet: plenty of prefix pairs
    # make the words distinct: write the word's index in base 20 ('g'..'z') over its last 6 bytes
    idx = np.arange(n, dtype=np.int64)
    for k in range(6):
        blob[offs[1:] - 1 - k] = 103 + (idx % 20)
        idx //= 20
    # every fifth word becomes the head of the word after it (where that one is long enough to keep its own tail):
    # prefix pairs for main.cu's merge rule, at every length
    src = np.arange(0, n - 1, 5)
    src = src[lens[src + 1] >= lens[src] + 6]
    for k in range(int(lens.max())):
        m = src[lens[src] > k]
        blob[offs[m + 1] + k] = blob[offs[m] + k]
    counts = rng.integers(1, 1000, size=n, dtype=np.int64)
    firsts = rng.permutation(n).astype(np.uint64) * 7 + 3
    out = C.c_void_p()
    lib.wc_debug_finalize.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_uint64, C.c_uint64,
                                      C.POINTER(C.c_void_p)]
    lib.wc_result_rows.restype = C.c_size_t
    lib.wc_result_rows.argtypes = [C.c_void_p]
    lib.wc_result_total.restype = C.c_int64
    lib.wc_result_total.argtypes = [C.c_void_p]
    lib.wc_result_checksum.argtypes = [C.c_void_p, C.POINTER(C.c_uint64 * 2)]
    t = time.perf_counter()
    rc = lib.wc_debug_finalize(blob.ctypes.data, lens.ctypes.data, counts.ctypes.data, firsts.ctypes.data, n, 0, 5, 2, C.byref(out))
    dt = time.perf_counter() - t
    assert rc == 0
    ck = (C.c_uint64 * 2)()
    lib.wc_result_checksum(out, C.byref(ck))
    print({"words": n, "seconds": round(dt, 2), "rows": lib.wc_result_rows(out), "total": lib.wc_result_total(out),
           "checksum": [ck[0], ck[1]], "threads": os.cpu_count()})


if __name__ == "__main__":
    main()
