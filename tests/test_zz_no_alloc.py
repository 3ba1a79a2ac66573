"""wc_count_device_static: the count kernel for callers that may not allocate.

Some users of libwc.so cannot call cudaMalloc -- the process gave the whole device heap to its own
pool, or the call sits on a path where the device-wide stall of cudaMalloc/cudaFree is not
acceptable.  For them the library offers the same tokenizer/count kernel with its working state in a
fixed array inside the library image; the caller passes the text and every result buffer.

This test calls that entry point through the C ABI twice in a row (the second text is different and
smaller, so anything the first call left in the shared array would show) and checks the tables it
fills, word for word, count for count and first offset for first offset, against the regular entry
point (wc_feed/wc_finish on an engine), bit-exact like every other parity test here.  It lives in
its own file, named to be collected last.
"""
import ctypes as C
import random

import numpy as np
import pytest

from conftest import load_pkg
from oracle import orc
from test_oracle import random_text

pytestmark = pytest.mark.gpu

S_LOG2, L_LOG2, ARENA = 16, 14, 4 << 20
NONE = (1 << 64) - 1


def decode_tables(short, long_, arena):
    """{word: (count, first_offset)} from the raw device tables (layout: include/wc.h)."""
    out = {}
    s_n, l_n = 1 << S_LOG2, 1 << L_LOG2
    keys, counts, finv = short[:2 * s_n], short[2 * s_n:3 * s_n], short[3 * s_n:4 * s_n]
    kraw = keys.tobytes()
    for i in np.nonzero(keys[0::2])[0]:
        if counts[i] == 0:
            continue
        word = kraw[16 * i:16 * i + 16].rstrip(b"\0")
        out[word] = (int(counts[i]), NONE - int(finv[i]))
    refs, counts, finv = long_[:l_n], long_[l_n:2 * l_n], long_[2 * l_n:]
    raw = arena.tobytes()
    for i in np.nonzero(refs)[0]:
        if counts[i] == 0:
            continue
        unit = int(refs[i]) & ((1 << 40) - 1)
        ln = int(arena[unit])
        out[raw[(unit + 1) * 8:(unit + 1) * 8 + ln]] = (int(counts[i]), NONE - int(finv[i]))
    return out


def count_static(pkg, torch, data, bufs, s_log2=S_LOG2, decode=True):
    short, long_, arena, counters = bufs
    text = torch.frombuffer(bytearray(data) if data else bytearray(16), dtype=torch.uint8).cuda()
    rc = pkg.lib().wc_count_device_static(0, C.c_void_p(text.data_ptr()), len(data), C.c_void_p(short.data_ptr()),
                                          s_log2, C.c_void_p(long_.data_ptr()), L_LOG2, C.c_void_p(arena.data_ptr()),
                                          ARENA, C.c_void_p(counters.data_ptr()), None)
    assert rc == 0, pkg.lib().wc_static_last_error()
    torch.cuda.synchronize()
    u = lambda t: t.cpu().numpy().view(np.uint64)
    cnt = [int(x) for x in u(counters)]
    return (decode_tables(u(short), u(long_), u(arena)) if decode else None), cnt


def test_static_entry_point_matches_engine_twice_in_a_row():
    import torch
    pkg = load_pkg()
    rng = random.Random(31)
    first = random_text(rng, 40000, weird=0.02)            # CR / NUL / blank lines included
    second = random_text(rng, 3000, weird=0.0).replace(b"\n\n", b"\n") + b"unterminated tail"
    bufs = (torch.empty(4 << S_LOG2, dtype=torch.int64, device="cuda"),
            torch.empty(3 << L_LOG2, dtype=torch.int64, device="cuda"),
            torch.empty(ARENA // 8, dtype=torch.int64, device="cuda"),
            torch.empty(12, dtype=torch.int64, device="cuda"))
    for t in bufs:
        t.fill_(-1)  # the entry point must not rely on the caller having cleared anything
    with pkg.Engine(mode=pkg.MODE_EXACT, short_table_log2=S_LOG2, long_table_log2=L_LOG2, arena_bytes=ARENA) as eng:
        for data in (first, second, first):
            got, counters = count_static(pkg, torch, data, bufs)
            eng.reset()
            eng.feed(data)
            ref = eng.finish()
            want = {w: (c, f) for (w, c), f in zip(ref.rows, ref.firsts)}
            assert counters[6] == 0, "table overflow flag"
            assert got == want
            # ... and against the checker itself (exact mode: plain equality counts, first = flush offset)
            exp = orc.fast(data, mode=orc.MODE_EXACT)
            assert got == {w: (c, f) for (w, c, f) in exp.rows}
            assert (counters[0], counters[1]) == (exp.n_flush, exp.n_empty)
            stop = counters[2] if counters[2] != NONE else len(data)
            assert (counters[0], counters[1], stop) == (ref.stats["n_flush"], ref.stats["n_empty"], ref.stats["stop"])


def test_static_entry_point_reports_a_full_table_and_stays_inside_its_buffers():
    """A word table too small for the text: the documented answer is err != 0 in counters[6]
    (there is no overflow log on this path), nothing outside the caller's buffers is touched, and
    the next call with a table of the right size is exact again."""
    import torch
    pkg = load_pkg()
    rng = random.Random(77)
    words = [b"w%05d" % i for i in range(6000)]          # 6000 distinct words against 1024 slots
    data = b"\n".join(b" ".join(rng.choice(words) for _ in range(12)) for _ in range(4000)) + b"\n"
    small_log2 = 10
    guard = 1 << 12                                       # int64 sentinels before and after every buffer
    def guarded(n):
        t = torch.full((n + 2 * guard,), -1, dtype=torch.int64, device="cuda")
        return t, t[guard:guard + n]
    g_short, short = guarded(4 << S_LOG2)
    g_long, long_ = guarded(3 << L_LOG2)
    g_arena, arena = guarded(ARENA // 8)
    g_cnt, counters = guarded(12)
    # dirty the host stack a little before the call (the entry point builds its launch structs there)
    _ = [rng.random() for _ in range(1000)]
    _, cnt = count_static(pkg, torch, data, (short[:4 << small_log2], long_, arena, counters), s_log2=small_log2, decode=False)
    assert cnt[6] != 0, "a 1024-slot table cannot hold 6000 words: err must be raised"
    for g, inner in ((g_short, 4 << S_LOG2), (g_long, 3 << L_LOG2), (g_arena, ARENA // 8), (g_cnt, 12)):
        assert bool((g[:guard] == -1).all()) and bool((g[guard + inner:] == -1).all()), "wrote outside a caller buffer"
    assert bool((short[4 << small_log2:] == -1).all()), "wrote past the 1024-slot table"
    got, cnt = count_static(pkg, torch, data, (short, long_, arena, counters))
    exp = orc.fast(data, mode=orc.MODE_EXACT)
    assert cnt[6] == 0 and got == {w: (c, f) for (w, c, f) in exp.rows}


def test_static_entry_point_refuses_oversize_and_bad_arguments():
    import torch
    pkg = load_pkg()
    L = pkg.lib()
    buf = torch.zeros(1 << 20, dtype=torch.int64, device="cuda")
    p = C.c_void_p(buf.data_ptr())
    too_big = (1 << 36) + 1
    assert L.wc_count_device_static(0, p, too_big, p, 10, p, 10, p, 4096, p, None) == -7  # WC_ERR_TOO_LARGE
    assert b"WC_STATIC_MAX_BYTES" in L.wc_static_last_error()
    assert L.wc_count_device_static(0, C.c_void_p(buf.data_ptr() + 1), 100, p, 10, p, 10, p, 4096, p, None) == -5
    assert L.wc_count_device_static(0, p, 100, None, 10, p, 10, p, 4096, p, None) == -1
