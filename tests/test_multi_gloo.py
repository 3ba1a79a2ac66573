"""World-size-2 CPU tests (gloo) of the multi-GPU host logic: shard planning, the shard-counter
reduction (wc_shard_stats_reduce, pure host code in libwc.so) and the key-owner exchange orchestration
(cuda-mapreduce_b200/multi.py) driven end to end with a host-memory stand-in for the engine.  The GPU
kernels behind export/import are covered by the -m gpu tests; here the stand-in keeps a dict and
speaks the same record format, so splits, offsets, ownership and the STOP / lost-empties rules are
all exercised without a GPU."""
import ctypes as C
import os
import random
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_pkg
from oracle import orc as O
from test_oracle import random_text

NONE = (1 << 64) - 1


class HostEngine:
    """Same interface as wc.Engine's multi-GPU methods, over a dict {word: [count, first]}."""

    def __init__(self, pkg, data, base):
        self.pkg, self.base, self.nbytes = pkg, base, len(data)
        ex = O.fast(data, O.MODE_EXACT)
        self.words = {w: [c, base + f] for (w, c, f) in ex.rows}
        self.st = pkg.ShardStats()
        self.st.base, self.st.bytes = base, len(data)
        self.st.n_flush, self.st.n_empty, self.st.lost_empty = ex.n_flush, ex.n_empty, ex.n_lost_empty
        self.st.p0_found = 1 if ex.rows else 0
        self.st.p0 = base + ex.rows[0][2] if ex.rows else NONE
        self.st.stop = base + ex.stop if ex.stop < len(data) else NONE
        self.job = None

    def shard_stats(self):
        return self.st

    def set_job_stats(self, job):
        self.job = job

    def clear_tables(self):
        self.words = {}

    @staticmethod
    def _key(w):
        p = w.ljust(16, b"\0")
        return int.from_bytes(p[:8], "little"), int.from_bytes(p[8:], "little")

    def _short(self):
        return {w: v for w, v in self.words.items() if len(w) <= 16}

    def _long(self):
        return {w: v for w, v in self.words.items() if len(w) > 16}

    def export_counts(self, n_parts):
        counts = [0] * n_parts
        for w in self._short():
            counts[self.pkg.lib().wc_key_owner(*self._key(w), n_parts)] += 1
        lg = self._long()
        return counts, len(lg), sum((len(w) + 7) // 8 * 8 for w in lg)

    def export_words(self, n_parts, offsets, ptr):
        cur = list(offsets)
        total = sum(self.export_counts(n_parts)[0])
        out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint64)), shape=(total * 4,))
        for w, (c, f) in self._short().items():
            k0, k1 = self._key(w)
            p = self.pkg.lib().wc_key_owner(k0, k1, n_parts)
            out[cur[p] * 4:cur[p] * 4 + 4] = [k0, k1, c, f]
            cur[p] += 1

    def export_long(self, pmeta, pblob):
        lg = self._long()
        meta = np.ctypeslib.as_array(C.cast(pmeta, C.POINTER(C.c_uint64)), shape=(len(lg) * 4,))
        nb = sum((len(w) + 7) // 8 * 8 for w in lg)
        blob = np.ctypeslib.as_array(C.cast(pblob, C.POINTER(C.c_uint8)), shape=(nb,))
        off = 0
        for i, (w, (c, f)) in enumerate(lg.items()):
            meta[4 * i:4 * i + 4] = [len(w), c, f, off]
            pad = (len(w) + 7) // 8 * 8
            blob[off:off + pad] = np.frombuffer(w.ljust(pad, b"\0"), dtype=np.uint8)
            off += pad

    def _add(self, w, c, f):
        e = self.words.setdefault(w, [0, NONE])
        e[0] += c
        e[1] = min(e[1], f)

    def import_words(self, ptr, n):
        if n == 0:
            return
        rec = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint64)), shape=(n * 4,)).reshape(n, 4)
        for k0, k1, c, f in rec.tolist():
            w = (int(k0).to_bytes(8, "little") + int(k1).to_bytes(8, "little")).rstrip(b"\0")
            self._add(w, c, f)

    def import_long(self, pmeta, n, pblob):
        meta = np.ctypeslib.as_array(C.cast(pmeta, C.POINTER(C.c_uint64)), shape=(n * 4,)).reshape(n, 4)
        for ln, c, f, off in meta.tolist():
            w = C.string_at(pblob + off, ln)
            self._add(w, c, f)


def _worker(rank, world, port, data, q, mode):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = load_pkg()
    from cuda_mapreduce_b200 import multi
    cuts = multi.line_aligned_cuts(data, world)
    lo, hi = cuts[rank], cuts[rank + 1]
    eng = HostEngine(pkg, data[lo:hi], lo)
    job, is_root = multi.exchange_and_gather(eng, torch.device("cpu"), mode=mode)
    if is_root:
        words = sorted(eng.words.items(), key=lambda kv: kv[1][1])
        res = pkg.wc.debug_finalize([w for w, _ in words], [v[0] for _, v in words], [v[1] for _, v in words],
                                    pkg.MODE_REF, job.n_empty, job.lost_empty if job.p0_found else job.n_empty)
        q.put((res.rows, res.total, job.n_flush, job.n_empty, job.stop, job.bytes))
    dist.barrier()
    dist.destroy_process_group()


def run_world(data, world=2, mode="auto"):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, data, q, mode)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


def check(data, world=2, modes=("direct", "owner")):
    """Both exchange protocols: entries straight to the root, and all-to-all by key owner first."""
    exp = O.fast(data)
    for mode in modes:
        rows, total, n_flush, n_empty, stop, nbytes = run_world(data, world, mode)
        assert rows == exp.pairs(), mode
        assert total == exp.total
        assert (n_flush, n_empty) == (exp.n_flush, exp.n_empty)
        assert (stop if stop != NONE else nbytes) == exp.stop and nbytes == len(data)


def test_two_ranks_plain_text():
    rng = random.Random(41)
    data = random_text(rng, 4000, weird=0.0).replace(b"\n\n", b"\n")
    check(data)


def test_two_ranks_with_long_words_and_leading_empties():
    rng = random.Random(42)
    body = random_text(rng, 3000, weird=0.0).replace(b"\n\n", b"\n")
    data = b"   \n  \n" + body + b" " + b"x" * 40 + b" tail words " + b"y" * 17 + b"\n" + body[:500]
    check(data)


def test_two_ranks_stop_in_first_and_in_second_shard():
    rng = random.Random(43)
    body = random_text(rng, 3000, weird=0.0).replace(b"\n\n", b"\n")
    check(body[:2000] + b"\n\n" + body[2000:], modes=("owner",))     # blank line in shard 0: shard 1 is ignored
    k = body.index(b"\n", len(body) * 3 // 4) + 1
    check(body[:k] + b"\n" + body[k:], modes=("direct",))            # blank line in shard 1
    check(b" \n" * 2000 + body, modes=("owner",))                     # shard 0 holds only empty words
    check(body[:2000] + b"\n\n" + body[2000:], modes=("direct",))


def test_shard_planning_and_stats_reduce():
    pkg = load_pkg()
    from cuda_mapreduce_b200 import multi
    total, block = 10 ** 9, 1 << 20
    edges = [multi.shard_range(total, block, r, 8) for r in range(8)]
    assert edges[0][0] == 0 and edges[-1][1] == total
    assert all(edges[i][1] == edges[i + 1][0] for i in range(7)) and all(lo % block == 0 for lo, _ in edges)
    # a gap between shards is refused
    a, b = pkg.ShardStats(), pkg.ShardStats()
    a.base, a.bytes, b.base, b.bytes = 0, 100, 101, 50
    arr = (pkg.ShardStats * 2)(a, b)
    assert pkg.lib().wc_shard_stats_reduce(arr, 2, C.byref(pkg.ShardStats())) != 0
