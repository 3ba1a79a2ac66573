"""Multi-GPU word count: one engine per rank over a byte-range shard, then the key-owner exchange
(SURVEY.md 8(e)).  The reference has a single GPU and nothing to mirror here (main.cu:133-162 is its
whole driver); the data path stays in libwc.so (export / import kernels, include/wc.h), this module
only moves the exported buffers with torch.distributed (NCCL all-to-all over NVLink on GPUs, gloo in
the CPU tests).

An "engine" is anything with the wc.Engine methods used below, so the CPU tests can drive the same
code with a host-memory stand-in.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import wc

NONE = (1 << 64) - 1
REC = 32  # bytes per exported record (wc.h: WC_WORD_RECORD_BYTES == WC_LONG_RECORD_BYTES)


def shard_range(total, block, rank, world):
    """Block-aligned [lo, hi) of this rank's shard.  Corpus blocks end with a newline, so every shard
    edge is a line start and no word or line state crosses it."""
    nblocks = (total + block - 1) // block
    b0 = nblocks * rank // world
    b1 = nblocks * (rank + 1) // world
    return b0 * block, min(b1 * block, total)


def line_aligned_cuts(data, world):
    """Cut points for arbitrary text: the first line start at or after each nominal boundary."""
    n = len(data)
    cuts = [0]
    for r in range(1, world):
        want = max(n * r // world, cuts[-1])
        nl = data.find(b"\n", want)
        cuts.append(n if nl < 0 else nl + 1)
    cuts.append(n)
    return cuts


STAT_FIELDS = ("base", "bytes", "n_flush", "n_empty", "lost_empty", "p0", "stop", "p0_found")


def stats_to_tensor(st, device):
    vals = [getattr(st, f) for f in STAT_FIELDS]
    return torch.tensor([v if v < (1 << 63) else v - (1 << 64) for v in vals], dtype=torch.int64, device=device)


def tensor_to_stats(t):
    st = wc.ShardStats()
    for f, v in zip(STAT_FIELDS, t.tolist()):
        setattr(st, f, v & NONE)
    return st


def reduce_stats(shards):
    """Job-wide counters from every shard's (host logic in libwc.so: wc_shard_stats_reduce)."""
    arr = (wc.ShardStats * len(shards))(*shards)
    job = wc.ShardStats()
    rc = wc.lib().wc_shard_stats_reduce(arr, len(shards), C.byref(job))
    if rc != 0:
        raise wc.WcError(rc, "wc_shard_stats_reduce: shards do not tile the input")
    return job


def _a2a(send, send_counts, device):
    """all_to_all of variable-size byte buffers; returns (recv buffer, recv byte counts)."""
    world = dist.get_world_size()
    sc = torch.tensor(send_counts, dtype=torch.int64, device=device)
    rc = torch.empty(world, dtype=torch.int64, device=device)
    dist.all_to_all_single(rc, sc)
    rcl = rc.tolist()
    recv = torch.empty(max(sum(rcl), 1), dtype=torch.uint8, device=device)[: sum(rcl)]
    dist.all_to_all_single(recv, send[: sum(send_counts)], output_split_sizes=rcl, input_split_sizes=list(send_counts))
    return recv, rcl


def exchange_and_gather(engine, device, root=0):
    """Runs steps 1-5 of the multi-GPU protocol (include/wc.h).  Returns (job_stats, is_root); on the
    root the engine then holds the whole job's tables and counters and wc_finish gives the rows."""
    world, rank = dist.get_world_size(), dist.get_rank()
    # 1. shard counters -> job counters; shards at or after STOP contribute nothing
    mine = engine.shard_stats()
    allst = [torch.empty(len(STAT_FIELDS), dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(allst, stats_to_tensor(mine, device))
    job = reduce_stats([tensor_to_stats(t) for t in allst])
    if job.stop != NONE and mine.base >= job.stop and mine.bytes > 0:
        engine.clear_tables()
    # words of 17+ bytes leave through their own (small) buffers before the tables are cleared
    counts, n_long, long_bytes = engine.export_counts(world)
    lmeta = torch.empty(max(n_long * REC, 1), dtype=torch.uint8, device=device)
    lblob = torch.empty(max(long_bytes, 8), dtype=torch.uint8, device=device)
    if n_long:
        engine.export_long(lmeta.data_ptr(), lblob.data_ptr())
    # 2. word records grouped by owner, all-to-all
    offsets, acc = [], 0
    for c in counts:
        offsets.append(acc); acc += c
    send = torch.empty(max(acc * REC, 1), dtype=torch.uint8, device=device)
    if acc:
        engine.export_words(world, offsets, send.data_ptr())
    recv, rcl = _a2a(send, [c * REC for c in counts], device)
    # 3. every word is now reduced on its owner
    engine.clear_tables()
    engine.import_words(recv.data_ptr(), sum(rcl) // REC)
    # 4. gather the owners' tables (and the long words) on the root
    owned, _, _ = engine.export_counts(1)
    n_owned = owned[0]
    out = torch.empty(max(n_owned * REC, 1), dtype=torch.uint8, device=device)
    to_root = [0] * world
    if rank != root and n_owned:
        engine.export_words(1, [0], out.data_ptr())
        to_root[root] = n_owned * REC
    got, gcl = _a2a(out, to_root, device)
    lm_send, lb_send = [0] * world, [0] * world
    lm_send[root], lb_send[root] = n_long * REC, long_bytes
    gmeta, gml = _a2a(lmeta, lm_send, device)
    gblob, gbl = _a2a(lblob, lb_send, device)
    if rank == root:
        engine.import_words(got.data_ptr(), sum(gcl) // REC)
        mo = bo = 0
        for r in range(world):  # blob offsets are relative to each sender's blob
            if gml[r]:
                engine.import_long(gmeta.data_ptr() + mo, gml[r] // REC, gblob.data_ptr() + bo)
            mo += gml[r]; bo += gbl[r]
        # 5. the root's counters become the job's
        engine.set_job_stats(job)
    return job, rank == root
