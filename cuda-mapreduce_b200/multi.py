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


def _a2a(send, send_counts, recv_counts, device):
    """all_to_all of variable-size byte buffers whose sizes both sides already know."""
    recv = torch.empty(max(sum(recv_counts), 1), dtype=torch.uint8, device=device)[: sum(recv_counts)]
    dist.all_to_all_single(recv, send[: sum(send_counts)], output_split_sizes=list(recv_counts),
                           input_split_sizes=list(send_counts))
    return recv


def _gather_ints(vals, device):
    """Every rank's list of integers on every rank: [world][len(vals)]."""
    world = dist.get_world_size()
    mine = torch.tensor([v if v < (1 << 63) else v - (1 << 64) for v in vals], dtype=torch.int64, device=device)
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    return [p.tolist() for p in parts]


# Below this many table entries in the whole job, every rank sends its entries straight to the root
# (one collective); above it they are first reduced on their key owners, so that the root receives
# each distinct word once instead of once per rank.
DIRECT_GATHER_MAX_RECORDS = 2_000_000


def exchange_and_gather(engine, device, root=0, mode="auto"):
    """Runs the multi-GPU protocol (include/wc.h).  Returns (job_stats, is_root); on the root the
    engine then holds the whole job's tables and counters and wc_finish gives the rows.

    One all-gather tells every rank every shard's counters and how many records each rank holds for
    each key owner, so no later collective needs a size exchange.  mode: "auto", "direct" (entries go
    straight to the root) or "owner" (all-to-all by key owner, owner reduce, then gather)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    # 1. shard counters and export sizes, once, to everybody
    mine = engine.shard_stats()
    counts, n_long, long_bytes = engine.export_counts(world)
    nst = len(STAT_FIELDS)
    table = _gather_ints([getattr(mine, f) for f in STAT_FIELDS] + list(counts) + [n_long, long_bytes], device)
    shards = []
    for row in table:
        st = wc.ShardStats()
        for f, v in zip(STAT_FIELDS, row[:nst]):
            setattr(st, f, v & NONE)
        shards.append(st)
    job = reduce_stats(shards)
    # a shard at or after STOP contributes nothing (the reference stopped reading before it)
    dead = [job.stop != NONE and sh.base >= job.stop and sh.bytes > 0 for sh in shards]
    per_owner = [[0 if dead[r] else c for c in table[r][nst:nst + world]] for r in range(world)]  # [sender][owner]
    longs = [(0, 0) if dead[r] else (table[r][nst + world], table[r][nst + world + 1]) for r in range(world)]
    if dead[rank]:
        engine.clear_tables()
        counts, n_long, long_bytes = [0] * world, 0, 0
    total_records = sum(sum(r) for r in per_owner)
    if mode == "auto":
        mode = "direct" if total_records <= DIRECT_GATHER_MAX_RECORDS else "owner"

    # words of 17+ bytes always travel straight to the root, in their own (small) buffers; they leave
    # the tables before step 3 clears them (in "owner" mode the root's own come back in at the end)
    any_long = any(n for n, _ in longs)
    if any_long:
        lmeta = torch.empty(max(n_long * REC, 1), dtype=torch.uint8, device=device)
        lblob = torch.empty(max(long_bytes, 8), dtype=torch.uint8, device=device)
        if n_long and (rank != root or mode == "owner"):
            engine.export_long(lmeta.data_ptr(), lblob.data_ptr())

    def to_root(nbytes_by_rank, buf):
        """Every rank's buffer (its size known to all) to the root; the root's own stays where it is."""
        send = [0] * world
        if rank != root:
            send[root] = nbytes_by_rank[rank]
        recv = [nbytes_by_rank[r] if (rank == root and r != root) else 0 for r in range(world)]
        return _a2a(buf, send, recv, device), recv

    if mode == "direct":
        # 2. each rank's entries, compacted, to the root; the root adds them to its own table
        n_mine = sum(counts)
        out = torch.empty(max(n_mine * REC, 1), dtype=torch.uint8, device=device)
        if rank != root and n_mine:
            engine.export_words(1, [0], out.data_ptr())
        got, gcl = to_root([sum(r) * REC for r in per_owner], out)
        if rank == root and sum(gcl):
            engine.import_words(got.data_ptr(), sum(gcl) // REC)
    else:
        # 2. word records grouped by owner, all-to-all (sizes known from step 1)
        offsets, acc = [], 0
        for c in counts:
            offsets.append(acc); acc += c
        send = torch.empty(max(acc * REC, 1), dtype=torch.uint8, device=device)
        if acc:
            engine.export_words(world, offsets, send.data_ptr())
        recv = _a2a(send, [c * REC for c in counts], [per_owner[r][rank] * REC for r in range(world)], device)
        # 3. every word is now reduced on its owner
        engine.clear_tables()
        engine.import_words(recv.data_ptr(), recv.numel() // REC)
        # 4. gather the owners' tables on the root
        owned, _, _ = engine.export_counts(1)
        owned_all = [r[0] for r in _gather_ints([owned[0]], device)]
        out = torch.empty(max(owned[0] * REC, 1), dtype=torch.uint8, device=device)
        if rank != root and owned[0]:
            engine.export_words(1, [0], out.data_ptr())
        got, gcl = to_root([n * REC for n in owned_all], out)
        if rank == root and sum(gcl):
            engine.import_words(got.data_ptr(), sum(gcl) // REC)

    if any_long:
        gmeta, gml = to_root([n * REC for n, _ in longs], lmeta)
        gblob, gbl = to_root([b for _, b in longs], lblob)
        if rank == root:
            if mode == "owner" and n_long:
                engine.import_long(lmeta.data_ptr(), n_long, lblob.data_ptr())
            mo = bo = 0
            for r in range(world):  # blob offsets are relative to each sender's blob
                if gml[r]:
                    engine.import_long(gmeta.data_ptr() + mo, gml[r] // REC, gblob.data_ptr() + bo)
                mo += gml[r]; bo += gbl[r]
    if rank == root:
        engine.set_job_stats(job)  # 5. the root's counters become the job's
    return job, rank == root


class PeerGather:
    """Single-node gather over NVLink peer memory (include/wc.h, wc_peer_*): set up once per engine,
    then one call per job.  Every non-root rank stores its table entries and shard counters straight
    into the root GPU's memory with one kernel.

    device_wait=True (the default; every rank must have a GPU of its own): nothing else happens on the
    senders, and the root queues a kernel that waits for the senders' headers in its own memory,
    reduces the shard counters there and imports what arrived -- no collective and no host wait in
    the job.  A sender that overflowed makes the root's wc_finish fail with WC_ERR_TOO_LARGE.

    device_wait=False: a one-word NCCL all-reduce behind the push is both the barrier and the "any
    sender overflowed?" vote, the root then waits on the host for the headers; if a sender did
    overflow, all ranks fall back to exchange_and_gather() together."""

    def __init__(self, engine, device, words_per_rank, long_per_rank, blob_bytes_per_rank, root=0, device_wait=True):
        self.engine, self.device, self.root, self.device_wait = engine, device, root, device_wait
        world, rank = dist.get_world_size(), dist.get_rank()
        caps = (words_per_rank, long_per_rank, blob_bytes_per_rank)
        handle = torch.zeros(64, dtype=torch.uint8, device=device)
        ok = torch.ones(1, dtype=torch.int32, device=device)
        self.why = ""
        try:
            if rank == root:
                raw = engine.peer_root_create(world, root, *caps)
                handle.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
        except wc.WcError as ex:
            ok.zero_(); self.why = str(ex)
        dist.broadcast(handle, src=root)
        try:
            if rank != root:
                engine.peer_open(rank, world, root, *caps, bytes(handle.cpu().numpy().tobytes()))
        except wc.WcError as ex:  # no CUDA IPC / no peer access on this machine
            ok.zero_(); self.why = str(ex)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        self.available = int(ok.item()) == 1  # False on every rank if it failed on any: use exchange_and_gather
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)

    def run(self):
        """Call on the engine's stream (torch.cuda.stream(ExternalStream(engine.stream()))).
        Returns (job_stats or None, is_root)."""
        rank = dist.get_rank()
        if not self.available:
            return exchange_and_gather(self.engine, self.device, self.root)
        if self.device_wait:
            if rank != self.root:
                self.engine.peer_push(None)
            else:
                self.engine.peer_pull_device()
            return None, rank == self.root
        self.flag.zero_()
        if rank != self.root:
            self.engine.peer_push(self.flag.data_ptr())
        dist.all_reduce(self.flag, op=dist.ReduceOp.MAX)  # behind the push on this stream: barrier + vote
        # Every rank learns the vote, so that all of them take the fallback together: the senders read
        # the flag, the root sees the same bits in the headers it pulls (one host wait instead of two).
        job = None
        if rank == self.root:
            try:
                job = self.engine.peer_pull()
                bad = False
            except wc.WcError as ex:
                if ex.code != -7:  # WC_ERR_TOO_LARGE: a sender said its tables or its landing area overflowed
                    raise
                bad = True
        else:
            bad = int(self.flag.item()) != 0
        if bad:
            return exchange_and_gather(self.engine, self.device, self.root, mode="owner")
        return job, rank == self.root
