#!/usr/bin/env python3
"""bench.py -- word-count throughput of libwc.so on BASELINE.json's configs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload config2|config3|config4|config5]

A "step" is one whole word count of the workload: reset the device tables, tokenize + count every
byte, leave exact per-word counts in HBM (N > 1: plus the gather of every rank's table on rank 0).
Prints ONE JSON line (contract in the task statement):

  value     input GB/s (1e9 B/s) over the K timed steps, text already resident in HBM, CUDA events on
            the engine's stream, max over ranks.
  e2e       the same metric through the reference-facing C ABI with HOST buffers: wc_reset + wc_feed
            (pinned host text, chunked cudaMemcpyAsync overlapped with the kernel) + wc_finish (the
            reference's reduce on the device, rows D2H).  Wall clock around synchronised calls.
  roofline  the tokenizer/count kernel against the measured HBM copy bandwidth (MEASURED_PEAKS.json);
            algorithmic bytes = input bytes, read once.
  cpu_baseline  the oracle (a single-threaded C port of main.cu's semantics) timed on this box.
  result_check  rows / total / SHA-256 of the job's "word<TAB>count" rows, and oracle_match: the same
            digest computed by the multi-threaded oracle (oracle/) over the same bytes on rank 0's host
            cores, outside every timed region.
  config3_strong / config4_strong  (N > 1, default workload only) north_star's strong-scaling corpora,
            a few steps each, with their own oracle_match.

--impl reference times the reference semantics on host cores (oracle port, all threads): the
reference's own main.cu cannot process more than 990 bytes (main.cu:12,175), so its CPU path at
this size is the capacity-lifted restatement in oracle/.
"""
import argparse
import hashlib
import json
import re
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "word-count input throughput (tokenize+count, device-timed)"
UNIT = "GB/s"


def load_pkg():
    import __graft_entry__ as ge
    return ge._load_pkg()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples nvidia-smi clocks and throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, val in zip(names, r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def workload_params(name, n_gpus):
    from tools.corpus import CONFIGS
    if name == "config2":
        per_gpu = CONFIGS[2]["bytes"]  # 1e9 B per GPU (weak scaling for N > 1)
        return dict(config=2, total=per_gpu * n_gpus, per_gpu=per_gpu, s_log2=18, l_log2=12,
                    label="config2: Zipf(1.1) English-like text, 50k-word vocabulary, 1e9 B per GPU")
    if name == "config3":
        total = CONFIGS[3]["bytes"]
        return dict(config=3, total=total, per_gpu=total // n_gpus, s_log2=22, l_log2=14,
                    label="config3: Zipf(1.1) text, 1M-word vocabulary, 1.6e10 B over all GPUs (strong scaling)")
    if name == "config4":
        total = CONFIGS[4]["bytes"]
        return dict(config=4, total=total, per_gpu=total // n_gpus, s_log2=22, l_log2=14,
                    label="config4: Zipf(1.1) text, 1M-word vocabulary, 1e11 B over all GPUs (strong scaling)")
    if name == "config5":
        per_gpu = CONFIGS[5]["bytes"]
        return dict(config=5, total=per_gpu * n_gpus, per_gpu=per_gpu, s_log2=30, l_log2=16,
                    label="config5: adversarial text, ~90 % unique 8-16 byte tokens, 4e9 B per GPU")
    raise SystemExit("unknown workload " + name)


def rows_digest(rows):
    """SHA-256 over "word<TAB>count<LF>" of every row, in output order."""
    h = hashlib.sha256()
    for r in rows:
        h.update(r[0] + b"\t%d\n" % r[1])
    return h.hexdigest()


def oracle_digest(config, total, host_view=None, live=False):
    """The whole job counted by the oracle (orc_fast_mt, all host threads): the text is regenerated on
    the host from the same seeds (or taken from host_view when the caller already holds all of it).
    Returns a dict, or a reason string if the host cannot hold the text."""
    import numpy as np
    from oracle import orc
    from tools.corpus import Corpus
    t0 = time.perf_counter()
    fixture = os.path.join(ROOT, "tests", "golden", "config%d_full_oracle.json" % config)
    if os.path.exists(fixture) and not live:  # counted ahead of time (tests/golden/make_config_fixtures.py): config 5 takes the oracle minutes
        fx = json.load(open(fixture))
        if fx["bytes"] == total:
            return {"rows": fx["rows"], "total": fx["total"], "n_flush": fx["n_flush"], "stop": fx["stop"], "rows_sha256": None,
                    "checksum": fx["checksum"], "seconds": None, "threads": None,
                    "source": "tests/golden/config%d_full_oracle.json" % config}
    if host_view is None:
        try:
            avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
        except Exception:
            avail = 0
        if avail and avail < total * 1.15 + (4 << 30):
            return "host RAM: %.0f GB available, %.0f GB needed" % (avail / 1e9, total * 1.15 / 1e9)
        c = Corpus(config, total_bytes=total)
        host_view = np.empty(total, dtype=np.uint8)
        step = 1 << 32  # (ctypes-sized pieces; block aligned: the generator works on whole blocks)
        step -= step % c.block
        for off in range(0, total, step):
            c.fill(host_view[off:off + step], off, min(step, total - off))
    t1 = time.perf_counter()
    keep = config != 5  # (2.9e8 rows there: compare the counters only)
    r = orc.fast_mt(host_view, keep_rows=keep)
    out = {"rows": r.n_rows, "total": r.total, "n_flush": r.n_flush, "stop": r.stop, "checksum": r.checksum,
           "source": "orc_fast_mt on this host, live",
           "rows_sha256": rows_digest(r.rows) if keep else None,
           "seconds": {"generate": round(t1 - t0, 2), "count": round(time.perf_counter() - t1, 2)},
           "threads": os.cpu_count()}
    return out


def run_reference(args):
    """--impl reference: the reference semantics on host cores (oracle port, all threads), on the same
    workload (every byte of it, unless K steps of that would run for more than about four minutes)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import orc
    from tools.corpus import Corpus
    wp = workload_params(args.workload, 1)
    c = Corpus(wp["config"], total_bytes=wp["per_gpu"])
    sample = wp["per_gpu"]
    data = c.bytes(0, sample)
    threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    orc.fast_mt(data, nthreads=threads, keep_rows=False)
    one = time.perf_counter() - t0
    budget = 240.0
    if one * (args.steps + args.warmup) > budget:  # a bounded sample: a prefix of whole blocks
        sample = max(c.block, int(sample * budget / (one * (args.steps + args.warmup))) // c.block * c.block)
        data = data[:sample]
    for _ in range(max(0, args.warmup - 1)):
        orc.fast_mt(data, nthreads=threads, keep_rows=False)
    t0 = time.perf_counter()
    words = 0
    for _ in range(args.steps):
        r = orc.fast_mt(data, nthreads=threads, keep_rows=False)
        words = r.n_flush
    dt = (time.perf_counter() - t0) / args.steps
    val = sample / dt / 1e9
    whole = sample == wp["per_gpu"]
    what = ("all %d bytes of the workload" % sample) if whole else ("first %d bytes of the workload" % sample)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wp["label"], "bytes_per_gpu": wp["per_gpu"], "sample": what + " per step"},
        "words_per_s": words / dt,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "%s, orc_fast_mt on %d threads" % (what, threads)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def ncu_traffic(workload, nbytes):
    """dram__bytes_read + dram__bytes_write of one forward count_kernel launch, from the newest ncu
    --set full capture committed under profiles/ (tools/ncu_summary.py); only for the workload and
    size that capture was taken on."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_v*_traffic.json")),
                   key=lambda f: [int(x) for x in re.findall(r"\d+", os.path.basename(f))])
    if not files or workload != "config2" or nbytes != 1_000_000_000:
        return {}, None
    d = json.load(open(files[-1]))
    return d, "profiles/" + os.path.basename(files[-1])


class Ctx:
    """What every workload run shares: torch / dist state of this rank."""

    def __init__(self, args):
        import torch
        self.torch = torch
        self.args = args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, self.world))
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.pkg = load_pkg()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def allreduce(self, vals, op):
        t = self.torch.tensor(vals, dtype=self.torch.float64, device="cuda")
        if self.dist:
            self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return t.tolist()


def run_workload(cx, workload, steps, warmup, primary):
    """Times one workload on all ranks.  Returns the fields of the JSON line that describe it (rank 0)."""
    torch, args, world, rank, local_rank = cx.torch, cx.args, cx.world, cx.rank, cx.local_rank
    from tools.corpus import Corpus
    from cuda_mapreduce_b200 import multi
    wp = workload_params(workload, world)
    if args.bytes and primary:
        wp["per_gpu"] = args.bytes
        wp["total"] = args.bytes * world
        wp["label"] += " (size overridden: %d B per GPU)" % args.bytes
    if args.table_log2 and primary:
        wp["s_log2"] = args.table_log2
    path = args.path if primary else "auto"
    if path == "sort":
        wp["s_log2"] = 12

    # ---- this rank's shard of the synthetic corpus
    corpus = Corpus(wp["config"], total_bytes=wp["total"])
    lo, hi = multi.shard_range(wp["total"], corpus.block, rank, world)
    nbytes = hi - lo
    device_corpus = (args.device_corpus and primary) or nbytes > (9 << 29) or not primary
    host = None
    shm = None
    # (the 100 GB corpus is always run that way when it is a sub-run of the default workload: "sharded across 8 x B200
    # with streamed host->device chunks" is how BASELINE.json states config 4)
    want_shm = wp["config"] != 5 and ((args.host_stream and primary) or (workload == "config4" and not primary and not args.no_e2e))
    if want_shm:  # only if /dev/shm and the host's memory can hold the corpus (decided by the worst rank's view)
        try:
            vfs = os.statvfs("/dev/shm")
            avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
            fits = float(vfs.f_bavail * vfs.f_frsize >= wp["total"] * 1.02 and avail >= wp["total"] * 1.25 + (16 << 30))
        except Exception:
            fits = 0.0
        want_shm = cx.allreduce([fits], "MIN")[0] > 0.5
    if want_shm:
        # the corpus sits in HOST memory and is streamed to the GPUs in chunks (BASELINE.json configs[3]): one
        # shared mapping for all ranks -- every rank generates and pins its own shard of it, and rank 0 later
        # counts the whole of it with the oracle
        import numpy as np
        shm = "/dev/shm/wc_bench_%s_%d.bin" % (workload, wp["total"])
        if rank == 0:
            with open(shm, "wb") as f:
                f.truncate(wp["total"])
        cx.barrier()
        whole_map = np.memmap(shm, dtype=np.uint8, mode="r+", shape=(wp["total"],))
        corpus.fill(whole_map[lo:hi], lo, nbytes, nthreads=max(1, (os.cpu_count() or 1) // world))
        rt = torch.cuda.cudart()
        rc_reg = rt.cudaHostRegister(whole_map.ctypes.data + lo, nbytes, 0)
        shm_pinned = int(rc_reg) == 0  # (if the shard cannot be pinned wc_feed stages it through its own pinned slots)
        if not shm_pinned:
            print("bench: cudaHostRegister of the shard failed (%s): feeding pageable memory" % rc_reg, file=sys.stderr)
        host = torch.from_numpy(whole_map[lo:hi])
        dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        dev.copy_(host)
        device_corpus = False
        no_e2e = args.no_e2e
    elif device_corpus and wp["config"] != 5:
        # configs 3 / 4: the shard is written by tools/corpus_gpu.cu (byte-identical to the host generator,
        # tests/test_gpu_parity.py); there is no host copy, so no e2e leg
        dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        corpus.fill_device(dev.data_ptr(), lo, nbytes)
        no_e2e = True
    else:
        device_corpus = False
        host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        corpus.fill(host.numpy(), lo, nbytes)
        dev = host.cuda(non_blocking=False)
        no_e2e = args.no_e2e or not primary

    eng = cx.pkg.Engine(device=local_rank, short_table_log2=wp["s_log2"], long_table_log2=wp["l_log2"], base_offset=lo,
                        path={"auto": 0 if world == 1 else 1, "hash": 1, "sort": 2}[path], strict_ref=args.strict_ref and primary,
                        lane_serial=args.lane_serial)
    ext = torch.cuda.ExternalStream(eng.stream(), device=cx.dev)
    # N > 1: table entries go to rank 0 through NVLink peer memory (wc_peer_*): one push kernel per sender, the
    # root waits for the headers on the device; --exchange nccl uses the all-to-all protocol instead.
    gather = None
    if world > 1 and args.exchange in ("peer", "peer-host"):
        gather = multi.PeerGather(eng, cx.dev, 1 << wp["s_log2"], 1 << wp["l_log2"], 64 << 20,
                                  device_wait=args.exchange == "peer")
    exchange_mode = {"nccl": "auto", "nccl-owner": "owner", "nccl-direct": "direct"}.get(args.exchange, "auto")

    def exchange():
        with torch.cuda.stream(ext):
            return gather.run() if gather else multi.exchange_and_gather(eng, cx.dev, mode=exchange_mode)

    def step_device():
        eng.reset()
        eng.count_device(dev.data_ptr(), nbytes, last_piece=(rank == world - 1))
        if path == "sort":
            eng.aggregate()  # radix sort + run reduce: the sort path's counterpart of the hash tables
        if world > 1:
            exchange()

    # ---- value: device-resident, CUDA events on the launching stream
    for _ in range(warmup):
        step_device()
    eng.sync()
    cx.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0 and primary:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(ext)
    for _ in range(steps):
        step_device()
    ev1.record(ext)
    eng.sync()
    cx.barrier()
    ms_total = ev0.elapsed_time(ev1)
    # the tokenizer kernel's own average duration: K more steps, reading each step's event pair
    ksum, launches = 0.0, 0
    for _ in range(steps):
        step_device()
        eng.sync()
        _, kms, nl = eng.device_ms()
        ksum += kms
        launches = nl
    kernel_ms = ksum / steps
    clocks = sampler.stop() if (rank == 0 and primary) else None
    cx.barrier()

    # ---- the whole job's result once more, outside any timed region: row count, total and a digest of
    # the rows, checked against the oracle's count of the same bytes
    result_check = None
    want_rows = wp["config"] != 5  # (2.9e8 rows there; decided the same way on every rank)
    finish_ms = None
    for _ in range(2):  # (the first wc_finish of an engine allocates its buffers: time the second)
        step_device()
        if rank == 0:
            eng.sync()
            t0 = time.perf_counter()
            full = eng.finish(keep_rows=False)  # the reference's reduce on the device + rows to the host (outside `value`)
            finish_ms = (time.perf_counter() - t0) * 1e3
            del full
        cx.barrier()
    step_device()
    if rank == 0:
        full = eng.finish(keep_rows=want_rows)
        result_check = {"rows": full.n_rows, "total": full.total, "n_flush": full.stats["n_flush"], "stop": full.stats["stop"],
                        "path_used": full.stats.get("path_used"), "tokenizing_warps": full.stats.get("tokenizing_warps"),
                        "checksum": full.checksum,
                        "rows_sha256": rows_digest(full.rows) if want_rows else None}
        del full
    cx.barrier()
    if rank == 0 and not args.no_oracle:
        whole = whole_map if shm else (host.numpy() if (host is not None and world == 1) else None)
        od = oracle_digest(wp["config"], wp["total"], whole, live=args.oracle_live)
        if isinstance(od, str):
            result_check["oracle_match"] = None
            result_check["oracle_skipped"] = od
        else:
            keys = ["rows", "total", "n_flush", "stop", "checksum"] + (["rows_sha256"] if od["rows_sha256"] else [])
            result_check["oracle_match"] = all(result_check[k] == od[k] for k in keys)
            result_check["oracle_rows_sha256"] = od["rows_sha256"]
            result_check["oracle"] = {k: od[k] for k in ("rows", "total", "checksum", "seconds", "threads", "source")}
    cx.barrier()
    eng.reset()
    eng.count_device(dev.data_ptr(), nbytes, last_piece=(rank == world - 1))
    res = eng.finish(keep_rows=False)  # this rank's shard alone: its words for the words/s figure
    words = res.stats["n_flush"]

    ms_step = cx.allreduce([ms_total], "MAX")[0] / steps
    all_bytes, all_words = cx.allreduce([float(nbytes), float(words)], "SUM")
    value = all_bytes / (ms_step * 1e-3) / 1e9

    # ---- e2e: host buffers through the C ABI, H2D inside the timed region, rows back on the host
    def step_e2e():
        eng.reset()
        eng.feed((host.data_ptr(), nbytes))
        if world > 1:
            _, is_root = exchange()
            if not is_root:
                eng.sync()
                return None
        return eng.finish(keep_rows=False)

    e2e = None
    if not no_e2e:
        e2e_steps = min(50, max(3, steps // 2))
        # what the platform gives for the copies alone: every rank's pinned shard -> its GPU, all ranks at once
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            dev.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        h2d_s = cx.allreduce([(time.perf_counter() - t0) / 3], "MAX")[0]
        step_e2e()
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step_e2e()
        torch.cuda.synchronize()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
        e2e_s = cx.allreduce([e2e_s], "MAX")[0]
        cx.barrier()
        e2e = {"value": all_bytes / e2e_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": eng.finish_d2h_bytes(), "ms_per_step": e2e_s * 1e3,
               "h2d_ceiling_gbs": all_bytes / h2d_s / 1e9, "frac_of_ceiling": h2d_s / e2e_s,
               "h2d_ceiling": "plain cudaMemcpyAsync of every rank's pinned shard to its GPU, all ranks at the same time"}
    stats = res.stats
    cx.barrier()  # (the root's landing areas are mapped by the other ranks until they are done)
    gather = None
    eng.close()
    if shm:
        if shm_pinned:
            torch.cuda.cudart().cudaHostUnregister(whole_map.ctypes.data + lo)
        host = None
        del whole_map
        cx.barrier()
        if rank == 0:
            os.remove(shm)
    del dev
    torch.cuda.empty_cache()
    if rank != 0:
        return None, None
    exch = "none"
    if world > 1:
        exch = ("gather on rank 0 through NVLink peer memory (one push kernel per sender; "
                + ("the root waits for the headers on the device, no collective)" if args.exchange == "peer"
                   else "one-word NCCL all-reduce as barrier, host wait on the root)")
                if (gather and gather.available) else "key-owner exchange (NCCL all-to-all, owner reduce, gather on rank 0)")
    out = {
        "value": value, "ms_per_step": ms_step, "words_per_s": all_words / (ms_step * 1e-3),
        "config": {"workload": wp["label"], "bytes_per_gpu": nbytes, "words_per_gpu": words,
                   "distinct_words_this_shard": stats["n_distinct"], "deferred_1to8B_words": stats["n_spilled"],
                   "words_9B_plus": stats["n_long_words"],
                   "l2": "input is larger than the 126 MB L2, every step re-reads it from HBM",
                   "step": "wc_reset (table memset) + tokenize/count kernel + roll-back/summary kernel"
                           + (" + " + exch if world > 1 else ""),
                   "exchange": exch, "vocab_seed": corpus.vocab_seed, "text_seed": corpus.text_seed},
        "data": "synthetic" + (" (generated on the device)" if device_corpus else "")
                + (" (generated into pinned host memory, streamed to the GPUs by wc_feed)" if shm else ""),
        "kernel_ms": kernel_ms, "nbytes": nbytes, "e2e": e2e, "result_check": result_check, "finish_ms": finish_ms,
        "gpu_launches": int(launches) * steps, "clocks": clocks,
    }
    return out, host


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="config2")
    ap.add_argument("--bytes", type=int, default=0, help="override the per-GPU input size (profiling / sweeps)")
    ap.add_argument("--path", default="auto", choices=["auto", "hash", "sort"], help="aggregation path (wc_config.path)")
    ap.add_argument("--table-log2", type=int, default=0, help="override the word table size")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: skip the host-buffer leg")
    ap.add_argument("--oracle-live", action="store_true", help="count the job with the oracle on this host even where tests/golden holds its result (config 5)")
    ap.add_argument("--no-oracle", action="store_true", help="profiling runs: skip the oracle comparison of the result")
    ap.add_argument("--no-extras", action="store_true", help="N > 1: skip the config3 / config4 strong-scaling sub-runs")
    ap.add_argument("--device-corpus", action="store_true",
                    help="generate the corpus straight into device memory (default for shards above 4.5 GiB); no e2e leg")
    ap.add_argument("--host-stream", action="store_true",
                    help="keep the corpus in pinned HOST memory (one shared mapping under /dev/shm, a shard per rank) and report "
                         "e2e: streamed host->device chunks through wc_feed (config 4 as BASELINE.json states it)")
    ap.add_argument("--strict-ref", action="store_true",
                    help="WC_FLAG_STRICT_REF: read lines in 99-byte pieces like the reference's fgets (same results on these corpora)")
    ap.add_argument("--lane-serial", action="store_true", help="WC_FLAG_LANE_SERIAL: the alternative forward kernel (count_kernel2)")
    ap.add_argument("--exchange", choices=["peer", "peer-host", "nccl", "nccl-owner", "nccl-direct"], default="peer",
                    help="N > 1: gather through NVLink peer memory (default: device-side wait; peer-host: all-reduce barrier "
                         "+ host wait) or the NCCL protocol (auto / key-owner all-to-all / direct gather)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    cx = Ctx(args)
    main_out, host = run_workload(cx, args.workload, args.steps, args.warmup, primary=True)
    extras = {}
    if cx.world > 1 and args.workload == "config2" and not args.no_extras and not args.bytes:
        # north_star's own scaling curves: the 16 GB corpus on every N, the 100 GB corpus on 8 GPUs
        subs = [("config3_strong", "config3")] + ([("config4_strong", "config4")] if cx.world == 8 else [])
        for key, wl in subs:
            sub, _ = run_workload(cx, wl, 10, 3, primary=False)
            if cx.rank == 0:
                extras[key] = {"value": sub["value"], "unit": UNIT, "ms_per_step": sub["ms_per_step"], "steps": 10, "warmup": 3,
                               "kernel_ms": sub["kernel_ms"], "bytes_total": sub["nbytes"] * cx.world, "scaling": "strong",
                               "exchange": sub["config"]["exchange"], "oracle_match": sub["result_check"].get("oracle_match"),
                               "e2e": sub["e2e"],
                               "result_check": sub["result_check"], "data": sub["data"]}
    if cx.rank != 0:
        return
    o = main_out
    peak, peak_src = measured_peak()
    ncu, traffic_src = ncu_traffic(args.workload, o["nbytes"])
    traffic = ncu.get("traffic_bytes_per_launch")
    winst, wave = ncu.get("warp_instructions_per_launch"), ncu.get("shared_wavefronts_per_launch")
    achieved = o["nbytes"] / (o["kernel_ms"] * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": o["value"], "unit": UNIT, "n_gpus": cx.world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": o["ms_per_step"], "higher_is_better": True,
        "scaling": "weak" if args.workload in ("config2", "config5") else "strong", "vs_baseline": None, "dtype": "u8",
        "data": o["data"], "config": o["config"], "words_per_s": o["words_per_s"],
        "frac_of_hbm_nominal_8TBs": o["value"] / cx.world / 8000.0,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "kernel": "wc::count_kernel2 (forward launch)" if args.lane_serial else "wc::count_kernel (forward launch)",
                     "kernel_ms": o["kernel_ms"], "algorithmic_bytes_per_launch": o["nbytes"],
                     # from the same ncu capture as `traffic`: lane-instructions and shared-memory wavefronts per input byte
                     "instr_per_byte": (winst * 32 / o["nbytes"]) if winst else None,
                     "shared_wavefronts_per_word": (wave / o["config"]["words_per_gpu"]) if wave else None},
        "e2e": o["e2e"], "result_check": o["result_check"], "gpu_launches": o["gpu_launches"], "clocks": o["clocks"],
        # `value` stops where the exact per-word counts are in HBM; the reference's reduce half (prefix merge, row order)
        # and the rows' way to the host are wc_finish: this many ms on top of a step (it is inside `e2e`)
        "finish_ms": o["finish_ms"],
    }
    line.update(extras)
    if not args.no_cpu_baseline and cx.world == 1 and host is not None:  # (the CPU baseline is timed at N = 1 only)
        from oracle import orc
        sample = min(o["nbytes"], 1 << 30, host.numel())
        view = host.numpy()[:sample]
        t0 = time.perf_counter()
        rb = orc.fast(view, keep_rows=False)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": sample / dt / 1e9, "unit": UNIT, "cores": 1, "kind": "port",
                                "sample": "first %d bytes of the workload, single-threaded orc_fast (%.1f s)" % (sample, dt),
                                "words_per_s": rb.n_flush / dt, "host_cores_available": os.cpu_count()}
    print(json.dumps(line), flush=True)


def _main_and_cleanup():
    try:
        main()
    finally:
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                dist.destroy_process_group()
        except Exception:
            pass


if __name__ == "__main__":
    _main_and_cleanup()
