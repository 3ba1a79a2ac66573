#!/usr/bin/env python3
"""bench.py -- word-count throughput of libwc.so on BASELINE.json's configs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload config2|config3]

A "step" is one whole word count of the workload: reset the device tables, tokenize + count every
byte, leave exact per-word counts in HBM.  Prints ONE JSON line (contract in the task statement):

  value     input GB/s (1e9 B/s) over the K timed steps, text already resident in HBM, CUDA events on
            the engine's stream, max over ranks.
  e2e       the same metric through the reference-facing C ABI with HOST buffers: wc_reset + wc_feed
            (pinned host text, chunked cudaMemcpyAsync overlapped with the kernel) + wc_finish (tables
            D2H + the host half of the reduce -> rows).  Wall clock around synchronised calls.
  roofline  the tokenizer/count kernel against the measured HBM copy bandwidth (MEASURED_PEAKS.json);
            algorithmic bytes = input bytes, read once.
  cpu_baseline  the oracle (a single-threaded C port of main.cu's semantics) timed on this box.

--impl reference times the reference semantics on host cores (oracle port, all threads): the
reference's own main.cu cannot process more than 990 bytes (main.cu:12,175), so its CPU path at
this size is the capacity-lifted restatement in oracle/.
"""
import argparse
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
        return dict(config=2, total=per_gpu * n_gpus, per_gpu=per_gpu, s_log2=20, l_log2=12,
                    label="config2: Zipf(1.1) English-like text, 50k-word vocabulary, 1e9 B per GPU")
    if name == "config3":
        total = CONFIGS[3]["bytes"]
        return dict(config=3, total=total, per_gpu=total // n_gpus, s_log2=22, l_log2=14,
                    label="config3: Zipf(1.1) text, 1M-word vocabulary, 1.6e10 B over all GPUs")
    if name == "config4":
        total = CONFIGS[4]["bytes"]
        return dict(config=4, total=total, per_gpu=total // n_gpus, s_log2=22, l_log2=14,
                    label="config4: Zipf(1.1) text, 1M-word vocabulary, 1e11 B over all GPUs (strong scaling)")
    if name == "config5":
        per_gpu = CONFIGS[5]["bytes"]
        return dict(config=5, total=per_gpu * n_gpus, per_gpu=per_gpu, s_log2=30, l_log2=16,
                    label="config5: adversarial text, ~90 % unique 8-16 byte tokens, 4e9 B per GPU")
    raise SystemExit("unknown workload " + name)


def run_reference(args):
    """--impl reference: the reference semantics on host cores (oracle port, all threads)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import orc
    from tools.corpus import Corpus
    wp = workload_params(args.workload, 1)
    sample = 256 << 20
    c = Corpus(wp["config"], total_bytes=wp["per_gpu"])
    data = c.bytes(0, sample)
    threads = os.cpu_count() or 1
    for _ in range(args.warmup):
        orc.fast_mt(data, nthreads=threads, keep_rows=False)
    t0 = time.perf_counter()
    words = 0
    for _ in range(args.steps):
        r = orc.fast_mt(data, nthreads=threads, keep_rows=False)
        words = r.n_flush
    dt = (time.perf_counter() - t0) / args.steps
    val = sample / dt / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wp["label"], "sample": "first %d bytes per step" % sample},
        "words_per_s": words / dt,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "first %d bytes of the workload, orc_fast_mt on %d threads" % (sample, threads)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def ncu_traffic(workload, nbytes):
    """dram__bytes_read + dram__bytes_write of one forward count_kernel launch, from the newest ncu
    --set full capture committed under profiles/ (tools/ncu_summary.py); only for the workload and
    size that capture was taken on."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")),
                   key=lambda f: [int(x) for x in re.findall(r"\d+", os.path.basename(f))])
    if not files or workload != "config2" or nbytes != 1_000_000_000:
        return None, None
    d = json.load(open(files[-1]))
    return d["traffic_bytes_per_launch"], "profiles/" + os.path.basename(files[-1])


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
    ap.add_argument("--device-corpus", action="store_true",
                    help="generate the corpus straight into device memory (default for shards above 4.5 GiB); no e2e leg")
    ap.add_argument("--strict-ref", action="store_true",
                    help="WC_FLAG_STRICT_REF: read lines in 99-byte pieces like the reference's fgets (same results on these corpora)")
    ap.add_argument("--exchange", choices=["peer", "nccl"], default="peer",
                    help="N > 1: gather through NVLink peer memory (default) or the NCCL all-to-all protocol")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch
    from tools.corpus import Corpus

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    pkg = load_pkg()
    from cuda_mapreduce_b200 import multi
    shard_range = multi.shard_range
    wp = workload_params(args.workload, world)
    if args.bytes:
        wp["per_gpu"] = args.bytes
        wp["total"] = args.bytes * world
        wp["label"] += " (size overridden: %d B per GPU)" % args.bytes
    if args.table_log2:
        wp["s_log2"] = args.table_log2
    if args.path == "sort":
        wp["s_log2"] = 12

    # ---- this rank's shard of the synthetic corpus, generated into pinned host memory
    corpus = Corpus(wp["config"], total_bytes=wp["total"])
    lo, hi = shard_range(wp["total"], corpus.block, rank, world)
    nbytes = hi - lo
    device_corpus = args.device_corpus or nbytes > (9 << 29)  # > 4.5 GiB per GPU: generate straight into HBM
    if device_corpus:
        # configs 3 / 4: the shard is written by tools/corpus_gpu.cu (byte-identical to the host generator,
        # tests/test_gpu_parity.py); only a 1 GB sample exists on the host (cpu_baseline), so there is no e2e leg
        torch.cuda.set_device(local_rank)
        dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        corpus.fill_device(dev.data_ptr(), lo, nbytes)
        hs = min(nbytes, 1 << 30)
        host = torch.empty(hs, dtype=torch.uint8).pin_memory()
        if world == 1 and not args.no_cpu_baseline:
            corpus.fill(host.numpy(), lo, hs)
        args.no_e2e = True
    else:
        host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        corpus.fill(host.numpy(), lo, nbytes)
        dev = host.cuda(non_blocking=False)

    eng = pkg.Engine(device=local_rank, short_table_log2=wp["s_log2"], long_table_log2=wp["l_log2"], base_offset=lo,
                     path={"auto": 0, "hash": 1, "sort": 2}[args.path], strict_ref=args.strict_ref)
    ext = torch.cuda.ExternalStream(eng.stream(), device=torch.device("cuda", local_rank))
    # N > 1: table entries go to rank 0 through NVLink peer memory (wc_peer_*), one kernel per sender and
    # a one-word all-reduce as the barrier; --exchange nccl uses the all-to-all protocol instead.
    gather = None
    if world > 1 and args.exchange == "peer":
        gather = multi.PeerGather(eng, torch.device("cuda", local_rank), 1 << wp["s_log2"], 1 << wp["l_log2"], 64 << 20)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    devt = torch.device("cuda", local_rank)

    def step_device():
        eng.reset()
        eng.count_device(dev.data_ptr(), nbytes, last_piece=(rank == world - 1))
        if args.path == "sort":
            eng.aggregate()  # radix sort + run reduce: the sort path's counterpart of the hash tables
        if world > 1:  # key-owner exchange: NCCL all-to-all, owner reduce, gather on rank 0
            with torch.cuda.stream(ext):
                gather.run() if gather else multi.exchange_and_gather(eng, devt)

    # ---- value: device-resident, CUDA events on the launching stream
    for _ in range(args.warmup):
        step_device()
    eng.sync()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms, launches = 0.0, 0
    ev0.record(ext)
    for _ in range(args.steps):
        step_device()
        # per-step kernel times are read after the timed region; wc_reset would drop them, so
        # collect them here without synchronising: they are summed by wc_sync at the end of the loop
    ev1.record(ext)
    eng.sync()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    # the tokenizer kernel's own average duration: re-run K steps, reading each step's event pair
    ksum = 0.0
    for _ in range(args.steps):
        step_device()
        eng.sync()
        _, kms, nl = eng.device_ms()
        ksum += kms
        launches = nl
    kernel_ms = ksum / args.steps
    clocks = sampler.stop() if rank == 0 else None

    # ---- the whole job's result once more, outside any timed region: row count, total and a digest of
    # the rows, so that runs of the same workload on different GPU counts can be compared line by line
    result_check = None
    want_digest = args.workload != "config5"  # (2.9e8 rows there; decided the same way on every rank)
    step_device()
    if rank == 0:
        full = eng.finish(keep_rows=False)
        result_check = {"rows": full.n_rows, "total": full.total, "n_flush": full.stats["n_flush"], "stop": full.stats["stop"]}
    if want_digest:
        step_device()
        if rank == 0:
            import hashlib
            full = eng.finish(keep_rows=True)
            h = hashlib.sha256()
            for w, c in full.rows:
                h.update(w + b"\t%d\n" % c)
            result_check["rows_sha256"] = h.hexdigest()
    if world > 1:
        barrier()
        eng.reset()
        eng.count_device(dev.data_ptr(), nbytes, last_piece=(rank == world - 1))
    res = eng.finish(keep_rows=False)  # this rank's shard alone: its words for the words/s figure
    words = res.stats["n_flush"]

    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(nbytes), float(words)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_step = t.item() / args.steps
    all_bytes, all_words = tot[0].item(), tot[1].item()
    value = all_bytes / (ms_step * 1e-3) / 1e9

    # ---- e2e: host buffers through the C ABI, H2D inside the timed region, rows back on the host
    def step_e2e():
        eng.reset()
        eng.feed((host.data_ptr(), nbytes))
        if world > 1:
            with torch.cuda.stream(ext):
                _, is_root = gather.run() if gather else multi.exchange_and_gather(eng, devt)
            if not is_root:
                eng.sync()
                return None
        return eng.finish(keep_rows=False)

    e2e_steps = 0 if args.no_e2e else min(50, max(3, args.steps // 2))
    if e2e_steps:
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / max(1, e2e_steps)
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = all_bytes / te.item() / 1e9 if e2e_steps else None
    d2h = eng.finish_d2h_bytes()  # what the last wc_finish copied: counters, used entries, their two orders

    if rank != 0:
        return
    peak, peak_src = measured_peak()
    traffic, traffic_src = ncu_traffic(args.workload, nbytes)
    achieved = nbytes / (kernel_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak" if args.workload in ("config2", "config5") else "strong", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic" + (" (generated on the device)" if device_corpus else ""),
        "config": {"workload": wp["label"], "bytes_per_gpu": nbytes, "words_per_gpu": words,
                   "distinct_words": res.stats["n_distinct"], "deferred_1to8B_words": res.stats["n_spilled"],
                   "words_9B_plus": res.stats["n_long_words"],
                   "l2": "input (1e9 B) is larger than the 126 MB L2, every step re-reads it from HBM",
                   "step": "wc_reset (table memset) + tokenize/count kernel + roll-back/summary kernel"
                           + ((" + gather on rank 0 through NVLink peer memory (one push kernel per sender, one-word NCCL all-reduce as barrier)"
                              if (gather and gather.available) else " + key-owner exchange (NCCL all-to-all, owner reduce, gather on rank 0)") if world > 1 else ""),
                   "vocab_seed": corpus.vocab_seed, "text_seed": corpus.text_seed},
        "words_per_s": all_words / (ms_step * 1e-3),
        "frac_of_hbm_nominal_8TBs": value / world / 8000.0,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel": "wc::count_kernel (forward launch)",
                     "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": nbytes},
        "e2e": ({"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": d2h,
                 "ms_per_step": te.item() * 1e3} if e2e_steps else None),
        "result_check": result_check,
        "gpu_launches": int(launches) * args.steps,
        "clocks": clocks,
    }
    if not args.no_cpu_baseline and world == 1:  # (the CPU baseline is timed at N = 1 only)
        from oracle import orc
        sample = min(nbytes, 1 << 30, host.numel())
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
