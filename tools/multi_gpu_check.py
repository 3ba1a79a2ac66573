#!/usr/bin/env python3
"""Parity of the multi-GPU path on real GPUs, run under torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 tools/multi_gpu_check.py [--config 2] [--mb 256]

Every rank counts its shard of a synthetic corpus, the key-owner exchange runs over NCCL, and rank 0
compares the rows with the oracle on the whole text (bit-exact).  Prints one JSON line per case
(exchange: the NCCL direct gather, the NCCL key-owner all-to-all, the NVLink peer gather with the
device-side wait, and the peer gather with the all-reduce barrier + host wait)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--mb", type=int, default=256)
    args = ap.parse_args()
    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    from oracle import orc
    from tools.corpus import Corpus
    pkg = ge._load_pkg()
    from cuda_mapreduce_b200 import multi

    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    dev = torch.device("cuda", lr)
    cases = []
    total = args.mb << 20
    c = Corpus(args.config, total_bytes=total)
    cases.append(("config%d_%dMB" % (args.config, args.mb), c.bytes().tobytes()))
    base = cases[0][1][: 8 << 20]
    k = base.index(b"\n", len(base) * 3 // 5) + 1
    cases.append(("stop_in_a_later_shard", base[:k] + b"\n" + base[k:]))
    cases.append(("leading_empties_and_long_words", b"  \n" * 1000 + base + b" " + b"q" * 100 + b" " + b"w" * 5000 + b"\n" + base[:9999]))
    ok = True
    for name, data, mode in [(n, d, m) for (n, d) in cases for m in ("direct", "owner", "peer", "peer-host")]:
        cuts = multi.line_aligned_cuts(data, world)
        lo, hi = cuts[rank], cuts[rank + 1]
        with pkg.Engine(device=lr, base_offset=lo, short_table_log2=21, long_table_log2=12) as eng:
            eng.feed(data[lo:hi])
            ext = torch.cuda.ExternalStream(eng.stream(), device=dev)
            with torch.cuda.stream(ext):
                if mode.startswith("peer"):  # "peer": the root waits on the device; "peer-host": all-reduce + host wait
                    job, is_root = multi.PeerGather(eng, dev, 1 << 21, 1 << 12, 64 << 20, device_wait=mode == "peer").run()
                else:
                    job, is_root = multi.exchange_and_gather(eng, dev, mode=mode)
            if is_root:
                got = eng.finish()
                exp = orc.fast_mt(np.frombuffer(data, dtype=np.uint8))
                same = got.pairs() == exp.pairs() and got.total == exp.total and \
                    got.firsts == [f for _, _, f in exp.rows] and \
                    (got.stats["n_flush"], got.stats["n_empty"], got.stats["n_lost_empty"], got.stats["stop"]) == \
                    (exp.n_flush, exp.n_empty, exp.n_lost_empty, exp.stop)
                ok = ok and same
                print(json.dumps({"case": name, "exchange": mode, "world": world, "bytes": len(data), "rows": got.n_rows,
                                  "total": got.total, "bit_exact_vs_oracle": same}), flush=True)
            else:
                eng.sync()
        dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
