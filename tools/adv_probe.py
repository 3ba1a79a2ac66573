import sys, time, json
sys.path.insert(0,'/root/repo')
import torch, numpy as np
import __graft_entry__ as ge
from tools.corpus import Corpus
pkg = ge._load_pkg()
n = int(sys.argv[1]); log2 = int(sys.argv[2])
c = Corpus(5, total_bytes=n)
host = torch.empty(n, dtype=torch.uint8).pin_memory(); c.fill(host.numpy()); dev = host.cuda()
for path in (1, 2):
    with pkg.Engine(short_table_log2=log2 if path == 1 else 12, long_table_log2=16, path=path) as e:
        for it in range(2):
            e.reset(); torch.cuda.synchronize(); t0 = time.perf_counter()
            e.count_device(dev.data_ptr(), n, True); e.aggregate(); e.sync(); dt = time.perf_counter() - t0
            ms, kms, nl = e.device_ms()
            print(json.dumps({"path": path, "bytes": n, "wall_ms": dt * 1e3, "count_kernel_ms": kms, "device_ms": ms}), flush=True)
        try:
            r = e.finish(keep_rows=False)
            print(json.dumps({"path": path, "distinct": r.stats["n_distinct"], "words": r.stats["n_flush"], "grows": r.stats["table_grows"]}), flush=True)
        except Exception as ex:
            print("finish failed:", ex, flush=True)
