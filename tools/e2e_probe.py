"""Where does the host-buffer (e2e) step spend its time?  raw H2D rate, wc_feed, wc_finish."""
import sys, time
sys.path.insert(0, '/root/repo')
import torch
import __graft_entry__ as ge
from tools.corpus import Corpus
pkg = ge._load_pkg()
n = 1_000_000_000
c = Corpus(2, total_bytes=n)
host = torch.empty(n, dtype=torch.uint8).pin_memory(); c.fill(host.numpy())
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2): dev.copy_(host, non_blocking=True); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3): dev.copy_(host, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 3
print("raw H2D 1 GB pinned: %.2f ms = %.1f GB/s" % (dt * 1e3, n / dt / 1e9))
for chunk in (0, 16 << 20, 32 << 20, 128 << 20):
    with pkg.Engine(short_table_log2=18, long_table_log2=12, chunk_bytes=chunk) as e:
        for it in range(4):
            t0 = time.perf_counter(); e.reset(); t1 = time.perf_counter()
            e.feed((host.data_ptr(), n)); t2 = time.perf_counter()
            e.sync(); t3 = time.perf_counter()
            r = e.finish(keep_rows=False); t4 = time.perf_counter()
        print("chunk %4d MiB: reset %.2f feed %.2f sync %.2f finish %.2f total %.2f ms" % (chunk >> 20, (t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3, (t4-t0)*1e3))
# the same from pageable host memory (what a caller with a malloc'd buffer or a std::vector gives wc_feed)
pageable = torch.empty(n, dtype=torch.uint8)
pageable.copy_(host)
with pkg.Engine(short_table_log2=18, long_table_log2=12) as e:
    for it in range(4):
        t0 = time.perf_counter(); e.reset(); e.feed((pageable.data_ptr(), n)); e.sync(); t1 = time.perf_counter()
        r = e.finish(keep_rows=False); t2 = time.perf_counter()
    print("pageable source: feed+sync %.2f ms (%.1f GB/s) finish %.2f ms" % ((t1-t0)*1e3, n/(t1-t0)/1e9, (t2-t1)*1e3))
import ctypes as C
rt = C.CDLL("libcudart.so.12")
t0 = time.perf_counter()
for _ in range(2):
    rt.cudaMemcpy(C.c_void_p(dev.data_ptr()), C.c_void_p(pageable.data_ptr()), C.c_size_t(n), 1)
dt = (time.perf_counter() - t0) / 2
print("plain cudaMemcpy from pageable memory: %.2f ms (%.1f GB/s)" % (dt*1e3, n/dt/1e9))

