import sys, ctypes as C
sys.path.insert(0,'/root/repo')
import torch
import __graft_entry__ as ge
from tools.corpus import Corpus
pkg = ge._load_pkg()
n = 1_000_000_000
c = Corpus(2, total_bytes=n)
host = torch.empty(n, dtype=torch.uint8).pin_memory(); c.fill(host.numpy()); dev = host.cuda()
L = pkg.lib(); L.wc_debug_counters.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
with pkg.Engine(short_table_log2=18, long_table_log2=12) as e:
    for it in range(3):
        e.reset(); e.count_device(dev.data_ptr(), n, True); e.sync()
        ms, kms, nl = e.device_ms()
        out = (C.c_uint64 * 8)(); L.wc_debug_counters(e.h, out)
        d = list(out)
        nb = max(d[3], 1)
        print("kernel_ms %.3f  batches %d  per batch cycles: wait %.0f claim %.0f global %.0f ; producer ring-full wait cycles total %.3g (per SM %.3g)" % (kms, d[3], d[0]/nb, d[1]/nb, d[2]/nb, d[4], d[4]/148))
