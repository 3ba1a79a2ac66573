"""ctypes binding of tools/libcorpus.so: the synthetic corpora of BASELINE.json's configs.

CONFIGS maps the config numbers of BASELINE.json / SURVEY.md 8(d) to generator
parameters (G7: vocab seed 0x5EED0000 + config, text seed 0x5EED1000 + config).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libcorpus.so")

CONFIGS = {
    2: dict(kind=0, vocab=50_000, s=1.1, bytes=1_000_000_000, name="zipf1.1_v50k_1GB"),
    3: dict(kind=0, vocab=1_000_000, s=1.1, bytes=16_000_000_000, name="zipf1.1_v1M_16GB"),
    4: dict(kind=0, vocab=1_000_000, s=1.1, bytes=100_000_000_000, name="zipf1.1_v1M_100GB"),
    5: dict(kind=1, vocab=0, s=0.0, bytes=4_000_000_000, name="adversarial_4GB"),
}

_lib = None


def build(force=False):
    src = os.path.join(HERE, "corpus.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "libcorpus.so"], check=True, capture_output=True)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        L.corpus_create.argtypes = [C.c_int, C.c_uint32, C.c_double, C.c_uint64, C.c_uint64]
        L.corpus_create.restype = C.c_void_p
        L.corpus_destroy.argtypes = [C.c_void_p]
        L.corpus_block_bytes.argtypes = [C.c_void_p]
        L.corpus_block_bytes.restype = C.c_size_t
        L.corpus_add_straddle.argtypes = [C.c_void_p, C.c_uint64]
        L.corpus_fill.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_size_t, C.c_uint64, C.c_int]
        L.corpus_word.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
        L.corpus_word.restype = C.c_void_p
        L.corpus_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        _lib = L
    return _lib


_gpu_lib = None


def gpu_lib():
    """tools/libcorpus_gpu.so (built by `make -C tools`; nvcc cross-compiles without a GPU)."""
    global _gpu_lib
    if _gpu_lib is None:
        so = os.path.join(HERE, "libcorpus_gpu.so")
        src = os.path.join(HERE, "corpus_gpu.cu")
        if (not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src)) and os.path.exists("/usr/local/cuda/bin/nvcc"):
            subprocess.run(["make", "-C", HERE, "libcorpus_gpu.so"], check=True, capture_output=True)
        L = C.CDLL(so)
        L.corpus_gpu_fill.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32,
                                      C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]
        _gpu_lib = L
    return _gpu_lib


class Corpus:
    def __init__(self, config=2, total_bytes=None, kind=None, vocab=None, s=None, straddle=()):
        p = dict(CONFIGS[config])
        if kind is not None:
            p["kind"] = kind
        if vocab is not None:
            p["vocab"] = vocab
        if s is not None:
            p["s"] = s
        self.config = config
        self.params = p
        self.total = int(total_bytes if total_bytes is not None else p["bytes"])
        self.vocab_seed = 0x5EED0000 + config
        self.text_seed = 0x5EED1000 + config
        self.h = lib().corpus_create(p["kind"], p["vocab"], p["s"], self.vocab_seed, self.text_seed)
        if not self.h:
            raise RuntimeError("corpus_create failed")
        self.block = lib().corpus_block_bytes(self.h)
        for off in straddle:
            lib().corpus_add_straddle(self.h, int(off))

    def fill(self, out, off=0, n=None, nthreads=0):
        """Write corpus bytes [off, off+n) into the uint8 buffer `out` (numpy array or address)."""
        if n is None:
            n = self.total - off
        if nthreads <= 0:
            nthreads = os.cpu_count() or 1
        addr = out.ctypes.data if isinstance(out, np.ndarray) else int(out)
        rc = lib().corpus_fill(self.h, C.c_void_p(addr), off, n, self.total, nthreads)
        if rc != 0:
            raise ValueError("corpus_fill rc=%d (range must be block aligned; block=%d)" % (rc, self.block))
        return out

    def fill_device(self, dev_ptr, off=0, n=None):
        """The same bytes written straight into device memory by tools/corpus_gpu.cu (Zipf kinds only):
        configs 3 and 4 are too large to generate and copy on the host for every run."""
        if n is None:
            n = self.total - off
        gl = gpu_lib()
        words, lens, by_len, alias, thresh = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        fill_max, text_seed, vocab = C.c_uint64(), C.c_uint64(), C.c_uint32()
        rc = lib().corpus_tables(self.h, C.byref(words), C.byref(lens), C.byref(by_len), C.byref(fill_max),
                                 C.byref(alias), C.byref(thresh), C.byref(vocab), C.byref(text_seed))
        if rc != 0:
            raise ValueError("fill_device: only the Zipf corpora have a device generator")
        rc = gl.corpus_gpu_fill(words, lens, by_len, fill_max, alias, thresh, vocab, text_seed, C.c_void_p(int(dev_ptr)),
                                off, n, self.total)
        if rc != 0:
            raise ValueError("corpus_gpu_fill rc=%d" % rc)

    def bytes(self, off=0, n=None, nthreads=0):
        if n is None:
            n = self.total - off
        out = np.empty(n, dtype=np.uint8)
        return self.fill(out, off, n, nthreads)

    def word(self, rank):
        ln = C.c_uint32()
        p = lib().corpus_word(self.h, rank, C.byref(ln))
        return C.string_at(p, ln.value)

    def close(self):
        if self.h:
            lib().corpus_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
