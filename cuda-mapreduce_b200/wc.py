"""ctypes view of libwc.so (include/wc.h).  Mirrors the C ABI one to one."""
import ctypes as C
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
MODE_REF = 0
MODE_EXACT = 1
PATH_AUTO, PATH_HASH, PATH_SORT = 0, 1, 2
FLAG_STRICT_REF = 1
FLAG_LANE_SERIAL = 2
JOB_WORDCOUNT, JOB_LINE_LENGTHS = 0, 1
MAP_IDENTITY, MAP_LOWER_ASCII = 0, 1


def lib_path():
    return os.environ.get("WC_LIB") or os.path.join(HERE, "libwc.so")


def build(force=False):
    """Compile libwc.so and the wordcount CLI in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(HERE, "csrc", f) for f in os.listdir(os.path.join(HERE, "csrc"))]
    srcs += [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    so = lib_path()
    stale = force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale:
        if not os.path.exists("/usr/local/cuda/bin/nvcc") and os.path.exists(so):
            return so
        r = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("building libwc.so failed:\n" + r.stdout + r.stderr)
    return so


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("device", C.c_int32), ("mode", C.c_int32), ("path", C.c_int32),
        ("chunk_bytes", C.c_uint64), ("short_table_log2", C.c_uint32), ("long_table_log2", C.c_uint32),
        ("arena_bytes", C.c_uint64), ("base_offset", C.c_uint64), ("flags", C.c_uint32), ("mapper", C.c_uint32),
        ("job", C.c_uint32), ("reserved", C.c_uint32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("bytes", C.c_uint64), ("n_flush", C.c_uint64), ("n_empty", C.c_uint64), ("n_lost_empty", C.c_uint64),
        ("stop", C.c_uint64), ("n_distinct", C.c_uint64), ("n_long_words", C.c_uint64), ("n_spilled", C.c_uint64),
        ("n_slow_tiles", C.c_uint64), ("table_grows", C.c_uint64), ("device_ms", C.c_double),
        ("count_kernel_ms", C.c_double), ("path_used", C.c_uint64), ("distinct_ratio", C.c_double),
        ("tokenizing_warps", C.c_uint64),
    ]


class ShardStats(C.Structure):
    _fields_ = [("base", C.c_uint64), ("bytes", C.c_uint64), ("n_flush", C.c_uint64), ("n_empty", C.c_uint64),
                ("lost_empty", C.c_uint64), ("p0", C.c_uint64), ("stop", C.c_uint64), ("p0_found", C.c_uint32),
                ("reserved", C.c_uint32)]


class WcError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("libwc: %s (%d): %s" % (lib().wc_strerror(code).decode(), code, text))
        self.code = code


_lib = None


def lib():
    """Loads libwc.so.  No fallback: if the CUDA extension is missing this raises."""
    global _lib
    if _lib is not None:
        return _lib
    so = lib_path()
    if not os.path.exists(so):
        build()
    L = C.CDLL(so)
    vp, sz, i64, u64 = C.c_void_p, C.c_size_t, C.c_int64, C.c_uint64
    L.wc_config_default.argtypes = [C.POINTER(Config)]
    L.wc_config_default.restype = None
    L.wc_create.argtypes = [C.POINTER(vp), C.POINTER(Config)]
    L.wc_destroy.argtypes = [vp]
    L.wc_destroy.restype = None
    L.wc_reset.argtypes = [vp]
    L.wc_feed.argtypes = [vp, vp, sz]
    L.wc_count_device.argtypes = [vp, vp, sz, C.c_int]
    L.wc_sync.argtypes = [vp]
    L.wc_end_input.argtypes = [vp]
    L.wc_aggregate.argtypes = [vp]
    L.wc_finish.argtypes = [vp, C.POINTER(vp)]
    L.wc_result_rows.argtypes = [vp]
    L.wc_result_rows.restype = sz
    L.wc_result_row.argtypes = [vp, sz, C.POINTER(vp), C.POINTER(sz), C.POINTER(i64)]
    L.wc_result_row_first.argtypes = [vp, sz]
    L.wc_result_row_first.restype = u64
    L.wc_result_total.argtypes = [vp]
    L.wc_result_total.restype = i64
    L.wc_result_stats.argtypes = [vp, C.POINTER(Stats)]
    L.wc_result_checksum.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.wc_result_free.argtypes = [vp]
    L.wc_result_free.restype = None
    L.wc_result_render.argtypes = [vp, C.c_char_p, sz]
    L.wc_result_render.restype = sz
    L.wc_strerror.argtypes = [C.c_int]
    L.wc_strerror.restype = C.c_char_p
    L.wc_last_error.argtypes = [vp]
    L.wc_last_error.restype = C.c_char_p
    L.wc_host_alloc.argtypes = [C.POINTER(vp), sz]
    L.wc_host_free.argtypes = [vp]
    L.wc_device_ms.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(u64)]
    L.wc_finish_d2h_bytes.argtypes = [vp]
    L.wc_finish_d2h_bytes.restype = u64
    L.wc_stream.argtypes = [vp]
    L.wc_stream.restype = vp
    L.wc_debug_finalize.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(i64), C.POINTER(u64), sz, C.c_int,
                                    u64, u64, C.POINTER(vp)]
    L.wc_debug_classify16.argtypes = [vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.wc_debug_smem_slot.argtypes = [u64]
    L.wc_debug_smem_slot.restype = C.c_uint32
    L.wc_set_base_offset.argtypes = [vp, u64]
    L.wc_shard_stats_get.argtypes = [vp, C.POINTER(ShardStats)]
    L.wc_shard_stats_reduce.argtypes = [C.POINTER(ShardStats), sz, C.POINTER(ShardStats)]
    L.wc_shard_stats_set.argtypes = [vp, C.POINTER(ShardStats)]
    L.wc_clear_tables.argtypes = [vp]
    L.wc_export_counts.argtypes = [vp, C.c_uint32, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
    L.wc_export_words.argtypes = [vp, C.c_uint32, C.POINTER(u64), vp]
    L.wc_export_long.argtypes = [vp, vp, vp]
    L.wc_import_words.argtypes = [vp, vp, u64]
    L.wc_import_long.argtypes = [vp, vp, u64, vp]
    L.wc_peer_root_create.argtypes = [vp, C.c_uint32, C.c_uint32, u64, u64, u64, vp]
    L.wc_peer_open.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, u64, u64, u64, vp]
    L.wc_peer_attach.argtypes = [vp, C.c_uint32, vp]
    L.wc_peer_push.argtypes = [vp, vp]
    L.wc_peer_pull.argtypes = [vp, C.POINTER(ShardStats)]
    L.wc_peer_pull_device.argtypes = [vp]
    L.wc_key_owner.argtypes = [u64, u64, C.c_uint32]
    L.wc_key_owner.restype = C.c_uint32
    L.wc_count_device_static.argtypes = [C.c_int, vp, sz, vp, C.c_uint32, vp, C.c_uint32, vp, u64, vp, vp]
    L.wc_static_last_error.restype = C.c_char_p
    L.runMapReduce.argtypes = [vp, C.c_int, vp, C.POINTER(C.c_int)]
    L.runMapReduce.restype = None
    _lib = L
    return L


def declared_symbols():
    """Every function name declared in include/*.h (the C ABI a maintainer binds)."""
    names = []
    inc = os.path.join(ROOT, "include")
    for f in sorted(os.listdir(inc)):
        text = open(os.path.join(inc, f)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b(wc_[a-z0-9_]+|runMapReduce)\s*\(", text)
    return sorted(set(n for n in names if n != "wc_status"))


class Result:
    """Rows [(word_bytes, count)], total and stats of one finished job (copied out of a wc_result)."""

    def __init__(self, handle, keep_rows=True):
        L = lib()
        self.total = L.wc_result_total(handle)
        self.n_rows = L.wc_result_rows(handle)
        st = Stats()
        L.wc_result_stats(handle, C.byref(st))
        self.stats = {k: getattr(st, k) for k, _ in Stats._fields_}
        ck = (C.c_uint64 * 2)()
        L.wc_result_checksum(handle, ck)
        self.checksum = [int(ck[0]), int(ck[1])]
        self.rows, self.firsts = [], []
        if keep_rows:
            need = L.wc_result_render(handle, None, 0)
            buf = C.create_string_buffer(need)
            L.wc_result_render(handle, buf, need)
            self.rendered = buf.raw[:need]
            w, ln, c = C.c_void_p(), C.c_size_t(), C.c_int64()
            for i in range(self.n_rows):
                L.wc_result_row(handle, i, C.byref(w), C.byref(ln), C.byref(c))
                self.rows.append((C.string_at(w.value, ln.value), c.value))
                self.firsts.append(L.wc_result_row_first(handle, i))
        L.wc_result_free(handle)

    def pairs(self):
        return self.rows


class Engine:
    """with Engine(mode=...) as e: e.feed(bytes); res = e.finish()"""

    def __init__(self, mode=MODE_REF, device=0, chunk_bytes=0, short_table_log2=0, long_table_log2=0,
                 arena_bytes=0, base_offset=0, path=0, strict_ref=False, mapper=0, lane_serial=False, job=0):
        L = lib()
        cfg = Config()
        L.wc_config_default(C.byref(cfg))
        cfg.mode, cfg.device, cfg.base_offset, cfg.path = mode, device, base_offset, path
        cfg.flags = (FLAG_STRICT_REF if strict_ref else 0) | (FLAG_LANE_SERIAL if lane_serial else 0)
        cfg.job = job
        cfg.mapper = mapper
        if chunk_bytes:
            cfg.chunk_bytes = chunk_bytes
        if short_table_log2:
            cfg.short_table_log2 = short_table_log2
        if long_table_log2:
            cfg.long_table_log2 = long_table_log2
        if arena_bytes:
            cfg.arena_bytes = arena_bytes
        self.h = C.c_void_p()
        rc = L.wc_create(C.byref(self.h), C.byref(cfg))
        if rc != 0:
            raise WcError(rc, L.wc_last_error(None).decode())

    def _ck(self, rc):
        if rc != 0:
            raise WcError(rc, lib().wc_last_error(self.h).decode())

    def reset(self, base_offset=None):
        self._ck(lib().wc_reset(self.h))
        if base_offset is not None:
            self._ck(lib().wc_set_base_offset(self.h, base_offset))

    def feed(self, data):
        """data: bytes / bytearray / numpy uint8 array / (address, nbytes)."""
        if isinstance(data, tuple):
            addr, n = data
        elif isinstance(data, bytes):
            addr, n = C.cast(C.c_char_p(data), C.c_void_p).value, len(data)
        elif isinstance(data, bytearray):
            addr, n = C.addressof((C.c_char * len(data)).from_buffer(data)), len(data)
        else:
            addr, n = data.ctypes.data, data.size
        self._ck(lib().wc_feed(self.h, C.c_void_p(addr), n))

    def count_device(self, dev_ptr, n, last_piece=True):
        self._ck(lib().wc_count_device(self.h, C.c_void_p(dev_ptr), n, 1 if last_piece else 0))

    def aggregate(self):
        self._ck(lib().wc_aggregate(self.h))

    def sync(self):
        self._ck(lib().wc_sync(self.h))

    def finish(self, keep_rows=True):
        r = C.c_void_p()
        self._ck(lib().wc_finish(self.h, C.byref(r)))
        return Result(r, keep_rows)

    # ---- multi-GPU protocol (include/wc.h)
    def shard_stats(self):
        st = ShardStats()
        self._ck(lib().wc_shard_stats_get(self.h, C.byref(st)))
        return st

    def set_job_stats(self, job):
        self._ck(lib().wc_shard_stats_set(self.h, C.byref(job)))

    def clear_tables(self):
        self._ck(lib().wc_clear_tables(self.h))

    def export_counts(self, n_parts):
        counts = (C.c_uint64 * n_parts)()
        nl, lb = C.c_uint64(), C.c_uint64()
        self._ck(lib().wc_export_counts(self.h, n_parts, counts, C.byref(nl), C.byref(lb)))
        return list(counts), nl.value, lb.value

    def export_words(self, n_parts, offsets, dev_ptr):
        offs = (C.c_uint64 * n_parts)(*offsets)
        self._ck(lib().wc_export_words(self.h, n_parts, offs, C.c_void_p(dev_ptr)))

    def export_long(self, dev_meta, dev_blob):
        self._ck(lib().wc_export_long(self.h, C.c_void_p(dev_meta), C.c_void_p(dev_blob)))

    def import_words(self, dev_ptr, n):
        self._ck(lib().wc_import_words(self.h, C.c_void_p(dev_ptr), n))

    def import_long(self, dev_meta, n, dev_blob):
        self._ck(lib().wc_import_long(self.h, C.c_void_p(dev_meta), n, C.c_void_p(dev_blob)))

    # ---- peer gather (include/wc.h): entries stored straight into the root GPU's memory
    def peer_root_create(self, n_ranks, root_rank, words_per_rank, long_per_rank, blob_bytes_per_rank):
        handle = (C.c_uint8 * 64)()
        self._ck(lib().wc_peer_root_create(self.h, n_ranks, root_rank, words_per_rank, long_per_rank,
                                           blob_bytes_per_rank, handle))
        return bytes(handle)

    def peer_open(self, my_rank, n_ranks, root_rank, words_per_rank, long_per_rank, blob_bytes_per_rank, handle):
        buf = (C.c_uint8 * 64).from_buffer_copy(handle)
        self._ck(lib().wc_peer_open(self.h, my_rank, n_ranks, root_rank, words_per_rank, long_per_rank,
                                    blob_bytes_per_rank, buf))

    def peer_attach(self, my_rank, root_engine):
        self._ck(lib().wc_peer_attach(self.h, my_rank, root_engine.h))

    def peer_push(self, dev_flag_ptr=None):
        self._ck(lib().wc_peer_push(self.h, C.c_void_p(dev_flag_ptr) if dev_flag_ptr else None))

    def peer_pull_device(self):
        self._ck(lib().wc_peer_pull_device(self.h))

    def peer_pull(self):
        job = ShardStats()
        self._ck(lib().wc_peer_pull(self.h, C.byref(job)))
        return job

    def finish_d2h_bytes(self):
        return lib().wc_finish_d2h_bytes(self.h)

    def device_ms(self):
        ms, kms, n = C.c_double(), C.c_double(), C.c_uint64()
        self._ck(lib().wc_device_ms(self.h, C.byref(ms), C.byref(kms), C.byref(n)))
        return ms.value, kms.value, n.value

    def stream(self):
        return lib().wc_stream(self.h)

    def close(self):
        if self.h:
            lib().wc_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def debug_finalize(words, counts, firsts, mode=MODE_REF, n_empty=0, n_lost_empty=0):
    """Host half of the reduce on a table of exact counts (no GPU).  words: list of bytes."""
    n = len(words)
    lens = (C.c_uint32 * n)(*[len(w) for w in words])
    cs = (C.c_int64 * n)(*counts)
    fs = (C.c_uint64 * n)(*firsts)
    r = C.c_void_p()
    rc = lib().wc_debug_finalize(b"".join(words), lens, cs, fs, n, mode, n_empty, n_lost_empty, C.byref(r))
    if rc != 0:
        raise WcError(rc, "wc_debug_finalize")
    return Result(r)


class _RefWord(C.Structure):
    _fields_ = [("text", C.c_char * 30)]


class _RefKeyPair(C.Structure):
    _fields_ = [("text", C.c_char * 30), ("count", C.c_int)]


class _RefInput(C.Structure):
    _fields_ = [("words", _RefWord * 20), ("n_words", C.c_int)]


def run_map_reduce(lines):
    """Calls the reference-signature entry point runMapReduce(InputData*, int, OutputData*, int*)
    (main.cu:133) with `lines` = list of lists of word bytes.  Returns [(word, count)]."""
    assert C.sizeof(_RefInput) == 604 and C.sizeof(_RefKeyPair) == 36
    n = len(lines)
    inp = (_RefInput * max(n, 1))()
    for i, ws in enumerate(lines):
        inp[i].n_words = len(ws)
        for j, w in enumerate(ws):
            inp[i].words[j].text = w
    out = (_RefKeyPair * 10)()
    cnt = C.c_int(0)
    lib().runMapReduce(C.cast(inp, C.c_void_p), n, C.cast(out, C.c_void_p), C.byref(cnt))
    return [(out[i].text, out[i].count) for i in range(cnt.value)]
