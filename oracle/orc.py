"""ctypes binding of oracle/liboracle.so -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this.  See oracle/wc_oracle.h for what each call restates
(/root/reference/main.cu line references are given there).
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")
REF_CPU = os.path.join(HERE, "_ref", "main_cpu")
REF_GPU = os.path.join(HERE, "_ref", "main_sm100a")

LINECAP_REF = 99
LINECAP_INF = 0
MODE_REF = 0
MODE_EXACT = 1


class Row(C.Structure):
    _fields_ = [("off", C.c_uint64), ("len", C.c_uint32), ("count", C.c_int64), ("first", C.c_uint64)]


class Result(C.Structure):
    _fields_ = [
        ("rows", C.POINTER(Row)), ("n_rows", C.c_size_t), ("total", C.c_int64),
        ("n_flush", C.c_uint64), ("n_empty", C.c_uint64), ("n_lost_empty", C.c_uint64),
        ("stop", C.c_uint64), ("n_chunks", C.c_uint64), ("n_cr", C.c_uint64),
        ("n_nul", C.c_uint64), ("max_line", C.c_uint64),
    ]


def build(force=False):
    """Compile the restatement (and, where /root/reference exists, oracle/_ref)."""
    if force or not os.path.exists(LIB) or \
            os.path.getmtime(LIB) < os.path.getmtime(os.path.join(HERE, "wc_oracle.c")):
        subprocess.run(["make", "-C", HERE, "liboracle.so"], check=True, capture_output=True)
    subprocess.run(["make", "-C", HERE, "ref"], check=False, capture_output=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        u8p = C.c_void_p
        L.orc_literal.argtypes = [u8p, C.c_size_t, C.c_int, C.POINTER(Result)]
        L.orc_fast.argtypes = [u8p, C.c_size_t, C.c_int, C.POINTER(Result)]
        L.orc_fast_mt.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.POINTER(Result)]
        L.orc_fast_cap.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.POINTER(Result)]
        L.orc_free.argtypes = [C.POINTER(Result)]
        L.orc_flush_formula.argtypes = [u8p, C.c_size_t, u8p]
        L.orc_flush_formula.restype = C.c_uint64
        L.orc_split_flushes.argtypes = [u8p, C.c_size_t, C.c_int, u8p]
        L.orc_split_flushes.restype = C.c_uint64
        L.orc_render_stdout.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.POINTER(Result),
                                        C.c_char_p, C.c_size_t]
        L.orc_render_stdout.restype = C.c_size_t
        L.orc_result_checksum.argtypes = [u8p, C.POINTER(Result), C.POINTER(C.c_uint64)]
        L.orc_line_lengths.argtypes = [u8p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]
        L.orc_line_lengths.restype = C.c_size_t
        _lib = L
    return _lib


class Counted:
    """Python view of an orc_result: rows as [(word_bytes, count, first)], plus the counters."""

    def __init__(self, data, res, keep_rows=True):
        ck = (C.c_uint64 * 2)()
        lib().orc_result_checksum(_ptr(data)[0], C.byref(res), ck)
        self.checksum = [int(ck[0]), int(ck[1])]
        self.total = res.total
        self.n_flush = res.n_flush
        self.n_empty = res.n_empty
        self.n_lost_empty = res.n_lost_empty
        self.stop = res.stop
        self.n_chunks = res.n_chunks
        self.n_cr = res.n_cr
        self.n_nul = res.n_nul
        self.max_line = res.max_line
        self.n_rows = res.n_rows
        self.rows = []
        if keep_rows:
            mv = memoryview(data)
            for i in range(res.n_rows):
                r = res.rows[i]
                self.rows.append((bytes(mv[r.off:r.off + r.len]), r.count, r.first))

    def pairs(self):
        return [(w, c) for (w, c, _) in self.rows]


def _ptr(data):
    """Address of a bytes / bytearray / numpy uint8 array without copying."""
    if isinstance(data, bytes):
        return C.cast(C.c_char_p(data), C.c_void_p), len(data)
    if isinstance(data, bytearray):
        return C.cast((C.c_char * len(data)).from_buffer(data), C.c_void_p), len(data)
    # numpy
    return C.c_void_p(data.ctypes.data), data.size


def _run(fn, data, *args, render_linecap=None, keep_rows=True):
    p, n = _ptr(data)
    res = Result()
    fn(p, n, *args, C.byref(res))
    out = Counted(data, res, keep_rows)
    if render_linecap is not None:
        need = lib().orc_render_stdout(p, n, render_linecap, 1, C.byref(res), None, 0)
        buf = C.create_string_buffer(need)
        lib().orc_render_stdout(p, n, render_linecap, 1, C.byref(res), buf, need)
        out.stdout = buf.raw[:need]
    lib().orc_free(C.byref(res))
    return out


def literal(data, linecap=LINECAP_REF, render=False):
    return _run(lib().orc_literal, data, linecap, render_linecap=linecap if render else None)


def fast(data, mode=MODE_REF, render=False, keep_rows=True):
    return _run(lib().orc_fast, data, mode, render_linecap=LINECAP_INF if render else None,
                keep_rows=keep_rows)


def fast_cap(data, mode=MODE_REF, linecap=LINECAP_REF, render=False, keep_rows=True):
    """orc_fast with the reference's own 99-byte chunking (the engine's strict-ref mode)."""
    return _run(lib().orc_fast_cap, data, mode, linecap, render_linecap=linecap if render else None,
                keep_rows=keep_rows)


def fast_mt(data, mode=MODE_REF, nthreads=0, keep_rows=True):
    if nthreads <= 0:
        nthreads = os.cpu_count() or 1
    p, n = _ptr(data)
    res = Result()
    lib().orc_fast_mt(p, n, mode, nthreads, C.byref(res))
    out = Counted(data, res, keep_rows)
    lib().orc_free(C.byref(res))
    return out


def line_lengths(data):
    """[(length, count)] in increasing length, and the number of lines (the second job, WC_JOB_LINE_LENGTHS)."""
    p, n = _ptr(data)
    nl = C.c_uint64()
    need = lib().orc_line_lengths(p, n, None, None, 0, C.byref(nl))
    lens, cnts = (C.c_uint64 * max(need, 1))(), (C.c_uint64 * max(need, 1))()
    lib().orc_line_lengths(p, n, C.cast(lens, C.c_void_p), C.cast(cnts, C.c_void_p), need, C.byref(nl))
    return [(lens[i], cnts[i]) for i in range(need)], nl.value


def flush_formula(data):
    p, n = _ptr(data)
    buf = C.create_string_buffer(max(n, 1))
    stop = lib().orc_flush_formula(p, n, C.cast(buf, C.c_void_p))
    return stop, buf.raw[:n]


def split_flushes(data, linecap=LINECAP_INF):
    p, n = _ptr(data)
    buf = C.create_string_buffer(max(n, 1))
    stop = lib().orc_split_flushes(p, n, linecap, C.cast(buf, C.c_void_p))
    return stop, buf.raw[:n]


def render_missing_file():
    res = Result()
    need = lib().orc_render_stdout(None, 0, 0, 0, C.byref(res), None, 0)
    buf = C.create_string_buffer(need)
    lib().orc_render_stdout(None, 0, 0, 0, C.byref(res), buf, need)
    return buf.raw[:need]


def run_reference_cpu(data, workdir, present=True):
    """stdout of the reference's own unmodified main.cu (host-core shim) on `data`."""
    path = os.path.join(workdir, "test.txt")
    if present:
        with open(path, "wb") as f:
            f.write(data)
    elif os.path.exists(path):
        os.remove(path)
    return subprocess.run([REF_CPU], cwd=workdir, capture_output=True, check=True, timeout=60).stdout
