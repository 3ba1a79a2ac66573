"""CPU tests of the product's host side: the C ABI loads and exports what include/*.h declares, the
host half of the reduce (order-free prefix merge) matches the oracle, and the kernel's SWAR byte
classification -- compiled for the host from the same source -- is exact.  No compute call needs a GPU."""
import ctypes as C
import os
import random

import numpy as np

from conftest import load_pkg
from oracle import orc as O
from test_oracle import random_text


def test_abi_exports_every_declared_symbol():
    pkg = load_pkg()
    names = pkg.declared_symbols()
    assert len(names) >= 20 and "runMapReduce" in names
    raw = C.CDLL(pkg.lib_path())
    for n in names:
        assert hasattr(raw, n), "libwc.so does not export %s" % n


def test_strerror_and_config_default():
    pkg = load_pkg()
    L = pkg.lib()
    assert L.wc_strerror(0) == b"ok"
    assert b"full" in L.wc_strerror(-4)
    cfg = pkg.Config()
    L.wc_config_default(C.byref(cfg))
    assert cfg.struct_size == C.sizeof(pkg.Config) and cfg.mode == pkg.MODE_REF and cfg.chunk_bytes == 64 << 20


def test_classify_is_exact_for_every_byte_value():
    L = load_pkg().lib()
    d, lf, k = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rng = np.random.default_rng(7)
    for value in range(256):
        for pos in range(16):
            buf = rng.integers(0x41, 0x7B, size=16, dtype=np.uint8)  # letters: none of the classes
            buf[pos] = value
            L.wc_debug_classify16(buf.ctypes.data, C.byref(d), C.byref(lf), C.byref(k))
            assert d.value == ((1 << pos) if value in (0x20, 0x0A, 0x0D) else 0), (value, pos)
            assert lf.value == ((1 << pos) if value == 0x0A else 0), (value, pos)
            assert k.value == ((1 << pos) if value in (0x0D, 0x00) else 0), (value, pos)


def test_classify_random_bytes():
    L = load_pkg().lib()
    d, lf, k = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rng = np.random.default_rng(11)
    pool = np.array([0x20, 0x0A, 0x0D, 0x00, 0x09, 0x61, 0x80, 0xA0, 0x8A, 0x8D, 0xFF, 0x1F, 0x21], dtype=np.uint8)
    for _ in range(4000):
        buf = pool[rng.integers(0, len(pool), size=16)]
        L.wc_debug_classify16(buf.ctypes.data, C.byref(d), C.byref(lf), C.byref(k))
        exp_d = sum(1 << i for i in range(16) if buf[i] in (0x20, 0x0A, 0x0D))
        exp_l = sum(1 << i for i in range(16) if buf[i] == 0x0A)
        exp_k = sum(1 << i for i in range(16) if buf[i] in (0x0D, 0x00))
        assert (d.value, lf.value, k.value) == (exp_d, exp_l, exp_k), bytes(buf)


def test_smem_slot_hash_spreads_short_words():
    L = load_pkg().lib()
    rng = random.Random(3)
    words = set()
    while len(words) < 4000:
        words.add(bytes(rng.choice(b"abcdefghijklmnopqrstuvwxyz") for _ in range(rng.randint(1, 8))))
    slots = set(L.wc_debug_smem_slot(int.from_bytes(w, "little")) for w in words)
    m, n = L.wc_debug_smem_entries(), len(words)
    assert max(slots) < m
    ideal = m * (1 - (1 - 1 / m) ** n)  # distinct slots a uniform hash would hit
    assert len(slots) > 0.97 * ideal


def finalize_via_product(data):
    """exact counts from the oracle -> the product's host reduce."""
    pkg = load_pkg()
    ex = O.fast(data, O.MODE_EXACT)
    words = [w for w, _, _ in ex.rows]
    return pkg.wc.debug_finalize(words, [c for _, c, _ in ex.rows], [f for _, _, f in ex.rows], pkg.MODE_REF,
                                 ex.n_empty, ex.n_lost_empty)


def test_host_reduce_matches_oracle_on_golden(golden):
    for name, data, _ in golden:
        if data is None:
            continue
        ref = O.fast(data)
        got = finalize_via_product(data)
        assert got.pairs() == ref.pairs(), name
        assert got.total == ref.total, name


def test_host_reduce_matches_oracle_fuzz():
    rng = random.Random(5)
    for _ in range(300):
        data = random_text(rng, rng.randint(0, 150), weird=rng.choice([0.0, 0.1, 0.3]))
        ref = O.fast(data)
        got = finalize_via_product(data)
        assert got.pairs() == ref.pairs() and got.total == ref.total, data


def test_host_reduce_large_input_uses_all_threads():
    """Above 2^18 distinct words the host reduce (the sort path's wc_finish) runs its passes on all cores --
    bucketed sorts, per-thread histograms -- and must still give the oracle's rows: 400 k distinct words of
    1..24 bytes over a three-letter alphabet (prefix chains at every length, heads equal on 16 bytes)."""
    rng = np.random.default_rng(11)
    n_words = 900_000
    lens = rng.integers(1, 25, size=n_words)
    lens[rng.random(n_words) < 0.3] = 12               # enough equal-length words for plenty of repeats
    total = int(lens.sum()) + n_words
    buf = np.empty(total, dtype=np.uint8)
    ends = np.cumsum(lens + 1) - 1                      # where each word's delimiter sits
    buf[:] = rng.integers(97, 100, size=total, dtype=np.uint8)
    buf[ends] = 32
    buf[ends[7::8]] = 10                                # a line feed every eighth word: no blank line, no STOP
    buf[ends[-1]] = 10
    data = buf.tobytes()
    ref = O.fast_mt(data)
    ex = O.fast(data, O.MODE_EXACT)
    assert len(ex.rows) > (1 << 18) and ref.n_rows < len(ex.rows)   # large enough, and the prefix rule bites
    got = finalize_via_product(data)
    assert got.total == ref.total
    assert got.pairs() == ref.pairs()


def test_render_matches_reference_tail(golden):
    """wc_result_render reproduces main.cu:210-218 byte for byte (golden stdout minus the echo)."""
    for name, data, expect in golden:
        if data is None or O.fast(data).max_line > 99:  # beyond 99 bytes fgets chunking differs (A.2 L5)
            continue
        got = finalize_via_product(data)
        assert expect.endswith(got.rendered), name


def test_cli_file_reader_buffered_and_direct(tmp_path):
    """wordcount's reader (csrc/wordcount_main.cpp: read_at, used by --no-echo and --gpus=N, with --direct-io
    through O_DIRECT): every byte range comes back exactly, whatever its alignment, the slab size and the
    place of the end of the file."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "cli_read_selftest")
    pk = os.path.join(root, "cuda-mapreduce_b200")
    subprocess.run(["g++", "-O1", "-std=c++17", "-pthread", "-o", exe, os.path.join(root, "tests", "native", "cli_read_selftest.cpp"),
                    "-L" + pk, "-lwc", "-Wl,-rpath," + pk], check=True)
    rng = random.Random(5)
    data = bytes(rng.getrandbits(8) for _ in range(70001))
    f = tmp_path / "blob.bin"
    f.write_bytes(data)
    n = len(data)
    ranges = [(0, n), (0, 0), (1, n), (4095, 4097), (4096, 8192), (4097, n - 1), (n - 5, n), (12345, 54321), (8192, n + 100)]
    direct_ran = 0
    for mode in (0, 1):
        for cap in (8192, 12288, 1 << 20):
            for lo, hi in ranges:
                r = subprocess.run([exe, str(f), str(mode), str(cap), str(lo), str(hi)], capture_output=True)
                if mode == 1 and r.returncode == 3:
                    continue  # this file system refuses O_DIRECT: the tool falls back to buffered reads
                assert r.returncode == 0, (mode, cap, lo, hi, r.stderr)
                assert r.stdout == data[lo:hi], (mode, cap, lo, hi, len(r.stdout))
                direct_ran += mode
    print("O_DIRECT cases run:", direct_ran)
