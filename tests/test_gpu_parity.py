"""GPU parity tests: the CUDA path, called through the C ABI (libwc.so), against the oracle and
against the golden stdout of the reference's own code.  Bit-exact: rows, order, counts, totals and
the splitter's counters.  Run on the B200 box with `pytest -m gpu`."""
import os
import random
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from conftest import ROOT, load_pkg
from oracle import orc as O
from test_oracle import ALPHABET, DELIMS, long_line_text, random_text

pytestmark = pytest.mark.gpu

SMALL = dict(short_table_log2=14, long_table_log2=14, arena_bytes=8 << 20)


@pytest.fixture(scope="module")
def pkg():
    return load_pkg()


@pytest.fixture(scope="module")
def eng_ref(pkg):
    e = pkg.Engine(mode=pkg.MODE_REF, chunk_bytes=1 << 20, **SMALL)
    yield e
    e.close()


@pytest.fixture(scope="module")
def eng_exact(pkg):
    e = pkg.Engine(mode=pkg.MODE_EXACT, chunk_bytes=1 << 20, **SMALL)
    yield e
    e.close()


def check_against_oracle(engine, data, mode=O.MODE_REF, feed=None, what=""):
    engine.reset()
    if feed is None:
        engine.feed(data)
    else:
        feed(engine, data)
    got = engine.finish()
    exp = O.fast(data, mode)
    assert got.stats["n_flush"] == exp.n_flush, what
    assert got.stats["n_empty"] == exp.n_empty, what
    assert got.stats["stop"] == exp.stop, what
    assert got.pairs() == exp.pairs(), what
    assert got.total == exp.total, what
    assert got.firsts == [f for _, _, f in exp.rows], what
    if mode == O.MODE_REF:
        assert got.stats["n_lost_empty"] == exp.n_lost_empty, what
    return got


def test_golden_vectors(eng_ref, golden):
    for name, data, expect in golden:
        if data is None:
            continue
        got = check_against_oracle(eng_ref, data, what=name)
        if O.fast(data).max_line <= 99:  # literal reference semantics == lifted semantics (A.2 L5)
            assert expect.endswith(got.rendered), name


def test_golden_vectors_exact_mode(eng_exact, golden):
    for name, data, _ in golden:
        if data is not None:
            check_against_oracle(eng_exact, data, O.MODE_EXACT, what=name)


def test_cli_stdout_is_byte_identical_to_reference(pkg, golden):
    exe = os.path.join(ROOT, "cuda-mapreduce_b200", "wordcount")
    assert os.path.exists(exe)
    picked = [g for g in golden if g[1] is None or O.fast(g[1]).max_line <= 99]
    picked = picked[:30] + picked[-10:]
    with tempfile.TemporaryDirectory() as td:
        for name, data, expect in picked:
            p = os.path.join(td, "test.txt")
            if data is None:
                if os.path.exists(p):
                    os.remove(p)
            else:
                with open(p, "wb") as f:
                    f.write(data)
            r = subprocess.run([exe], cwd=td, capture_output=True, timeout=120)
            assert r.returncode == 0, (name, r.stderr)
            assert r.stdout == expect, name


@pytest.mark.skipif(not os.path.exists(O.REF_GPU), reason="oracle/_ref/main_sm100a not built")
def test_real_reference_binary_on_this_gpu(golden):
    """L0: the reference's unmodified main.cu, compiled by nvcc for sm_100a, run on this GPU."""
    by = {n: (d, e) for n, d, e in golden}
    with tempfile.TemporaryDirectory() as td:
        for name in ("test_txt", "c2_01_prefix", "c2_02_prefix_order", "c2_03_empty_to_row0", "c2_07_crlf"):
            data, expect = by[name]
            with open(os.path.join(td, "test.txt"), "wb") as f:
                f.write(data)
            r = subprocess.run([O.REF_GPU], cwd=td, capture_output=True, timeout=120)
            assert r.stdout == expect, name


def test_cli_force_path(pkg):
    """wordcount --force-path=hash|sort: the same stdout from either aggregation path."""
    exe = os.path.join(ROOT, "cuda-mapreduce_b200", "wordcount")
    rng = random.Random(6)
    data = random_text(rng, 50000, weird=0.01)
    with tempfile.TemporaryDirectory() as td:
        with open(os.path.join(td, "test.txt"), "wb") as f:
            f.write(data)
        outs = [subprocess.run([exe] + fl, cwd=td, capture_output=True, timeout=300) for fl in ([], ["--force-path=hash"], ["--force-path=sort"])]
        assert all(o.returncode == 0 for o in outs), [o.stderr for o in outs]
        assert outs[0].stdout == outs[1].stdout == outs[2].stdout == O.fast(data, render=True).stdout


def test_cli_gpus_flag(pkg, golden):
    """wordcount --gpus=N: one process, one engine per GPU, shards cut at line starts, tables gathered on
    GPU 0 over peer memory -- stdout must be byte-identical to the single-GPU run (and to the oracle)."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    from tools.corpus import Corpus
    exe = os.path.join(ROOT, "cuda-mapreduce_b200", "wordcount")
    rng = random.Random(5)
    body = random_text(rng, 120000, weird=0.0).replace(b"\n\n", b"\n")
    k = body.index(b"\n", len(body) * 3 // 5) + 1
    cases = [("test_txt", [d for nm, d, _ in golden if nm == "test_txt"][0], True),
             ("random", body + b"x" * 40 + b" tail", True),
             ("stop_in_second_half", body[:k] + b"\n" + body[k:], False),
             ("config2_32MB", Corpus(2, total_bytes=32 << 20).bytes().tobytes(), False)]
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "test.txt")
        for name, data, echo in cases:
            with open(p, "wb") as f:
                f.write(data)
            flags = [] if echo else ["--no-echo"]
            one = subprocess.run([exe, p] + flags, cwd=td, capture_output=True, timeout=300)
            many = subprocess.run([exe, p, "--gpus=%d" % min(n, 8)] + flags, cwd=td, capture_output=True, timeout=300)
            assert one.returncode == 0 and many.returncode == 0, (name, one.stderr, many.stderr)
            assert many.stdout == one.stdout, name
            exp = O.fast(data, render=True)
            if echo:
                assert one.stdout == exp.stdout, name


def test_cli_direct_io(pkg):
    """wordcount --no-echo --direct-io: the reader thread fills the pinned slabs with O_DIRECT reads (no page
    cache copy; buffered where the file system refuses it).  Same stdout as the buffered run and the oracle's
    rows, on one GPU and -- shards starting at unaligned line starts -- on several."""
    import torch
    from tools.corpus import Corpus
    exe = os.path.join(ROOT, "cuda-mapreduce_b200", "wordcount")
    data = Corpus(2, total_bytes=(96 << 20) + 12345).bytes().tobytes()
    data = data[:data.rindex(b"\n") + 1]
    with tempfile.TemporaryDirectory(dir=ROOT) as td:  # (the repository's file system, not a tmpfs)
        p = os.path.join(td, "corpus.txt")
        with open(p, "wb") as f:
            f.write(data)
        plain = subprocess.run([exe, p, "--no-echo"], cwd=td, capture_output=True, timeout=300)
        direct = subprocess.run([exe, p, "--no-echo", "--direct-io"], cwd=td, capture_output=True, timeout=300)
        assert plain.returncode == 0 and direct.returncode == 0, (plain.stderr, direct.stderr)
        assert direct.stdout == plain.stdout
        exp = O.fast(data, render=True)
        assert plain.stdout == b"Input Data:\n" + exp.stdout[exp.stdout.index(b"-----"):]
        if torch.cuda.device_count() >= 2:
            many = subprocess.run([exe, p, "--no-echo", "--direct-io", "--gpus=2"], cwd=td, capture_output=True, timeout=300)
            assert many.returncode == 0, many.stderr
            assert many.stdout == plain.stdout


REF_LINKED = os.path.join(ROOT, "oracle", "_ref", "main_linked")


@pytest.mark.skipif(not os.path.exists(REF_LINKED), reason="oracle/_ref/main_linked not built")
def test_reference_main_linked_against_libwc(golden):
    """The drop-in boundary as a binary: the reference's own main()
    (main.cu:164-222, its splitter and printer untouched) with main.cu:37-162 cut out, linked against
    libwc.so's runMapReduce (oracle/Makefile, target _ref/main_linked).  Its stdout must be the
    unmodified reference's, byte for byte."""
    by = {n: (d, e) for n, d, e in golden}
    names = ["test_txt", "c2_01_prefix", "c2_02_prefix_order", "c2_03_empty_to_row0", "c2_07_crlf"]
    names += [n for n in by if n.startswith("c2_") and n not in names][:8]
    with tempfile.TemporaryDirectory() as td:
        for name in names:
            data, expect = by[name]
            if data is None:
                continue
            with open(os.path.join(td, "test.txt"), "wb") as f:
                f.write(data)
            r = subprocess.run([REF_LINKED], cwd=td, capture_output=True, timeout=120)
            assert r.returncode == 0, (name, r.stderr)
            assert r.stdout == expect, name


def test_fuzz_small(eng_ref, eng_exact):
    rng = random.Random(2024)
    for i in range(250):
        data = random_text(rng, rng.randint(0, 300), weird=rng.choice([0.0, 0.05, 0.3]))
        check_against_oracle(eng_ref, data, what=data)
        if i % 5 == 0:
            check_against_oracle(eng_exact, data, O.MODE_EXACT, what=data)


def test_fuzz_arbitrary_bytes(eng_ref):
    rng = np.random.default_rng(99)
    pool = np.frombuffer(b"ab \n\r\0\tz,\xc3\xa9 \n  a", dtype=np.uint8)
    for i in range(120):
        n = int(rng.integers(0, 5000))
        data = pool[rng.integers(0, len(pool), size=n)].tobytes()
        check_against_oracle(eng_ref, data, what=i)


def test_lane_serial_kernel_matches_oracle(pkg, golden):
    """WC_FLAG_LANE_SERIAL: the alternative forward kernel (count_kernel2) on the golden vectors, on fuzzed
    text with control bytes, on a tile-straddling long word and on a config-2 / config-5 sample."""
    from tools.corpus import Corpus
    rng = random.Random(77)
    with pkg.Engine(mode=pkg.MODE_REF, chunk_bytes=1 << 20, lane_serial=True, **SMALL) as e:
        for name, data, _ in golden:
            if data is not None:
                check_against_oracle(e, data, what=name)
        for i in range(120):
            data = random_text(rng, rng.randint(0, 2500), weird=rng.choice([0.0, 0.05, 0.3]))
            check_against_oracle(e, data, what=i)
        check_against_oracle(e, b"ab " + b"x" * 3000 + b" cd\n" + b"y" * 17 + b" " + b"z" * 16 + b"\n", what="long words")
    for cfg, n, kw in ((2, 24 << 20, dict(short_table_log2=18, long_table_log2=14)),
                       (5, 6_000_000, dict(short_table_log2=20, long_table_log2=21, arena_bytes=96 << 20))):
        data = Corpus(cfg, total_bytes=n).bytes()
        exp = O.fast_mt(data)
        with pkg.Engine(chunk_bytes=8 << 20, lane_serial=True, **kw) as e:
            e.feed(data)
            got = e.finish()
        assert got.pairs() == exp.pairs() and got.total == exp.total and got.stats["n_flush"] == exp.n_flush
        assert got.firsts == [f for _, _, f in exp.rows]


def long_text(rng, n_tokens, weird):
    return random_text(rng, n_tokens, weird)


def test_tile_and_chunk_boundaries(pkg):
    """Sizes around the 512-byte warp tile, the 8 KiB CTA step and a tiny streaming chunk; words and
    delimiter runs placed to straddle them."""
    rng = random.Random(77)
    with pkg.Engine(chunk_bytes=4096, **SMALL) as e:
        for n in [1, 2, 15, 16, 17, 511, 512, 513, 1023, 1024, 1025, 4095, 4096, 4097, 8191, 8192, 8193,
                  12288, 65536 + 3]:
            body = bytearray()
            while len(body) < n:
                w = rng.choice(ALPHABET + [b"x" * rng.randint(1, 40)])
                body += w + rng.choice([b" ", b" ", b"\n", b"  "])
            data = bytes(body[:n])
            check_against_oracle(e, data, what=n)
        # a delimiter exactly at every tile edge, then a word ending exactly at every tile edge
        for shift in range(0, 20):
            data = (b"a" * (511 - shift) + b" " + b"bcd ef\n" * 80)[: 512 * 3 + shift] + b"\n"
            check_against_oracle(e, data, what=("edge", shift))


def test_feed_in_ragged_calls(pkg):
    rng = random.Random(5)
    data = random_text(rng, 60000, weird=0.01)
    data = data.replace(b"\n\n", b"\n").replace(b"\0", b"")

    def ragged(engine, d):
        pos = 0
        while pos < len(d):
            k = rng.choice([1, 2, 7, 100, 3000, 5000, 70000])
            engine.feed(d[pos:pos + k])
            pos += k

    with pkg.Engine(chunk_bytes=8192, **SMALL) as e:
        for _ in range(3):
            check_against_oracle(e, data, feed=ragged)


def test_line_longer_than_chunk(pkg):
    rng = random.Random(6)
    line = b" ".join(rng.choice(ALPHABET) for _ in range(20000))  # ~100 KB without a newline
    data = b"head words here\n" + line + b"\nlast line\n" + line  # the last one is never terminated
    with pkg.Engine(chunk_bytes=4096, **SMALL) as e:
        check_against_oracle(e, data)


def test_stop_rules_across_chunks(pkg):
    rng = random.Random(8)
    base = random_text(rng, 30000, weird=0.0).replace(b"\n\n", b"\n")
    assert O.fast(base).stop == len(base)
    with pkg.Engine(chunk_bytes=4096, **SMALL) as e:
        for cut in (0, 1, 5000, len(base) // 2):
            lines = base.split(b"\n")
            k = max(1, len(lines) * cut // max(len(base), 1))
            data = b"\n".join(lines[:k]) + b"\n\n" + b"\n".join(lines[k:]) + b"\n"
            exp = O.fast(data)
            assert exp.stop < len(data)
            check_against_oracle(e, data, what=("blank", cut))
        check_against_oracle(e, b"\n" + base, what="blank first")
        check_against_oracle(e, base.rstrip(b"\n") + b"\nq", what="1-byte tail")
        check_against_oracle(e, base.rstrip(b"\n") + b"\n ", what="1-byte space tail")
        check_against_oracle(e, base + b"ab\0cd ef\n" + base[:500], what="nul mid line")
        check_against_oracle(e, base + b"a\0cd ef\n" + base[:500], what="nul at offset 1: stop")
        check_against_oracle(e, base + b"\0cd ef\n" + base[:500], what="nul at offset 0: stop")


def test_crlf_and_nul_text(pkg):
    rng = random.Random(9)
    unix = random_text(rng, 80000, weird=0.0).replace(b"\n\n", b"\n")
    dos = unix.replace(b"\n", b"\r\n")
    with pkg.Engine(chunk_bytes=1 << 16, **SMALL) as e:
        got = check_against_oracle(e, dos, what="crlf")
        assert got.stats["n_slow_tiles"] > 0
        stray = bytearray(unix)
        for _ in range(200):
            stray[rng.randrange(len(stray))] = rng.choice([0x0D, 0x00])
        check_against_oracle(e, bytes(stray), what="stray cr/nul")


def test_device_resident_input(pkg):
    import torch
    rng = random.Random(10)
    data = random_text(rng, 200000, weird=0.0).replace(b"\n\n", b"\n")
    dev = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    with pkg.Engine(**SMALL) as e:
        for _ in range(2):  # twice: reset leaves no residue
            e.reset()
            e.count_device(dev.data_ptr(), dev.numel(), last_piece=True)
            got = e.finish()
            exp = O.fast(data)
            assert got.pairs() == exp.pairs() and got.total == exp.total
            assert (got.stats["n_flush"], got.stats["n_empty"], got.stats["stop"]) == (exp.n_flush, exp.n_empty, exp.stop)
        # two pieces, cut at a line start
        cut = data.index(b"\n", len(data) // 2) + 1
        e.reset()
        e.count_device(dev.data_ptr(), cut, last_piece=False)
        tail = dev[cut:].clone()  # 16-byte aligned copy
        e.count_device(tail.data_ptr(), tail.numel(), last_piece=True)
        got = e.finish()
        assert got.pairs() == exp.pairs() and got.total == exp.total


def test_compat_runMapReduce(pkg, golden):
    """The reference's own entry point (main.cu:133) with its struct layouts, served by libwc.so."""
    lines = [[b"Hello", b"World", b"EveryOne"], [b"World", b"Good", b"News"], [b"Good", b"Morning", b"Hello"]]
    rows = pkg.run_map_reduce(lines)
    assert rows == [(b"Hello", 2), (b"World", 2), (b"EveryOne", 1), (b"Good", 2), (b"News", 1), (b"Morning", 1)]
    # prefix rule and empty words, as the reference's reducer treats them (App. C.2 #1-#3)
    assert pkg.run_map_reduce([[b"Hello", b"He", b"He"]]) == [(b"Hello", 3)]
    assert pkg.run_map_reduce([[b"He", b"Hello", b"He"]]) == [(b"He", 2), (b"Hello", 1)]
    assert pkg.run_map_reduce([[b"a", b"", b"b", b"a"]]) == [(b"a", 3), (b"b", 1)]
    assert pkg.run_map_reduce([[b"", b"", b"a", b"b"]]) == [(b"a", 1), (b"b", 1)]
    assert pkg.run_map_reduce([]) == []


def corpus_check(pkg, config, nbytes, engine_kw, straddle=(), feed_chunks=None):
    from tools.corpus import Corpus
    c = Corpus(config, total_bytes=nbytes, straddle=straddle)
    data = c.bytes()
    exp = O.fast_mt(data)
    with pkg.Engine(**engine_kw) as e:
        e.feed(data)
        got = e.finish()
    assert got.stats["n_flush"] == exp.n_flush and got.stats["n_empty"] == exp.n_empty
    assert got.stats["stop"] == exp.stop == data.size
    assert got.total == exp.total
    assert got.pairs() == exp.pairs()
    return got, exp, data


def test_config2_sample_zipf_50k(pkg):
    got, exp, data = corpus_check(pkg, 2, 48 * (1 << 20), dict(chunk_bytes=16 << 20, short_table_log2=18, long_table_log2=16))
    assert got.stats["n_slow_tiles"] == 0 and exp.max_line <= 99  # generator invariants G2/G3
    assert got.stats["n_distinct"] > 40000


def test_config3_sample_zipf_1m(pkg):
    got, exp, _ = corpus_check(pkg, 3, 64 * (1 << 20), dict(chunk_bytes=32 << 20, short_table_log2=21, long_table_log2=21))
    assert got.stats["n_distinct"] > 300000


def test_config5_sample_adversarial(pkg):
    n = 12_000_000
    got, exp, _ = corpus_check(pkg, 5, n, dict(chunk_bytes=4 << 20, short_table_log2=21, long_table_log2=22, path=pkg.PATH_HASH,
                                               arena_bytes=128 << 20), straddle=(4 << 20, 6_000_000))
    assert exp.max_line > 99 and got.stats["n_long_words"] > 500000


def test_config2_full_size_1GB(pkg):
    """BASELINE.json config 2 at full size: 10^9 bytes, V = 50k, Zipf 1.1 -- bit-exact against the
    multi-threaded oracle, plus size-independent properties."""
    import torch
    from tools.corpus import Corpus
    c = Corpus(2)
    host = torch.empty(c.total, dtype=torch.uint8).pin_memory()
    c.fill(host.numpy())
    dev = host.cuda()
    with pkg.Engine(short_table_log2=18, long_table_log2=16) as e:
        e.count_device(dev.data_ptr(), dev.numel(), last_piece=True)
        got = e.finish()
        assert got.stats["n_empty"] == 0 and got.stats["stop"] == c.total and got.stats["n_slow_tiles"] == 0
        assert got.total == got.stats["n_flush"]                      # every word lands in exactly one row
        assert got.stats["n_flush"] == int((dev == 32).sum() + (dev == 10).sum())  # one word per delimiter
        assert got.firsts == sorted(got.firsts)
        # streaming the same bytes from host memory gives the same rows
        e.reset()
        e.feed((host.data_ptr(), host.numel()))
        got2 = e.finish()
        assert got2.pairs() == got.pairs() and got2.total == got.total
    exp = O.fast_mt(host.numpy())
    assert got.pairs() == exp.pairs() and got.total == exp.total and got.stats["n_flush"] == exp.n_flush


def test_key_owner_exchange_two_engines_one_gpu(pkg):
    """The multi-GPU protocol (include/wc.h) with its real export / import kernels:
    two engines on this one GPU stand for two ranks, device-to-device copies for the NCCL all-to-all."""
    import torch
    from cuda_mapreduce_b200 import multi
    rng = random.Random(12)
    body = random_text(rng, 120000, weird=0.0).replace(b"\n\n", b"\n")
    long_words = b" ".join(bytes(rng.choice(b"abcdefgh") for _ in range(rng.randint(17, 60))) for _ in range(50))
    texts = [
        body + b" " + long_words + b"\n" + body[:3000],
        b"  \n \n" + body[:50000] + long_words + b"\n",
        body[:40000] + b"\n\n" + body[40000:],          # STOP inside the first shard
    ]
    world, REC = 2, 32
    for data in texts:
        cuts = multi.line_aligned_cuts(data, world)
        engs = [pkg.Engine(base_offset=cuts[r], chunk_bytes=1 << 20, **SMALL) for r in range(world)]
        try:
            for r, e in enumerate(engs):
                e.feed(data[cuts[r]:cuts[r + 1]])
            stats = [e.shard_stats() for e in engs]
            job = multi.reduce_stats(stats)
            for st, e in zip(stats, engs):
                if job.stop != multi.NONE and st.base >= job.stop and st.bytes > 0:
                    e.clear_tables()
            sends, counts, longs = [], [], []
            for e in engs:
                c, n_long, lb = e.export_counts(world)
                meta = torch.zeros(max(n_long * REC, 1), dtype=torch.uint8, device="cuda")
                blob = torch.zeros(max(lb, 8), dtype=torch.uint8, device="cuda")
                if n_long:
                    e.export_long(meta.data_ptr(), blob.data_ptr())
                longs.append((meta, n_long, blob))
                offs = [0, c[0]]
                buf = torch.zeros(max(sum(c) * REC, 1), dtype=torch.uint8, device="cuda")
                if sum(c):
                    e.export_words(world, offs, buf.data_ptr())
                sends.append(buf); counts.append(c)
            for dst, e in enumerate(engs):                      # "all-to-all"
                e.clear_tables()
                for src in range(world):
                    off = sum(counts[src][:dst]) * REC
                    part = sends[src][off:off + counts[src][dst] * REC].clone()
                    e.import_words(part.data_ptr(), counts[src][dst])
            root = engs[0]                                     # gather on rank 0
            c1, _, _ = engs[1].export_counts(1)
            buf = torch.zeros(max(c1[0] * REC, 1), dtype=torch.uint8, device="cuda")
            if c1[0]:
                engs[1].export_words(1, [0], buf.data_ptr())
            root.import_words(buf.data_ptr(), c1[0])
            for meta, n_long, blob in longs:
                root.import_long(meta.data_ptr(), n_long, blob.data_ptr())
            root.set_job_stats(job)
            got = root.finish()
        finally:
            for e in engs:
                e.close()
        exp = O.fast(data)
        assert got.pairs() == exp.pairs() and got.total == exp.total
        assert got.firsts == [f for _, _, f in exp.rows]
        assert (got.stats["n_flush"], got.stats["n_empty"], got.stats["n_lost_empty"], got.stats["stop"]) == \
               (exp.n_flush, exp.n_empty, exp.n_lost_empty, exp.stop)


def test_peer_gather_three_engines_one_gpu(pkg):
    """wc_peer_*: the senders' kernels store their table entries and shard counters straight into the
    root engine's landing area (on a multi-GPU node: over NVLink through a CUDA IPC mapping; here three
    engines of one process on this one GPU, attached directly).  Same engines for several jobs in a
    row: the push/pull numbering and the landing areas are reused."""
    from cuda_mapreduce_b200 import multi
    rng = random.Random(21)
    body = random_text(rng, 150000, weird=0.0).replace(b"\n\n", b"\n")
    long_words = b" ".join(bytes(rng.choice(b"abcdefgh") for _ in range(rng.randint(17, 60))) for _ in range(50))
    k2 = body.index(b"\n", len(body) * 4 // 5) + 1
    texts = [
        body + b" " + long_words + b"\n" + body[:3000],
        b"  \n \n" + body[:50000] + long_words + b"\n",   # leading empty words: lost until the first real word
        body[:40000] + b"\n\n" + body[40000:],               # STOP inside the first shard: the others are dead
        body[:k2] + b"\n" + body[k2:],                        # STOP inside the last shard
        body,
    ]
    world = 3
    engs = [pkg.Engine(chunk_bytes=1 << 20, **SMALL) for _ in range(world)]
    try:
        engs[0].peer_root_create(world, 0, 1 << SMALL["short_table_log2"], 1 << SMALL["long_table_log2"],
                                 SMALL["arena_bytes"])
        for r in (1, 2):
            engs[r].peer_attach(r, engs[0])
        for i, data in enumerate(texts + texts[:3]):
            cuts = multi.line_aligned_cuts(data, world)
            for r, e in enumerate(engs):
                e.reset(base_offset=cuts[r])
                e.feed(data[cuts[r]:cuts[r + 1]])
            for e in engs[1:]:
                e.peer_push()
            for e in engs[1:]:
                e.sync()                                   # (stands for the ranks' barrier)
            # odd jobs: the root waits for the headers and reduces the shard counters on the device
            # (wc_peer_pull_device; safe on one GPU here because the pushes have completed)
            job = engs[0].peer_pull() if i % 2 == 0 else None
            if job is None:
                engs[0].peer_pull_device()
            got = engs[0].finish()
            exp = O.fast(data)
            assert got.pairs() == exp.pairs() and got.total == exp.total
            assert got.firsts == [f for _, _, f in exp.rows]
            assert (got.stats["n_flush"], got.stats["n_empty"], got.stats["n_lost_empty"], got.stats["stop"]) == \
                (exp.n_flush, exp.n_empty, exp.n_lost_empty, exp.stop)
            assert (job.bytes if job is not None else got.stats["bytes"]) == len(data)
    finally:
        for e in engs:
            e.close()


def test_large_result_storage(pkg, monkeypatch):
    """Results of hundreds of MB leave wc_finish in plain arrays filled by several threads; the
    threshold is lowered here so that an ordinary text takes that path."""
    rng = random.Random(33)
    data = random_text(rng, 60000, weird=0.01)
    monkeypatch.setenv("WC_BIG_RESULT_BYTES", "1")
    with pkg.Engine(chunk_bytes=1 << 20, **SMALL) as e:
        check_against_oracle(e, data)
        check_against_oracle(e, data[:1000])


def check_strict(engine, data, what=""):
    """strict-ref engine against the oracle with the reference's own 99-byte chunking."""
    engine.reset()
    engine.feed(data)
    got = engine.finish()
    exp = O.fast_cap(data, O.MODE_REF, O.LINECAP_REF)
    assert (got.stats["n_flush"], got.stats["n_empty"], got.stats["stop"]) == (exp.n_flush, exp.n_empty, exp.stop), what
    assert got.pairs() == exp.pairs(), what
    assert got.total == exp.total, what
    assert got.firsts == [f for _, _, f in exp.rows], what
    assert got.stats["n_lost_empty"] == exp.n_lost_empty, what
    return got


def test_strict_ref_mode(pkg, golden):
    """WC_FLAG_STRICT_REF: lines are read in pieces of 99 bytes like main.cu's
    fgets(szLine, 100) reads them.  With it the engine reproduces the reference's stdout on EVERY golden
    vector, the long-line quirks included, and matches the chunked oracle on fuzzed long lines."""
    rng = random.Random(515)
    with pkg.Engine(strict_ref=True, chunk_bytes=1 << 20, **SMALL) as e:
        for name, data, expect in golden:
            if data is None:
                continue
            got = check_strict(e, data, what=name)
            assert expect.endswith(got.rendered), name
        for i in range(200):
            data = long_line_text(rng, rng.randint(1, 500), weird=rng.choice([0.0, 0.03, 0.15]))
            check_strict(e, data, what=("fuzz", i, data))
        # piece starts at every position relative to the 32-byte lane and the 1 KiB tile: a short first
        # line of `shift` bytes moves a long line (and its piece edges) across them
        word = b"abcdefghij"
        for shift in list(range(2, 70)) + [99, 100, 198, 924, 925, 926, 1023, 1024, 1025]:
            first = (b"x" * (shift - 1) + b"\n") if shift <= 99 else (b"y " * (shift // 2))[: shift - 1] + b"\n"
            long_line = b" ".join(word[: rng.randint(1, 10)] for _ in range(400))
            data = first + long_line + b"\n" + b"tail words here\n"
            check_strict(e, data, what=("shift", shift))
        # a line whose length is 1 mod 99 ends in a one-byte piece: the reference stops there (main.cu:185)
        for extra in (0, 1, 2, 98, 99, 100):
            data = b"ok line\n" + b"a" * 50 + b" " + b"b" * (48 + extra) + b"\nafter the long line\n"
            check_strict(e, data, what=("len", extra))
        # kills are per piece: a CR / NUL in the first piece of a long line does not hide the later pieces
        data = b"w1 w2\rhidden " + b"c" * 90 + b" seen again\nnext\n" + b"n1\0hid " + b"d" * 95 + b" back\n"
        check_strict(e, data, what="kills per piece")
    # ragged feeds and small streaming chunks
    with pkg.Engine(strict_ref=True, chunk_bytes=4096, **SMALL) as e:
        data = long_line_text(rng, 20000, weird=0.01).replace(b"\n\n", b"\n").replace(b"\0", b"")
        check_strict(e, data, what="chunked")


def test_strict_ref_cli(pkg, golden):
    """wordcount --strict-ref prints what the reference prints on the long-line golden vectors too."""
    exe = os.path.join(ROOT, "cuda-mapreduce_b200", "wordcount")
    picked = [g for g in golden if g[1] is not None and O.fast(g[1]).max_line > 99][:12]
    assert picked
    with tempfile.TemporaryDirectory() as td:
        for name, data, expect in picked:
            with open(os.path.join(td, "test.txt"), "wb") as f:
                f.write(data)
            r = subprocess.run([exe, "--strict-ref"], cwd=td, capture_output=True, timeout=120)
            assert r.returncode == 0, (name, r.stderr)
            assert r.stdout == expect, name


def test_feed_from_pageable_memory_goes_through_pinned_staging(pkg):
    """wc_feed of ordinary (pageable) host memory: pieces of 1 MiB and more are copied into a pinned
    slot by several threads and sent from there; the slots are reused across pieces and jobs."""
    rng = random.Random(44)
    base = random_text(rng, 400000, weird=0.0).replace(b"\n\n", b"\n")
    data = base * 4 + b"last words without newline"
    assert len(data) > (6 << 20)
    with pkg.Engine(chunk_bytes=2 << 20, **SMALL) as e:
        for _ in range(2):
            check_against_oracle(e, data)


def test_cli_streams_large_files_without_echo(pkg):
    """wordcount FILE --no-echo never holds the file in memory: a reader thread fills pinned slabs while
    the previous one is fed.  160 MB of config-2 text (several slabs), output against the oracle."""
    from tools.corpus import Corpus
    exe = os.path.join(ROOT, "cuda-mapreduce_b200", "wordcount")
    n = 160 << 20
    data = Corpus(2, total_bytes=n).bytes()
    exp = O.fast_mt(data)
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "big.txt")
        data.tofile(path)
        r = subprocess.run([exe, path, "--no-echo"], capture_output=True, timeout=300)
        assert r.returncode == 0, r.stderr
    lines = r.stdout.split(b"\n")
    assert lines[0] == b"Input Data:" and lines[1] == b"-" * 26 and lines[-2] == b"Total Count:%d" % exp.total
    rows = [tuple(l.rsplit(b"\t", 1)) for l in lines[2:-3]]
    assert rows == [(w, b"%d" % c) for w, c in exp.pairs()]


def test_device_corpus_generator_is_byte_identical_to_the_host_one(pkg):
    """tools/corpus_gpu.cu (bench tooling for the 16 GB / 100 GB corpora) against tools/corpus.c."""
    import torch
    from tools.corpus import Corpus
    for config, total, off, n in ((2, (9 << 20) + 12345, 0, (9 << 20) + 12345), (3, 64 << 20, 5 << 20, 6 << 20),
                                  (4, (3 << 20) + 2, 1 << 20, (2 << 20) + 2)):
        c = Corpus(config, total_bytes=total)
        dev = torch.zeros(n, dtype=torch.uint8, device="cuda")
        c.fill_device(dev.data_ptr(), off, n)
        assert np.array_equal(dev.cpu().numpy(), c.bytes(off, n)), config


def test_mapper_policy_lower_ascii(pkg):
    """wc_config.mapper: the reference's mapper emits (word, 1) for the word as it
    is; WC_MAP_LOWER_ASCII folds 'A'..'Z' inside the kernel wherever word bytes are read (tile keys,
    bytes 8..15, 17+ byte words).  A byte-wise map that leaves delimiters alone commutes with the
    splitter, so the engine on the text must equal the oracle on the mapped text -- rows, first
    offsets, prefix merges and all."""
    rng = random.Random(77)
    words = [b"The", b"the", b"THE", b"Then", b"tHEN", b"International", b"INTERNATIONAL", b"international",
             b"AVeryVeryVeryLongWord_Over_16_Bytes", b"averyveryverylongword_over_16_bytes", b"MiXeD9\xc3\x89", b"A", b"a,",
             b"Z" * 40, b"z" * 33, b"z" * 40, b"x"]
    body = bytearray()
    for _ in range(60000):
        body += rng.choice(words) + rng.choice([b" ", b" ", b"\n", b"  ", b"\r\n"])
    data = bytes(body).replace(b"\n\n", b"\n")
    with pkg.Engine(mapper=pkg.MAP_LOWER_ASCII, chunk_bytes=1 << 20, **SMALL) as e:
        for text in (data, data[:5000] + b"\n\n" + data[5000:9000], b"UPPER lower Upper\n"):
            e.reset()
            e.feed(text)
            got = e.finish()
            exp = O.fast(text.lower())  # bytes.lower() folds ASCII letters only
            assert got.pairs() == exp.pairs() and got.total == exp.total
            assert got.firsts == [f for _, _, f in exp.rows]
            assert (got.stats["n_flush"], got.stats["n_empty"], got.stats["stop"]) == (exp.n_flush, exp.n_empty, exp.stop)
    with pkg.Engine(mapper=pkg.MAP_LOWER_ASCII, mode=pkg.MODE_EXACT, path=pkg.PATH_SORT, chunk_bytes=1 << 20, **SMALL) as e:
        e.feed(data)
        got = e.finish()
        exp = O.fast(data.lower(), O.MODE_EXACT)
        assert got.pairs() == exp.pairs() and got.total == exp.total


def test_word_table_grows_instead_of_overflowing(pkg):
    """A word table far too small for the vocabulary: words it cannot take go to the overflow log, the
    host re-hashes into a larger table and replays the log -- nothing is lost (wc_engine.cu
    settle_tables)."""
    rng = random.Random(21)
    vocab = [bytes(rng.choice(b"abcdefghijklmnopqrstuvwxyz") for _ in range(rng.randint(2, 14))) for _ in range(60000)]
    words = [rng.choice(vocab) for _ in range(400000)]
    lines, cur = [], []
    for w in words:
        cur.append(w)
        if len(cur) == 12:
            lines.append(b" ".join(cur)); cur = []
    data = b"\n".join(lines) + b"\n"
    with pkg.Engine(chunk_bytes=1 << 20, short_table_log2=10, long_table_log2=10, arena_bytes=1 << 20) as e:
        got = check_against_oracle(e, data, what="grow")
        assert got.stats["table_grows"] >= 1 and got.stats["n_distinct"] > 40000
        # device-resident input in one piece: the same machinery at wc_finish
        import torch
        dev = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    with pkg.Engine(short_table_log2=12, long_table_log2=10, arena_bytes=1 << 20) as e:
        e.count_device(dev.data_ptr(), dev.numel(), last_piece=True)
        got = e.finish()
        exp = O.fast(data)
        assert got.pairs() == exp.pairs() and got.total == exp.total and got.stats["table_grows"] >= 1


def test_second_job_line_lengths(pkg):
    """WC_JOB_LINE_LENGTHS: another map/reduce pair through the same engine calls --
    every line mapped to its length, counted per length -- against its own oracle function: fuzzed text,
    lines longer than the histogram's bins, no final line feed, ragged wc_feed calls, device-resident input."""
    import torch
    from tools.corpus import Corpus
    rng = random.Random(8)
    cases = [b"", b"\n", b"a", b"ab\n\nabc\nab\nxyz", b"x" * 5000 + b"\n" + b"y" * 4096 + b"\n\n" + b"z" * 70000,
             Corpus(2, total_bytes=8 << 20).bytes().tobytes(), Corpus(5, total_bytes=3_000_000).bytes().tobytes()]
    cases += [random_text(rng, rng.randint(0, 3000), weird=rng.choice([0.0, 0.3])) for _ in range(40)]
    with pkg.Engine(job=pkg.JOB_LINE_LENGTHS, chunk_bytes=1 << 20, **SMALL) as e:
        for i, data in enumerate(cases):
            exp, n_lines = O.line_lengths(data)
            e.reset()
            if i % 3 == 0:      # ragged pieces
                pos = 0
                while pos < len(data):
                    k = rng.randint(1, 200000)
                    e.feed(data[pos:pos + k]); pos += k
            elif i % 3 == 1 and data:
                buf = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
                e.count_device(buf.data_ptr(), len(data), last_piece=True)
            else:
                e.feed(data)
            got = e.finish()
            assert [(int(w), c) for w, c in got.rows] == exp, i
            assert got.total == n_lines


def test_auto_path_switches_on_distinct_ratio(pkg, monkeypatch):
    """WC_PATH_AUTO: an engine's first job is probed (the head of its first piece is counted on the hash
    path and the counters read), later jobs follow the ratio of the job before them.  Above the threshold
    -> the sort path; Zipf text -> the hash tables; either way the oracle's rows.  (The shipped threshold
    is "never" -- the hash path won every measurement -- so the test lowers it to see the switch.)"""
    from tools.corpus import Corpus
    monkeypatch.setenv("WC_AUTO_SORT_RATIO", "0.5")
    adv = Corpus(5, total_bytes=12_000_000).bytes()
    zipf = Corpus(2, total_bytes=24 << 20).bytes()
    exp_adv, exp_zipf = O.fast_mt(adv), O.fast_mt(zipf)

    def same(got, exp):
        return got.pairs() == exp.pairs() and got.total == exp.total and got.stats["n_flush"] == exp.n_flush

    with pkg.Engine(path=pkg.PATH_AUTO, chunk_bytes=4 << 20, short_table_log2=20, long_table_log2=21,
                    arena_bytes=128 << 20) as e:
        e.feed(adv)                       # first job of the engine: probed
        got = e.finish()
        assert got.stats["path_used"] == pkg.PATH_SORT and got.stats["distinct_ratio"] > 0.5 and same(got, exp_adv)
        e.reset()
        e.feed(zipf)                      # follows the job before it (sort) ...
        got = e.finish()
        assert got.stats["path_used"] == pkg.PATH_SORT and got.stats["distinct_ratio"] < 0.1 and same(got, exp_zipf)
        e.reset()
        e.feed(zipf)                      # ... whose ratio sends this one back to the hash tables
        got = e.finish()
        assert got.stats["path_used"] == pkg.PATH_HASH and same(got, exp_zipf)
    with pkg.Engine(path=pkg.PATH_AUTO, chunk_bytes=4 << 20, short_table_log2=18, long_table_log2=14) as e:
        e.feed(zipf)                      # probed: few distinct words
        got = e.finish()
        assert got.stats["path_used"] == pkg.PATH_HASH and same(got, exp_zipf)
        import torch
        dev = torch.from_numpy(adv).cuda()
        e.reset()
        e.count_device(dev.data_ptr(), dev.numel(), last_piece=True)  # (device-resident input goes the same way)
        got = e.finish()
        assert got.stats["path_used"] == pkg.PATH_HASH and same(got, exp_adv)  # follows the Zipf job: hash, still exact


def test_warp_split_follows_the_text(pkg):
    """count_kernel's split into tokenizing and inserting warps (Launch::producers) follows the share of words
    the previous launches handed to the inserting warps: 20 + 8 by default, 18 + 10 after text that sends
    22-60 % of its words to the device tables -- learned at wc_finish, after the first pieces of a fed job, or
    (no host wait) from the counters wc_reset copies out.  Whatever the split, the rows are the oracle's."""
    import torch
    from tools.corpus import Corpus
    heavy = Corpus(3, total_bytes=64 << 20).bytes()
    light = np.frombuffer(b"the quick brown fox jumps over the lazy dog\n" * 400000, dtype=np.uint8).copy()
    exp_h, exp_l = O.fast_mt(heavy), O.fast_mt(light)
    dev_h, dev_l = torch.from_numpy(heavy).cuda(), torch.from_numpy(light).cuda()

    def load(r):
        return (r.stats["n_spilled"] + r.stats["n_long_words"]) / (r.stats["n_flush"] - r.stats["n_empty"])

    with pkg.Engine(short_table_log2=21, long_table_log2=18, path=pkg.PATH_HASH) as e:
        e.count_device(dev_h.data_ptr(), dev_h.numel())
        r1 = e.finish()
        assert r1.stats["tokenizing_warps"] == 20 and r1.pairs() == exp_h.pairs() and r1.total == exp_h.total
        assert 0.22 <= load(r1) < 0.60, load(r1)       # the sample is what this test needs it to be
        e.reset()
        e.count_device(dev_h.data_ptr(), dev_h.numel())
        r2 = e.finish()                                 # learned from the job before, at its wc_finish
        assert r2.stats["tokenizing_warps"] == 18 and r2.pairs() == exp_h.pairs() and r2.total == exp_h.total
        e.reset()
        e.count_device(dev_l.data_ptr(), dev_l.numel())
        e.sync()
        e.reset()                                       # a job dropped without wc_finish: its counters are copied out ...
        e.sync()
        e.count_device(dev_l.data_ptr(), dev_l.numel()) # ... and read here
        r3 = e.finish()
        assert load(r3) < 0.22, load(r3)
        assert r3.stats["tokenizing_warps"] == 20 and r3.pairs() == exp_l.pairs() and r3.total == exp_l.total
    # a fed job switches between its pieces (the counters are read after each of the first four): the first
    # piece is the text above, so the second one runs with what that one showed
    twice = np.concatenate([heavy, heavy])
    exp_t = O.fast_mt(twice)
    with pkg.Engine(short_table_log2=21, long_table_log2=18, chunk_bytes=heavy.size, path=pkg.PATH_HASH) as e:
        e.feed(twice)
        r4 = e.finish()
        assert r4.stats["tokenizing_warps"] == 18 and r4.pairs() == exp_t.pairs() and r4.total == exp_t.total


def test_forced_warp_split_passes_the_parity_cases():
    """The 18 + 10 split on every input of the small parity tests (WC_PRODUCERS=18 is read once per process,
    hence the child pytest)."""
    env = dict(os.environ, WC_PRODUCERS="18")
    sel = "golden_vectors or fuzz or tile_and_chunk or stop_rules or crlf or ragged or config2_sample or config5_sample or mapper_policy"
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_parity.py"), "-x", "-q", "-k", sel,
                        "-p", "no:cacheprovider"], env=env, cwd=ROOT, capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


def test_sort_path_matches_oracle(pkg, golden):
    """WC_PATH_SORT: records -> cub radix sort by hash -> runs of equal words -> (count, first).  Same
    rows as the hash path and the oracle, on the golden vectors, fuzz, STOP cases and corpus samples."""
    from tools.corpus import Corpus
    rng = random.Random(77)
    with pkg.Engine(path=pkg.PATH_SORT, chunk_bytes=1 << 20, **SMALL) as e:
        for name, data, _ in golden:
            if data is not None:
                check_against_oracle(e, data, what=("sort", name))
        for i in range(60):
            data = random_text(rng, rng.randint(0, 400), weird=rng.choice([0.0, 0.05, 0.3]))
            check_against_oracle(e, data, what=("sort fuzz", i))
        base = random_text(rng, 30000, weird=0.0).replace(b"\n\n", b"\n")
        check_against_oracle(e, base[:40000] + b"\n\n" + base[40000:], what="sort stop")
    for cfg, nbytes in ((2, 24 << 20), (5, 9_000_000)):
        data = Corpus(cfg, total_bytes=nbytes).bytes()
        exp = O.fast_mt(data)
        with pkg.Engine(path=pkg.PATH_SORT, chunk_bytes=4 << 20, short_table_log2=12, long_table_log2=14,
                        arena_bytes=64 << 20) as e:
            e.feed(data)
            got = e.finish()
        assert got.pairs() == exp.pairs() and got.total == exp.total and got.stats["n_flush"] == exp.n_flush
