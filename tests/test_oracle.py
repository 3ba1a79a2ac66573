"""CPU tests that pin the oracle (oracle/) to the reference.

Chain: reference's own main.cu on host cores (golden
fixtures, and live where oracle/_ref/main_cpu exists) == orc_literal(linecap 99)
== orc_literal(lifted) on short lines == orc_fast == orc_fast_mt, and the
byte-parallel flush formula == the line-by-line splitter.
"""
import os
import random
import tempfile

import pytest
from hypothesis import given, settings, strategies as st

from oracle import orc as O

ALPHABET = [b"a", b"b", b"ab", b"abc", b"abcd", b"A", b"a,", b"the", b"then", b"x", b"longerword9",
            b"averyveryverylongword_over_16", b"\t", b"\xc3\xa9"]
DELIMS = [b" ", b" ", b" ", b"\n", b"\n", b"\r\n", b"  ", b"\r", b"\0", b"\n\n"]


def random_text(rng, n_tokens, weird=0.15):
    out = bytearray()
    for _ in range(n_tokens):
        out += rng.choice(ALPHABET)
        if rng.random() < weird:
            out += rng.choice(DELIMS)
        else:
            out += rng.choice([b" ", b" ", b" ", b"\n"])
    if rng.random() < 0.3:
        out += rng.choice(ALPHABET)
    return bytes(out)


def test_golden_literal_matches_reference(golden):
    for name, data, expect in golden:
        if data is None:
            got = O.render_missing_file()
        else:
            got = O.literal(data, O.LINECAP_REF, render=True).stdout
        assert got == expect, name


def test_survey_c2_known_answers():
    # SURVEY.md App. C.2 rows, as (input, rows, total)
    table = [
        (b"Hello He He\n", [(b"Hello", 3)], 3),
        (b"He Hello He\n", [(b"He", 2), (b"Hello", 1)], 3),
        (b"a  b a\n", [(b"a", 3), (b"b", 1)], 4),
        (b"  a b\n", [(b"a", 1), (b"b", 1)], 2),
        (b"a b \n", [(b"a", 2), (b"b", 1)], 3),
        (b"a b\nc d", [(b"a", 1), (b"b", 1), (b"c", 1)], 3),
        (b"a b\r\nc d\r\n", [(b"a", 1), (b"b", 1), (b"c", 1), (b"d", 1)], 4),
        (b"a\tb c\n", [(b"a\tb", 1), (b"c", 1)], 2),
        (b"a b\n\nc d\n", [(b"a", 1), (b"b", 1)], 2),
        (b"a b\nc", [(b"a", 1), (b"b", 1)], 2),
        (b"a, a a. a\n", [(b"a,", 3), (b"a.", 1)], 4),
        (b"a\rb c\nd\n", [(b"a", 1), (b"d", 1)], 2),
        (b"ab\0cd ef\ngh\n", [(b"gh", 1)], 1),
        (b" \n \n", [], 0),
    ]
    for data, rows, total in table:
        for r in (O.literal(data, O.LINECAP_REF), O.literal(data, O.LINECAP_INF), O.fast(data)):
            assert r.pairs() == rows, data
            assert r.total == total, data


def test_test_txt_known_answer(golden):
    data = dict((n, d) for n, d, _ in golden)["test_txt"]
    r = O.fast(data)
    assert r.pairs() == [(b"Hello", 2), (b"World", 2), (b"EveryOne", 1), (b"Good", 2), (b"News", 1),
                         (b"Morning", 1)]
    assert r.total == 9 and r.n_flush == 9 and r.n_empty == 0 and r.stop == len(data)


def check_chain(data):
    lit = O.literal(data, O.LINECAP_INF)
    fast = O.fast(data)
    assert fast.pairs() == lit.pairs()
    assert fast.total == lit.total
    assert (fast.n_flush, fast.n_empty, fast.n_lost_empty, fast.stop) == \
           (lit.n_flush, lit.n_empty, lit.n_lost_empty, lit.stop)
    assert [f for _, _, f in fast.rows] == [f for _, _, f in lit.rows]
    if lit.max_line <= 99:  # chunking never splits a line: literal == lifted (App. A.2 L5)
        cap = O.literal(data, O.LINECAP_REF)
        assert cap.pairs() == lit.pairs() and cap.total == lit.total
    stop_f, marks_f = O.flush_formula(data)
    stop_s, marks_s = O.split_flushes(data)
    assert stop_f == stop_s == lit.stop
    assert marks_f == marks_s
    assert sum(marks_s) == lit.n_flush
    # exact mode: plain equality counts of non-empty words
    ex = O.fast(data, O.MODE_EXACT)
    assert ex.total == lit.n_flush - lit.n_empty
    return lit


def test_chain_on_golden_inputs(golden):
    for name, data, _ in golden:
        if data is not None:
            check_chain(data)


def test_chain_fuzz_seeded():
    rng = random.Random(1234)
    nontrivial = 0
    for _ in range(400):
        data = random_text(rng, rng.randint(0, 120), weird=rng.choice([0.0, 0.1, 0.4]))
        lit = check_chain(data)
        ex = O.fast(data, O.MODE_EXACT)
        nontrivial += ex.pairs() != lit.pairs()
    assert nontrivial > 100  # the prefix rule must actually bite in this corpus


@settings(max_examples=300, deadline=None)
@given(st.lists(st.sampled_from(ALPHABET + DELIMS), max_size=200))
def test_chain_hypothesis(tokens):
    check_chain(b"".join(tokens))


@settings(max_examples=200, deadline=None)
@given(st.binary(max_size=300))
def test_chain_arbitrary_bytes(data):
    check_chain(data)


def test_chain_long_words_sharing_prefixes():
    """The prefix sweep of orc_fast orders and tests words through their first 16 bytes held as two key words and
    reads the text only past them: words of 1..40 bytes over a two-letter alphabet put every case of that on the
    table (equal on 8, on 16, on more; a word and its extensions on either side of the 8- and 16-byte edges)."""
    rng = random.Random(4242)
    for trial in range(60):
        stems = [bytes(rng.choice(b"ab") for _ in range(rng.choice([7, 8, 9, 15, 16, 17, 24, 40]))) for _ in range(4)]
        words = []
        for _ in range(rng.randint(1, 60)):
            s = rng.choice(stems)
            w = s[:rng.randint(1, len(s))]
            if rng.random() < 0.3:
                w += bytes(rng.choice(b"abc\xff") for _ in range(rng.randint(1, 4)))
            words.append(w)
        lines, i = [], 0
        while i < len(words):
            k = rng.randint(1, 5)
            lines.append(b" ".join(words[i:i + k]))
            i += k
        check_chain(b"\n".join(lines) + b"\n")


def test_fast_mt_equals_fast():
    rng = random.Random(99)
    for trial in range(6):
        weird = [0.0, 0.0, 0.02, 0.0, 0.1, 0.0][trial]
        data = random_text(rng, 60000, weird=weird)
        if trial == 3:  # a blank line in the middle stops every later slice
            data = data[:len(data) // 2] + b"\n\n" + data[len(data) // 2:]
        a = O.fast(data)
        for nt in (2, 3, 8):
            b = O.fast_mt(data, nthreads=nt)
            assert b.pairs() == a.pairs() and b.total == a.total
            assert (b.n_flush, b.n_empty, b.n_lost_empty, b.stop) == (a.n_flush, a.n_empty, a.n_lost_empty, a.stop)


@pytest.mark.skipif(not os.path.exists(O.REF_CPU), reason="oracle/_ref/main_cpu not built (no /root/reference)")
def test_live_reference_fuzz():
    """orc_literal vs the reference's own code, on fresh inputs inside main.cu's capacities."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "make_golden", os.path.join(os.path.dirname(__file__), "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    with tempfile.TemporaryDirectory() as td:
        for name, data in mg.fuzz_vectors(150, 777):
            ref = O.run_reference_cpu(data, td)
            assert O.literal(data, O.LINECAP_REF, render=True).stdout == ref, (name, data)


def long_line_text(rng, n_tokens, weird=0.05):
    """Text whose lines often exceed the reference's 99-byte fgets buffer (main.cu:175,179)."""
    out = bytearray()
    for _ in range(n_tokens):
        out += rng.choice(ALPHABET + [b"w" * rng.randint(1, 120), b"zz", b"q"])
        r = rng.random()
        if r < weird:
            out += rng.choice([b"\r", b"\0", b"\r\n", b"\n\n"])
        elif r < weird + 0.04:
            out += b"\n"
        else:
            out += rng.choice([b" ", b" ", b"  "])
    return bytes(out)


def test_fast_cap_matches_literal_with_the_reference_line_buffer(golden):
    """The strict-ref oracle (99-byte chunks): orc_fast_cap == orc_literal(99) on
    every golden vector -- whose stdout it must also reproduce, long lines included -- and on fuzzed
    text with long lines, CR, NUL and blank lines."""
    for name, data, expect in golden:
        if data is None:
            continue
        a, b = O.literal(data, O.LINECAP_REF), O.fast_cap(data, O.MODE_REF, O.LINECAP_REF, render=True)
        assert a.pairs() == b.pairs() and a.total == b.total, name
        assert (a.n_flush, a.n_empty, a.n_lost_empty, a.stop) == (b.n_flush, b.n_empty, b.n_lost_empty, b.stop), name
        assert b.stdout == expect, name
    rng = random.Random(4242)
    for i in range(150):
        data = long_line_text(rng, rng.randint(1, 400), weird=rng.choice([0.0, 0.03, 0.15]))
        a, b = O.literal(data, O.LINECAP_REF), O.fast_cap(data, O.MODE_REF, O.LINECAP_REF)
        assert a.pairs() == b.pairs() and a.total == b.total, i
        assert (a.n_flush, a.n_empty, a.n_lost_empty, a.stop) == (b.n_flush, b.n_empty, b.n_lost_empty, b.stop), i



def test_line_lengths_oracle_against_python():
    """The second job's oracle (orc_line_lengths) against a two-line Python statement of it."""
    import collections
    rng = random.Random(12)
    cases = [b"", b"\n", b"a", b"a\n", b"\n\n", b"ab\n\nabc\nab\nxyz", b"x" * 5000 + b"\n" + b"y" * 4096]
    cases += [random_text(rng, rng.randint(0, 400), weird=rng.choice([0.0, 0.3])) for _ in range(200)]
    for data in cases:
        lines = data.split(b"\n")
        if lines and lines[-1] == b"":
            lines.pop()  # (what follows the last line feed is a line only if it is not empty)
        exp = sorted(collections.Counter(len(l) for l in lines).items())
        got, n_lines = O.line_lengths(data)
        assert got == exp and n_lines == len(lines), data
