#!/usr/bin/env python3
"""Regenerates tests/golden/ref_vectors.json by executing the reference's OWN
unmodified source (/root/reference/main.cu, built as oracle/_ref/main_cpu by
oracle/Makefile) on each input, inside this container.

The GPU box has no /root/reference, so the outputs are committed as fixtures;
this script is the record of how they were made.  Every input stays inside
main.cu's fixed capacities (main.cu:9-33,170-184: <10 chunks, <=20 words per
chunk, words <=19 bytes, <10 rows), because outside them the reference has
undefined behaviour, not an answer.

    python tests/golden/make_golden.py
"""
import json
import os
import random
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402


def quirk_vectors():
    v = []
    add = lambda name, data: v.append((name, data))
    add("test_txt", open("/root/reference/test.txt", "rb").read())
    # SURVEY.md App. C.2, one per clause of App. A.1
    add("c2_01_prefix", b"Hello He He\n")
    add("c2_02_prefix_order", b"He Hello He\n")
    add("c2_03_empty_to_row0", b"a  b a\n")
    add("c2_04_leading_empties_lost", b"  a b\n")
    add("c2_05_trailing_space", b"a b \n")
    add("c2_06_tail_drop", b"a b\nc d")
    add("c2_07_crlf", b"a b\r\nc d\r\n")
    add("c2_08_tab_is_word_byte", b"a\tb c\n")
    add("c2_09_blank_line_stops", b"a b\n\nc d\n")
    add("c2_10_two_byte_line", b"a b\nc\n")
    add("c2_11_one_byte_tail", b"a b\nc")
    add("c2_12_case", b"A a A a\n")
    add("c2_13_punct", b"a, a a. a\n")
    add("c2_14_stray_cr", b"a\rb c\nd\n")
    add("c2_15_chunk99", b" ".join([b"abcdefghi"] * 12) + b"\n")
    add("c2_16_straddler", b"abcdefg " * 12 + b"0123456789 zz\n")
    add("c2_17_space_crlf", b"a \r\nb\n")
    add("c2_18_leading_crlf", b"\r\na\n")
    add("c2_19_nul", b"ab\0cd ef\ngh\n")
    add("c2_20_only_empties", b" \n \n")
    add("c2_22_missing_file", None)
    # a few more edges of the same clauses
    add("x_empty_file", b"")
    add("x_single_lf", b"\n")
    add("x_one_byte", b"a")
    add("x_one_space_tail", b"a b\n ")
    add("x_nul_at_1", b"a b\nc\0d e\nf\n")
    add("x_nul_at_0", b"a b\n\0 d e\nf\n")
    add("x_cr_only_line", b"a\n\r\nb\n")
    add("x_prefix_chain", b"abcd abc ab a abcd abc ab a\n")
    add("x_prefix_chain_rev", b"a ab abc abcd a ab abc abcd\n")
    add("x_high_bytes", b"\xc3\xa9t\xc3\xa9 \xc3\xa9 \xff\xfe x\n")
    add("x_exact_99", b"abcdefghi " * 9 + b"12345678\n")          # 99 bytes with the newline
    add("x_exact_100", b"abcdefghi " * 9 + b"123456789\n")        # newline lands in chunk 2
    add("x_no_final_newline_space", b"a b c ")
    return v


VOCAB = [b"a", b"ab", b"abc", b"abcd", b"b", b"ba", b"bab", b"c", b"ca", b"A", b"a,", b"The", b"the",
         b"then", b"x.", b"x"]


def fuzz_vectors(n, seed):
    rng = random.Random(seed)
    out = []
    for k in range(n):
        words = rng.sample(VOCAB, rng.randint(1, 8))  # <=8 distinct rows
        nlines = rng.randint(1, 8)
        buf = bytearray()
        for li in range(nlines):
            line = bytearray()
            ndelim = 0
            if rng.random() < 0.15:
                line += b" " * rng.randint(1, 2); ndelim += 2
            for wi in range(rng.randint(1, 12)):
                w = rng.choice(words)
                sep = b"  " if rng.random() < 0.12 else b" "
                if len(line) + len(w) + len(sep) > 90 or ndelim + 2 > 16:
                    break
                line += w + sep; ndelim += len(sep)
            if rng.random() < 0.6 and line.endswith(b" "):
                line = line[:-1]
            if len(line) < 1:
                line = bytearray(rng.choice(words))
            end = rng.choice([b"\n", b"\n", b"\n", b"\r\n", b" \n"])
            if li == nlines - 1 and rng.random() < 0.2:
                end = b""
            buf += line + end
        out.append(("fuzz_%03d" % k, bytes(buf)))
    return out


def main():
    orc.build()
    assert os.path.exists(orc.REF_CPU), "oracle/_ref/main_cpu missing (needs /root/reference)"
    vectors = quirk_vectors() + fuzz_vectors(120, 20261106)
    recs = []
    with tempfile.TemporaryDirectory() as td:
        for name, data in vectors:
            so = orc.run_reference_cpu(data if data is not None else b"", td, present=data is not None)
            recs.append({"name": name, "input_hex": None if data is None else data.hex(),
                         "stdout_hex": so.hex()})
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_vectors.json")
    with open(path, "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py",
                   "reference": "zimisoho/cuda-mapreduce main.cu sha256 66e9676782db..., run through oracle/shim",
                   "vectors": recs}, f, indent=0)
    print("wrote", path, len(recs), "vectors")


if __name__ == "__main__":
    main()
