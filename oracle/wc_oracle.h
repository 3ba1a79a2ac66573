/*
 * oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A CPU restatement of the word-count path of zimisoho/cuda-mapreduce
 * (/root/reference/main.cu).  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it, and only as
 * the checker or as the timed CPU baseline.  The product (libwc.so) never
 * links or calls anything in this directory.
 *
 * Parity status: PINNED.  orc_literal() is checked against the reference's own
 * unmodified source executed on host cores (oracle/_ref/main_cpu, built by
 * oracle/Makefile from /root/reference/main.cu) on the bundled test.txt, on the
 * quirk vectors of SURVEY.md App. C.2 and under fuzzing; the outputs are
 * committed under tests/golden/.  orc_fast() is checked against orc_literal().
 */
#ifndef WC_ORACLE_H
#define WC_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_LINECAP_REF 99 /* fgets(szLine, 100, ...)  main.cu:175,179 */
#define ORC_LINECAP_INF 0  /* capacity lifted: one chunk == one whole line */

#define ORC_MODE_REF 0   /* first-match-by-prefix reduce, main.cu:57-67,81-89 */
#define ORC_MODE_EXACT 1 /* plain equality counts (no prefix merge, empties dropped) */

typedef struct orc_row {
    uint64_t off;   /* byte offset in the input of one occurrence of the word */
    uint32_t len;   /* word length in bytes */
    int64_t count;  /* what the reference would print after the tab */
    uint64_t first; /* offset of the delimiter that flushed the first occurrence */
} orc_row;

typedef struct orc_result {
    orc_row *rows; /* in output (insertion) order, main.cu:212-215 */
    size_t n_rows;
    int64_t total;         /* "Total Count", main.cu:214,218 */
    uint64_t n_flush;      /* words emitted by the splitter, empty ones included */
    uint64_t n_empty;      /* of which empty (main.cu:190-194 on a delimiter run) */
    uint64_t n_lost_empty; /* empties seen before the first non-empty word (S13) */
    uint64_t stop;         /* offset where input processing ended (== n if it ran out) */
    uint64_t n_chunks;     /* fgets chunks that were split into words */
    uint64_t n_cr;         /* 0x0D bytes seen inside processed chunks */
    uint64_t n_nul;        /* 0x00 bytes seen inside processed chunks */
    uint64_t max_line;     /* longest line (bytes, newline included) */
} orc_result;

/* Verbatim restatement: splitter main.cu:174-206, mapper 37-54, compare 57-67,
 * reducer 69-108, with the fixed capacities lifted.  O(words x rows). */
int orc_literal(const uint8_t *data, size_t n, int linecap, orc_result *out);

/* Same results through exact counts + first occurrence + the order-free
 * post-pass.  linecap is always ORC_LINECAP_INF here. */
int orc_fast(const uint8_t *data, size_t n, int mode, orc_result *out);

/* The same with the reference's own chunking: linecap = ORC_LINECAP_REF reads lines in pieces of at
 * most 99 bytes like fgets(szLine, 100) does (main.cu:175,179), so a word that straddles a piece
 * edge loses its head, CR / NUL hide only the rest of the piece, and a piece of one byte stops the
 * input (main.cu:185).  Checked against orc_literal(linecap) in tests/test_oracle.py. */
int orc_fast_cap(const uint8_t *data, size_t n, int mode, int linecap, orc_result *out);

/* orc_fast over `nthreads` line-aligned slices, merged.  Same results. */
int orc_fast_mt(const uint8_t *data, size_t n, int mode, int nthreads, orc_result *out);

void orc_free(orc_result *r);

/* A digest of all rows, for results too large to compare row by row: out[0] = sum over rows i of
 * mix(fnv1a64(word_i)) * count_i, out[1] = sum of mix(fnv1a64(word_i) + i) * (first_i + 1), modulo 2^64
 * (libwc.so's wc_result_checksum computes the same two numbers from its rows). */
void orc_result_checksum(const uint8_t *data, const orc_result *r, uint64_t out[2]);

/* The byte-parallel statement of the splitter: for every byte
 * writes 1 into flush[i] iff byte i emits a word; returns STOP.  Used to test
 * that the formula the GPU kernels implement equals the line-by-line code. */
uint64_t orc_flush_formula(const uint8_t *data, size_t n, uint8_t *flush);

/* The same marks produced by the line-by-line splitter (main.cu:174-206). */
uint64_t orc_split_flushes(const uint8_t *data, size_t n, int linecap, uint8_t *flush);

/* Renders what main.cu prints (main.cu:166,180,210-218): header, echo of every
 * chunk read, separator, rows, separator, total.  Returns bytes written (or
 * needed, if cap is too small). */
size_t orc_render_stdout(const uint8_t *data, size_t n, int linecap, int file_present,
                         const orc_result *r, char *buf, size_t cap);

/* The second job (WC_JOB_LINE_LENGTHS, include/wc.h): the lines of the input the way main.cu's reader sees
 * them before its 99-byte cap -- fgets delivers everything up to and including a line feed, and what is
 * left at the end of the file without one (main.cu:179) -- mapped to their length without the line feed.
 * Writes at most `cap` (length, count) pairs in increasing length; returns how many distinct lengths
 * there are.  *n_lines gets the number of lines. */
size_t orc_line_lengths(const uint8_t *data, size_t n, uint64_t *lengths, uint64_t *counts, size_t cap, uint64_t *n_lines);

#ifdef __cplusplus
}
#endif
#endif
