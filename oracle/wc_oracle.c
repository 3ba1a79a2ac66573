/*
 * oracle/wc_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see wc_oracle.h).
 *
 * CPU restatement of /root/reference/main.cu's word count.  Parity: PINNED
 * against the reference's own source run on host cores (oracle/_ref/main_cpu)
 * -- tests/test_oracle.py and tests/golden/.
 *
 * Each function cites the reference lines it restates.  Nothing is copied: the
 * reference works on fixed-size structs (main.cu:16-33); this file works on
 * (offset, length) views of the raw input so that none of its capacities
 * (10 chunks, 20 words per chunk, 19-byte words, 10 rows) exist here.
 */
#define _GNU_SOURCE
#include "wc_oracle.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------- */
/* Splitter: main.cu:174-206                                                 */
/* ------------------------------------------------------------------------- */

typedef void (*orc_emit_fn)(void *ctx, uint64_t off, uint32_t len, uint64_t flush_off);

typedef struct {
    size_t pos;   /* where the next fgets starts */
    size_t clen;  /* bytes fgets consumed (0: nothing left)          main.cu:179 */
    size_t slen;  /* strlen() of the zero-filled line buffer          main.cu:185,187 */
} orc_chunk;

/* One `memset(szLine,0,100); fgets(szLine,100,pFile)` step (main.cu:178-179):
 * at most linecap bytes, ending after the first '\n'.  The buffer is zeroed
 * first, so a failed read leaves "" and strlen() stops at the first NUL. */
static int next_chunk(const uint8_t *d, size_t n, int linecap, orc_chunk *c) {
    c->pos += c->clen;
    if (c->pos >= n) { c->clen = 0; c->slen = 0; return 0; }
    size_t room = n - c->pos;
    if (linecap > 0 && room > (size_t)linecap) room = (size_t)linecap;
    const uint8_t *nl = memchr(d + c->pos, '\n', room);
    c->clen = nl ? (size_t)(nl - (d + c->pos)) + 1 : room;
    const uint8_t *z = memchr(d + c->pos, 0, c->clen);
    c->slen = z ? (size_t)(z - (d + c->pos)) : c->clen;
    return 1;
}

static void split(const uint8_t *d, size_t n, uint64_t base, int linecap, orc_emit_fn emit,
                  void *ctx, orc_result *st) {
    orc_chunk c = {0, 0, 0};
    st->stop = base + n;
    for (;;) { /* while(!feof(pFile))  main.cu:176 -- an exhausted file yields "" below */
        if (!next_chunk(d, n, linecap, &c)) break;
        if (c.clen > st->max_line) st->max_line = c.clen;
        if (c.slen < c.clen) st->n_nul++;
        if (c.slen < 2) { st->stop = base + c.pos; break; } /* main.cu:185-186 */
        st->n_chunks++;
        size_t wstart = c.pos; /* nIndex = 0 per chunk, main.cu:183 */
        for (size_t i = 0; i < c.slen; i++) { /* main.cu:187 */
            uint8_t b = d[c.pos + i];
            if (b == ' ' || b == 0x0d || b == 0x0A) { /* main.cu:188 */
                size_t at = c.pos + i;
                emit(ctx, base + wstart, (uint32_t)(at - wstart), base + at); /* main.cu:190-194 */
                wstart = at + 1;
                if (b == 0x0d) { st->n_cr++; break; } /* main.cu:195-196 */
            }
        }
        /* bytes after the last delimiter are dropped: no flush after the loop,
         * main.cu:198-203 */
    }
}

/* ------------------------------------------------------------------------- */
/* Literal reducer: main.cu:37-54 (map), 57-67 (compare), 69-108 (reduce)     */
/* ------------------------------------------------------------------------- */

typedef struct {
    const uint8_t *data;
    orc_result *r;
    size_t cap;
    int seen_nonempty;
} lit_ctx;

/* compare(sz1 = row word, sz2 = new word): true iff sz2 is a prefix of sz1;
 * sz1 is NUL-terminated, so a longer sz2 fails at the terminator. main.cu:57-67 */
static int lit_is_prefix(const uint8_t *row, uint32_t row_len, const uint8_t *w, uint32_t w_len) {
    if (w_len > row_len) return 0;
    return memcmp(row, w, w_len) == 0;
}

static void lit_emit(void *vctx, uint64_t off, uint32_t len, uint64_t flush_off) {
    lit_ctx *c = (lit_ctx *)vctx;
    orc_result *r = c->r;
    r->n_flush++;
    if (len == 0) r->n_empty++;
    /* scan rows j = 0..nIndex inclusive, main.cu:81-89 */
    for (size_t j = 0; j < r->n_rows; j++) {
        if (lit_is_prefix(c->data + r->rows[j].off, r->rows[j].len, c->data + off, len)) {
            r->rows[j].count++; /* main.cu:85 */
            return;
        }
    }
    /* j == nIndex: the slot past the rows is all zero (cudaMemset, main.cu:154):
     * only the empty word matches it, and that increment is overwritten when
     * the slot is later filled (main.cu:103) -> the word is lost. */
    if (len == 0) { r->n_lost_empty++; return; }
    if (r->n_rows == c->cap) {
        c->cap = c->cap ? c->cap * 2 : 64;
        r->rows = (orc_row *)realloc(r->rows, c->cap * sizeof(orc_row));
    }
    orc_row *row = &r->rows[r->n_rows++]; /* append, main.cu:93-104 */
    row->off = off; row->len = len; row->count = 1; row->first = flush_off;
}

int orc_literal(const uint8_t *data, size_t n, int linecap, orc_result *out) {
    memset(out, 0, sizeof(*out));
    lit_ctx c = {data, out, 0, 0};
    split(data, n, 0, linecap, lit_emit, &c, out);
    for (size_t j = 0; j < out->n_rows; j++) out->total += out->rows[j].count; /* main.cu:214 */
    return 0;
}

/* ------------------------------------------------------------------------- */
/* Fast path: exact counts + first occurrence, then SURVEY.md A.3 post-pass    */
/* ------------------------------------------------------------------------- */

typedef struct {
    uint64_t k0, k1; /* the word's first 16 bytes, zero padded (a word never contains 0x00) */
    uint64_t off, first;
    int64_t count;
    uint32_t len, used;
} fent;

typedef struct {
    const uint8_t *data;
    size_t n_data;
    fent *e;
    size_t cap, n; /* cap is a power of two; a key's home slot is the TOP log2(cap) bits of its hash, so that
                    * walking one table in slot order visits another (of any size) in slot order too: merging the
                    * per-thread tables and growing are then sequential sweeps, not a pile-up of probe runs */
    int shift;
    uint64_t n_flush, n_empty, n_lost_empty, p0;
    int seen_nonempty;
} fmap;

static inline uint64_t mix_key(uint64_t k0, uint64_t k1, uint32_t len) {
    uint64_t h = (k0 ^ (len * 0xff51afd7ed558ccdull)) * 0x9FB21C651E98DF25ull;
    h ^= h >> 32;
    h = (h ^ k1) * 0x9E3779B97F4A7C15ull;
    h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
    return h;
}

static inline uint64_t hash_tail(uint64_t h, const uint8_t *p, size_t len) {
    while (len >= 8) {
        uint64_t v; memcpy(&v, p, 8);
        h = (h ^ v) * 0x9FB21C651E98DF25ull; h ^= h >> 32;
        p += 8; len -= 8;
    }
    if (len) {
        uint64_t v = 0; memcpy(&v, p, len);
        h = (h ^ v) * 0x9FB21C651E98DF25ull; h ^= h >> 32;
    }
    return h;
}

static inline void load_key(const fmap *m, uint64_t off, uint32_t len, uint64_t *k0, uint64_t *k1) {
    const uint8_t *p = m->data + off;
    uint64_t a = 0, b = 0;
    if (off + 16 <= m->n_data) {
        memcpy(&a, p, 8); memcpy(&b, p + 8, 8);
        if (len < 8) { a &= (1ull << (8 * len)) - 1; b = 0; }
        else if (len < 16) b &= (1ull << (8 * (len - 8))) - 1; /* len == 8 -> mask 0 */
    } else {
        memcpy(&a, p, len < 8 ? len : 8);
        if (len > 8) memcpy(&b, p + 8, (len < 16 ? len : 16) - 8);
    }
    *k0 = a; *k1 = b;
}

static void fmap_init(fmap *m, const uint8_t *data, size_t n_data, size_t cap) {
    memset(m, 0, sizeof(*m));
    m->data = data; m->n_data = n_data; m->cap = cap;
    m->shift = 64 - __builtin_ctzll((unsigned long long)cap);
    m->e = (fent *)calloc(cap, sizeof(fent));
    m->p0 = UINT64_MAX;
}

static void fmap_add(fmap *m, uint64_t k0, uint64_t k1, uint64_t off, uint32_t len, int64_t count, uint64_t first);

static void fmap_grow(fmap *m) {
    fent *old = m->e; size_t ocap = m->cap;
    m->cap = ocap * 2; m->n = 0; m->shift--;
    m->e = (fent *)calloc(m->cap, sizeof(fent));
    for (size_t i = 0; i < ocap; i++)
        if (old[i].used) fmap_add(m, old[i].k0, old[i].k1, old[i].off, old[i].len, old[i].count, old[i].first);
    free(old);
}

static void fmap_add(fmap *m, uint64_t k0, uint64_t k1, uint64_t off, uint32_t len, int64_t count, uint64_t first) {
    uint64_t hash = mix_key(k0, k1, len);
    if (len > 16) hash = hash_tail(hash, m->data + off + 16, len - 16);
    size_t mask = m->cap - 1, i = (size_t)(hash >> m->shift);
    for (;;) {
        fent *e = &m->e[i];
        if (!e->used) {
            e->used = 1; e->k0 = k0; e->k1 = k1; e->off = off; e->len = len; e->count = count; e->first = first;
            if (++m->n * 10 > m->cap * 7) fmap_grow(m);
            return;
        }
        if (e->k0 == k0 && e->k1 == k1 && e->len == len &&
            (len <= 16 || memcmp(m->data + e->off + 16, m->data + off + 16, len - 16) == 0)) {
            e->count += count;
            if (first < e->first) e->first = first;
            return;
        }
        i = (i + 1) & mask;
    }
}

static void fast_emit(void *vctx, uint64_t off, uint32_t len, uint64_t flush_off) {
    fmap *m = (fmap *)vctx;
    m->n_flush++;
    if (len == 0) {
        m->n_empty++;
        if (!m->seen_nonempty) m->n_lost_empty++;
        return;
    }
    if (!m->seen_nonempty) { m->seen_nonempty = 1; m->p0 = flush_off; }
    uint64_t k0, k1;
    load_key(m, off, len, &k0, &k1);
    fmap_add(m, k0, k1, off, len, 1, flush_off);
}

/* Bytewise order of two words (shorter first among a word and its extensions).  The first 16 bytes are
 * compared through the zero-padded key words the table already holds -- a word never contains 0x00, so
 * the padding sorts a word before every extension of it -- and the text is only touched for two words
 * that agree on 16 bytes (at 3e8 distinct words the sort is otherwise one cache miss per comparison). */
static int cmp_lex(const void *a, const void *b, void *vdata) {
    const uint8_t *d = (const uint8_t *)vdata;
    const fent *x = (const fent *)a, *y = (const fent *)b;
    uint64_t p = __builtin_bswap64(x->k0), q = __builtin_bswap64(y->k0);
    if (p != q) return p < q ? -1 : 1;
    p = __builtin_bswap64(x->k1); q = __builtin_bswap64(y->k1);
    if (p != q) return p < q ? -1 : 1;
    if (x->len > 16 && y->len > 16) {
        uint32_t l = (x->len < y->len ? x->len : y->len) - 16;
        int c = memcmp(d + x->off + 16, d + y->off + 16, l);
        if (c) return c;
    }
    return (x->len > y->len) - (x->len < y->len);
}

/* is word t a prefix of (or equal to) word w? */
static inline int is_prefix_of(const uint8_t *d, const fent *t, const fent *w) {
    if (t->len > w->len) return 0;
    if (t->len <= 8) return (w->k0 & (t->len == 8 ? ~0ull : (1ull << (8 * t->len)) - 1)) == t->k0;
    if (w->k0 != t->k0) return 0;
    if (t->len <= 16) return (w->k1 & (t->len == 16 ? ~0ull : (1ull << (8 * (t->len - 8))) - 1)) == t->k1;
    return w->k1 == t->k1 && memcmp(d + t->off + 16, d + w->off + 16, t->len - 16) == 0;
}

static int cmp_first(const void *a, const void *b) {
    const orc_row *x = (const orc_row *)a, *y = (const orc_row *)b;
    return (x->first > y->first) - (x->first < y->first);
}

/* ---- sorting the large jobs on all threads ------------------------------------------------------------------
 * A bucket function that is monotone in the sort order cuts the array into ranges that can be sorted
 * independently: count, scatter into a second array, qsort_r every bucket on its own thread.  Returns the
 * sorted array (the second one; the first is freed), or the input sorted in place when it is small. */
typedef uint32_t (*orc_bucket_fn)(const void *elem);
typedef int (*orc_cmp_fn)(const void *, const void *, void *);
typedef struct {
    uint8_t *base; size_t size; const size_t *start; uint32_t nb; volatile uint32_t *next;
    orc_cmp_fn cmp; void *arg;
} psort_job;

static void *psort_main(void *vp) {
    psort_job *j = (psort_job *)vp;
    for (;;) {
        uint32_t b = __sync_fetch_and_add(j->next, 1u);
        if (b >= j->nb) return NULL;
        size_t n = j->start[b + 1] - j->start[b];
        if (n > 1) qsort_r(j->base + j->start[b] * j->size, n, j->size, j->cmp, j->arg);
    }
}

static void *parallel_sort(void *base, size_t n, size_t size, uint32_t nb, orc_bucket_fn bucket,
                           orc_cmp_fn cmp, void *arg, int nthreads) {
    if (nthreads <= 1 || n < (1u << 18)) { qsort_r(base, n, size, cmp, arg); return base; }
    uint8_t *src = (uint8_t *)base, *dst = (uint8_t *)malloc(n * size);
    size_t *start = (size_t *)calloc((size_t)nb + 2, sizeof(size_t));
    if (!dst || !start) { free(dst); free(start); qsort_r(base, n, size, cmp, arg); return base; }
    for (size_t i = 0; i < n; i++) start[bucket(src + i * size) + 2]++;
    for (uint32_t b = 0; b < nb; b++) start[b + 2] += start[b + 1]; /* start[b + 1] = where bucket b begins */
    for (size_t i = 0; i < n; i++) memcpy(dst + start[bucket(src + i * size) + 1]++ * size, src + i * size, size);
    free(src);                                                      /* now start[b + 1] = where bucket b ends */
    volatile uint32_t next = 0;
    psort_job job = {dst, size, start, nb, &next, cmp, arg};
    pthread_t th[64];
    if (nthreads > 64) nthreads = 64;
    int started = 0;
    for (int t = 0; t < nthreads; t++) started += pthread_create(&th[started], NULL, psort_main, &job) == 0;
    if (!started) psort_main(&job);
    for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
    free(start);
    return dst;
}

static uint32_t bucket_lex(const void *e) { /* the word's first two bytes */
    return (uint32_t)(__builtin_bswap64(((const fent *)e)->k0) >> 48);
}
static int g_first_shift; /* set by finalize() before the rows are sorted: first >> shift < 65536 */
static uint32_t bucket_first(const void *e) { return (uint32_t)(((const orc_row *)e)->first >> g_first_shift); }
static int cmp_first_r(const void *a, const void *b, void *unused) { (void)unused; return cmp_first(a, b); }

/* ORC_TIMING=1: phase times of the large jobs on stderr */
static void orc_tick(const char *what) {
    static struct timespec last;
    struct timespec now;
    if (!getenv("ORC_TIMING")) return;
    clock_gettime(CLOCK_MONOTONIC, &now);
    if (what) fprintf(stderr, "[orc] %-18s %8.2f s\n", what, (double)(now.tv_sec - last.tv_sec) + 1e-9 * (double)(now.tv_nsec - last.tv_nsec));
    last = now;
}

/* Turn a table of exact counts into the rows main.cu would print. */
static void finalize(fmap *m, int mode, int nthreads, orc_result *out) {
    size_t nd = 0;
    fent *v = (fent *)malloc((m->n ? m->n : 1) * sizeof(fent));
    for (size_t i = 0; i < m->cap; i++)
        if (m->e[i].used) v[nd++] = m->e[i];
    free(m->e); m->e = NULL; /* the large jobs need the room for the sort */
    out->n_flush = m->n_flush; out->n_empty = m->n_empty; out->n_lost_empty = m->n_lost_empty;
    out->rows = NULL; out->n_rows = 0; out->total = 0;
    if (nd == 0) { free(v); return; }

    if (mode == ORC_MODE_EXACT) {
        out->rows = (orc_row *)malloc(nd * sizeof(orc_row));
        for (size_t i = 0; i < nd; i++) {
            orc_row r = {v[i].off, v[i].len, v[i].count, v[i].first};
            out->rows[i] = r; out->total += v[i].count;
        }
        out->n_rows = nd;
        qsort(out->rows, nd, sizeof(orc_row), cmp_first);
        free(v);
        return;
    }

    /* A.3: owner(w) = the earliest-first-seen distinct word that has w as a
     * prefix.  In bytewise order the words having prefix w are a contiguous
     * run that starts at w, so one sweep with a stack of open prefixes gives
     * every run's minimum. */
    orc_tick(NULL);
    v = (fent *)parallel_sort(v, nd, sizeof(fent), 65536, bucket_lex, cmp_lex, (void *)m->data, nthreads);
    orc_tick("sort by bytes");
    size_t *owner = (size_t *)malloc(nd * sizeof(size_t));
    size_t *stk = (size_t *)malloc(nd * sizeof(size_t));
    size_t *best = (size_t *)malloc(nd * sizeof(size_t)); /* argmin(first) of the open run */
    size_t sp = 0;
    for (size_t i = 0; i <= nd; i++) {
        while (sp > 0) {
            size_t t = stk[sp - 1];
            int open = i < nd && is_prefix_of(m->data, &v[t], &v[i]);
            if (open) break;
            owner[t] = best[sp - 1];
            if (sp > 1 && v[best[sp - 1]].first < v[best[sp - 2]].first) best[sp - 2] = best[sp - 1];
            sp--;
        }
        if (i < nd) { stk[sp] = i; best[sp] = i; sp++; }
    }
    orc_tick("prefix sweep");
    int64_t *acc = (int64_t *)calloc(nd, sizeof(int64_t));
    size_t nrows = 0;
    for (size_t i = 0; i < nd; i++) { acc[owner[i]] += v[i].count; nrows += owner[i] == i; }
    out->rows = (orc_row *)malloc(nrows * sizeof(orc_row));
    for (size_t i = 0; i < nd; i++)
        if (owner[i] == i) {
            orc_row r = {v[i].off, v[i].len, acc[i], v[i].first};
            out->rows[out->n_rows++] = r;
        }
    for (g_first_shift = 0; (m->n_data >> g_first_shift) >= 65536; g_first_shift++) {}
    out->rows = (orc_row *)parallel_sort(out->rows, out->n_rows, sizeof(orc_row), 65536, bucket_first, cmp_first_r, NULL, nthreads);
    orc_tick("rows by first");
    /* S13: empty words after the first non-empty one increment row 0 */
    out->rows[0].count += (int64_t)(m->n_empty - m->n_lost_empty);
    for (size_t i = 0; i < out->n_rows; i++) out->total += out->rows[i].count;
    free(acc); free(best); free(stk); free(owner); free(v);
}

int orc_fast_cap(const uint8_t *data, size_t n, int mode, int linecap, orc_result *out) {
    memset(out, 0, sizeof(*out));
    fmap m; fmap_init(&m, data, n, 1u << 16);
    split(data, n, 0, linecap, fast_emit, &m, out);
    finalize(&m, mode, 1, out);
    free(m.e);
    return 0;
}

int orc_fast(const uint8_t *data, size_t n, int mode, orc_result *out) {
    return orc_fast_cap(data, n, mode, ORC_LINECAP_INF, out);
}

/* ---- multi-threaded variant: line-aligned slices, per-thread tables ------- */

typedef struct {
    const uint8_t *data; size_t n, lo, hi;
    fmap m; orc_result st;
} slice_job;

static void *slice_main(void *arg) {
    slice_job *j = (slice_job *)arg;
    /* the table is keyed on absolute offsets into data */
    fmap_init(&j->m, j->data, j->n, 1u << 16);
    memset(&j->st, 0, sizeof(j->st));
    /* split() reports offsets relative to its own buffer start plus `base`;
     * passing the slice pointer with base = lo keeps them absolute. */
    split(j->data + j->lo, j->hi - j->lo, j->lo, ORC_LINECAP_INF, fast_emit, &j->m, &j->st);
    return NULL;
}

int orc_fast_mt(const uint8_t *data, size_t n, int mode, int nthreads, orc_result *out) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads == 1 || n < (1u << 16)) return orc_fast(data, n, mode, out);
    memset(out, 0, sizeof(*out));
    slice_job *jobs = (slice_job *)calloc((size_t)nthreads, sizeof(slice_job));
    pthread_t *th = (pthread_t *)calloc((size_t)nthreads, sizeof(pthread_t));
    size_t prev = 0;
    int nj = 0;
    for (int t = 0; t < nthreads && prev < n; t++) {
        size_t want = (t + 1 == nthreads) ? n : (size_t)((double)n * (t + 1) / nthreads);
        if (want < prev) want = prev;
        size_t hi = n;
        if (want < n) { /* move the cut to the next line start: no state crosses it */
            const uint8_t *nl = memchr(data + want, '\n', n - want);
            hi = nl ? (size_t)(nl - data) + 1 : n;
        }
        if (hi == prev) continue;
        jobs[nj].data = data; jobs[nj].n = n; jobs[nj].lo = prev; jobs[nj].hi = hi;
        prev = hi; nj++;
    }
    orc_tick(NULL);
    for (int t = 0; t < nj; t++) pthread_create(&th[t], NULL, slice_main, &jobs[t]);
    for (int t = 0; t < nj; t++) pthread_join(th[t], NULL);
    orc_tick("slices");

    /* The merged table is sized for every slice's entries before the first goes in: the slices are walked in
     * slot order = hash order, and a table that was still small would take such a run as one long pile of
     * collisions at its start (quadratic); at its final size the run lands slot after slot. */
    size_t want_cap = 1u << 16, n_all = 0;
    for (int t = 0; t < nj; t++) n_all += jobs[t].m.n;
    while (want_cap * 7 < (n_all + 1) * 10) want_cap <<= 1;
    fmap m; fmap_init(&m, data, n, want_cap);
    out->stop = n;
    for (int t = 0; t < nj; t++) {
        slice_job *j = &jobs[t];
        if (!m.seen_nonempty) {
            /* empties before the job-wide first non-empty word are lost (S13) */
            m.n_lost_empty += j->m.seen_nonempty ? j->m.n_lost_empty : j->m.n_empty;
            if (j->m.seen_nonempty) { m.seen_nonempty = 1; m.p0 = j->m.p0; }
        }
        m.n_flush += j->m.n_flush; m.n_empty += j->m.n_empty;
        for (size_t i = 0; i < j->m.cap; i++)
            if (j->m.e[i].used)
                fmap_add(&m, j->m.e[i].k0, j->m.e[i].k1, j->m.e[i].off, j->m.e[i].len, j->m.e[i].count, j->m.e[i].first);
        out->n_chunks += j->st.n_chunks; out->n_cr += j->st.n_cr; out->n_nul += j->st.n_nul;
        if (j->st.max_line > out->max_line) out->max_line = j->st.max_line;
        if (j->st.stop < j->hi) { out->stop = j->st.stop; break; } /* main.cu:185: later slices never ran */
    }
    for (int t = 0; t < nj; t++) free(jobs[t].m.e);
    orc_tick("merge of slices");
    finalize(&m, mode, nthreads, out);
    free(m.e);
    free(jobs); free(th);
    return 0;
}

static void mark_emit(void *vctx, uint64_t off, uint32_t len, uint64_t flush_off) {
    (void)off; (void)len;
    ((uint8_t *)vctx)[flush_off] = 1;
}

uint64_t orc_split_flushes(const uint8_t *data, size_t n, int linecap, uint8_t *flush) {
    orc_result st; memset(&st, 0, sizeof(st));
    memset(flush, 0, n);
    split(data, n, 0, linecap, mark_emit, flush, &st);
    return st.stop;
}

void orc_free(orc_result *r) {
    free(r->rows);
    r->rows = NULL; r->n_rows = 0;
}

/* ------------------------------------------------------------------------- */
/* Byte-parallel statement of the splitter                     */
/* ------------------------------------------------------------------------- */

uint64_t orc_flush_formula(const uint8_t *b, size_t n, uint8_t *flush) {
    /* STOP: first line start i at which the line's strlen() is < 2:
     *   b[i]=='\n' (blank line), b[i]==0, b[i+1]==0, or a 1-byte unterminated tail */
    uint64_t stop = n;
    int64_t last_lf = -1, last_kill = -1;
    memset(flush, 0, n);
    for (size_t i = 0; i < n; i++) {
        int line_start = (i == 0) || b[i - 1] == '\n';
        if (line_start) {
            int short_line = b[i] == '\n' || b[i] == 0 || (i + 1 < n && b[i + 1] == 0) || (i + 1 == n);
            if (short_line) { stop = i; break; }
        }
        int delim = b[i] == ' ' || b[i] == 0x0d || b[i] == 0x0A;
        int live = last_kill <= last_lf; /* no CR / NUL earlier on this line */
        if (delim && live) flush[i] = 1;
        if (b[i] == 0x0d || b[i] == 0) last_kill = (int64_t)i;
        if (b[i] == '\n') last_lf = (int64_t)i;
    }
    return stop;
}

/* ------------------------------------------------------------------------- */
/* What main.cu prints: main.cu:166,180,210-218                                */
/* ------------------------------------------------------------------------- */

typedef struct { char *buf; size_t cap, len; } sbuf;
static void sb_put(sbuf *s, const void *p, size_t n) {
    if (s->len + n <= s->cap) memcpy(s->buf + s->len, p, n);
    s->len += n;
}

size_t orc_render_stdout(const uint8_t *data, size_t n, int linecap, int file_present,
                         const orc_result *r, char *buf, size_t cap) {
    sbuf s = {buf, cap, 0};
    static const char sep[] = "--------------------------\n";
    sb_put(&s, "Input Data:\n", 12); /* main.cu:166 */
    if (file_present) {
        orc_chunk c = {0, 0, 0};
        for (;;) {
            int got = next_chunk(data, n, linecap, &c);
            if (got) sb_put(&s, data + c.pos, c.slen); /* printf("%s", szLine), main.cu:180 */
            if (!got || c.slen < 2) break;
        }
    }
    sb_put(&s, sep, sizeof(sep) - 1); /* main.cu:210 */
    char num[48];
    for (size_t i = 0; i < r->n_rows; i++) { /* main.cu:212-215 */
        sb_put(&s, data + r->rows[i].off, r->rows[i].len);
        int k = snprintf(num, sizeof(num), "\t%lld\n", (long long)r->rows[i].count);
        sb_put(&s, num, (size_t)k);
    }
    sb_put(&s, sep, sizeof(sep) - 1); /* main.cu:217 */
    int k = snprintf(num, sizeof(num), "Total Count:%lld\n", (long long)r->total);
    sb_put(&s, num, (size_t)k); /* main.cu:218 */
    return s.len;
}

/* ---- second job: line lengths (main.cu:179 is the line reader; the cap of 99 bytes per piece is lifted) */
static int cmp_u64(const void *a, const void *b) {
    const uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : x > y;
}

size_t orc_line_lengths(const uint8_t *data, size_t n, uint64_t *lengths, uint64_t *counts, size_t cap, uint64_t *n_lines) {
    size_t n_l = 0, alloc = 1024;
    uint64_t *lens = (uint64_t *)malloc(alloc * sizeof(uint64_t));
    size_t start = 0;
    for (size_t i = 0; i <= n; i++) {
        const int at_end = i == n;
        if (at_end ? (start < n) : (data[i] == 0x0A)) { /* a line feed ends a line; so does the end of the file */
            if (n_l == alloc) { alloc *= 2; lens = (uint64_t *)realloc(lens, alloc * sizeof(uint64_t)); }
            lens[n_l++] = (uint64_t)(i - start);
            start = i + 1;
        }
    }
    if (n_lines) *n_lines = n_l;
    qsort(lens, n_l, sizeof(uint64_t), cmp_u64);
    size_t n_distinct = 0;
    for (size_t i = 0; i < n_l;) {
        size_t j = i;
        while (j < n_l && lens[j] == lens[i]) j++;
        if (n_distinct < cap) { lengths[n_distinct] = lens[i]; counts[n_distinct] = (uint64_t)(j - i); }
        n_distinct++;
        i = j;
    }
    free(lens);
    return n_distinct;
}

/* ---- digest of a whole result (compared with wc_result_checksum of libwc.so) */
static uint64_t mix_u64(uint64_t x) {
    x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32;
    return x;
}

void orc_result_checksum(const uint8_t *data, const orc_result *r, uint64_t out[2]) {
    uint64_t sa = 0, sb = 0;
    for (size_t i = 0; i < r->n_rows; i++) {
        uint64_t h = 0xCBF29CE484222325ull;
        const uint8_t *w = data + r->rows[i].off;
        for (uint32_t k = 0; k < r->rows[i].len; k++) { h ^= w[k]; h *= 0x100000001B3ull; }
        sa += mix_u64(h) * (uint64_t)r->rows[i].count;
        sb += mix_u64(h + (uint64_t)i) * (r->rows[i].first + 1);
    }
    out[0] = sa; out[1] = sb;
}
