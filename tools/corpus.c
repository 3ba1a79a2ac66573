/*
 * tools/corpus.c -- synthetic text for the BASELINE.json configs (bench + tests).
 *
 * Not part of the product library and not derived from the reference (which
 * ships one 56-byte fixture).  It realises the generator invariants G1-G7 of
 * SURVEY.md section 8(d):
 *
 *   kind 0  "zipf": English-like text, V distinct words drawn Zipf(s) by rank.
 *           Single 0x20 between words, 0x0A line ends, every line 2..99 bytes
 *           including its newline, no blank lines / CR / NUL / tab, words of
 *           1..19 bytes.  On such text main.cu's 99-byte fgets chunking never
 *           splits a line, so the literal reference semantics and the
 *           capacity-lifted restatement coincide.
 *   kind 1  "adversarial": ~90 % of bytes in (almost always) unique 8..16 byte
 *           tokens over [a-z0-9], ~5 % one-character words, ~5 % in very long
 *           words (100 B .. 1 MiB); lines of hundreds of bytes.
 *
 * The corpus is a concatenation of fixed-size blocks; block b depends only on
 * (text_seed, b), so any range can be produced by any number of threads and
 * always has the same bytes.  Every block starts at a line start and ends
 * with 0x0A.
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ZIPF_BLOCK (1u << 20)
#define ADV_BLOCK 3000000u
#define MAX_WORD 19

typedef struct corpus {
    int kind;
    uint32_t vocab;
    double s;
    uint64_t vocab_seed, text_seed;
    /* vocabulary: packed words */
    uint8_t (*words)[MAX_WORD + 1]; /* [vocab][20], words[r][19] unused, len in lens */
    uint8_t *lens;
    uint32_t by_len[MAX_WORD + 1]; /* lowest rank having that length, or UINT32_MAX */
    size_t fill_max;               /* lengths 1..fill_max all exist in the vocabulary */
    /* alias tables for Zipf sampling */
    uint32_t *alias;
    uint32_t *thresh; /* probability scaled to 2^32 */
    /* adversarial: offsets that a very long word must cover */
    uint64_t straddle[16];
    int n_straddle;
} corpus;

static inline uint64_t mix64(uint64_t x) { /* splitmix64 finaliser: a bijection */
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

typedef struct { uint64_t key, ctr; } rng_t;
static inline uint64_t rng_next(rng_t *r) { return mix64(r->key + (r->ctr++) * 0xD1342543DE82EF95ull); }

/* ---- vocabulary -------------------------------------------------------- */

static const char LETTER_FREQ[] = /* rough English letter frequencies, 100 slots */
    "eeeeeeeeeeeetttttttttaaaaaaaaooooooooiiiiiiinnnnnnnsssssshhhhhhrrrrrrddddlllluuucccmmmwwffggyyppbbvkjxqz";

static uint64_t word_key(const uint8_t *w, int len) {
    uint64_t h = 1469598103934665603ull;
    for (int i = 0; i < len; i++) h = (h ^ w[i]) * 1099511628211ull;
    return mix64(h ^ (uint64_t)len);
}

static int build_vocab(corpus *c) {
    uint32_t V = c->vocab;
    c->words = calloc(V, sizeof(*c->words));
    c->lens = calloc(V, 1);
    size_t cap = 1; while (cap < (size_t)V * 4) cap <<= 1;
    uint64_t *set = calloc(cap, sizeof(uint64_t)); /* open addressing on word_key, 0 = empty */
    if (!c->words || !c->lens || !set) return -1;
    for (int l = 0; l <= MAX_WORD; l++) c->by_len[l] = UINT32_MAX;
    size_t nfreq = strlen(LETTER_FREQ);
    for (uint32_t r = 0; r < V; r++) {
        rng_t g = {mix64(c->vocab_seed ^ 0xA5A5u) + r * 0x9E3779B97F4A7C15ull, 0};
        for (int attempt = 0;; attempt++) {
            uint8_t w[MAX_WORD + 1]; int len = 0;
            uint64_t x = rng_next(&g);
            if (r == 1 && attempt == 0) { w[0] = 'a'; len = 1; goto have_word; } /* "a": a 1-letter word must exist */
            /* 15 %: extend an earlier word (the -> then, the, the.) so prefix relations exist */
            if (r >= 8 && (x & 0xFF) < 38) {
                uint32_t base = (uint32_t)((x >> 8) % r);
                len = c->lens[base];
                memcpy(w, c->words[base], (size_t)len);
                uint8_t last = w[len - 1];
                int kind = (int)((x >> 40) % 4);
                if (last == ',' || last == '.') kind = 3; /* never extend past punctuation */
                if (kind == 0 && len < MAX_WORD) w[len++] = ',';
                else if (kind == 1 && len < MAX_WORD) w[len++] = '.';
                else if (kind == 2 || kind == 3) {
                    if (last == ',' || last == '.') len--; /* then the, -> the + letters */
                    int extra = 1 + (int)((x >> 44) % 3);
                    uint64_t y = rng_next(&g);
                    for (int k = 0; k < extra && len < MAX_WORD; k++, y >>= 8)
                        w[len++] = (uint8_t)LETTER_FREQ[(y & 0xFF) % nfreq];
                }
            } else {
                /* length grows with the log of the rank: short function words first */
                double want = 2.0 + 0.62 * log((double)r + 2.0) + ((double)((x >> 8) & 0xFFFF) / 65536.0 - 0.5) * 3.0;
                len = (int)(want + 0.5) + attempt / 4;
                if (len < 2) len = 2; /* the only 1-letter word is rank 1 ("a") */
                if (r == 0) len = 3;  /* rank 0 plays "the" */
                if (len > MAX_WORD - 1) len = MAX_WORD - 1;
                uint64_t y = 0;
                for (int k = 0; k < len; k++) {
                    if ((k & 7) == 0) y = rng_next(&g);
                    w[k] = (uint8_t)LETTER_FREQ[(y & 0xFF) % nfreq];
                    y >>= 8;
                }
                if (((x >> 24) & 0xFF) < 13) w[0] = (uint8_t)(w[0] - 'a' + 'A'); /* ~5 % capitalised */
            }
        have_word:
            if (len < 1 || len > MAX_WORD) continue;
            uint64_t k = word_key(w, len); if (k == 0) k = 1;
            size_t i = (size_t)k & (cap - 1); int dup = 0;
            while (set[i]) { if (set[i] == k) { dup = 1; break; } i = (i + 1) & (cap - 1); }
            if (dup) continue; /* (a 64-bit key clash only makes us skip a candidate) */
            set[i] = k;
            memcpy(c->words[r], w, (size_t)len);
            c->lens[r] = (uint8_t)len;
            if (c->by_len[len] == UINT32_MAX) c->by_len[len] = r;
            break;
        }
    }
    free(set);
    c->fill_max = 0;
    while (c->fill_max < 8 && c->by_len[c->fill_max + 1] != UINT32_MAX) c->fill_max++;
    return c->fill_max >= 2 ? 0 : -2; /* the block filler needs lengths 1..2 at least */
}

/* Vose alias method over p(k) ~ (k+1)^-s */
static int build_alias(corpus *c) {
    uint32_t V = c->vocab;
    double *p = malloc(sizeof(double) * V);
    uint32_t *small = malloc(sizeof(uint32_t) * V), *large = malloc(sizeof(uint32_t) * V);
    c->alias = malloc(sizeof(uint32_t) * V);
    c->thresh = malloc(sizeof(uint32_t) * V);
    if (!p || !small || !large || !c->alias || !c->thresh) return -1;
    double z = 0;
    for (uint32_t k = 0; k < V; k++) { p[k] = pow((double)k + 1.0, -c->s); z += p[k]; }
    uint32_t ns = 0, nl = 0;
    for (uint32_t k = 0; k < V; k++) { p[k] = p[k] / z * V; if (p[k] < 1.0) small[ns++] = k; else large[nl++] = k; }
    while (ns && nl) {
        uint32_t a = small[--ns], b = large[--nl];
        double t = p[a] * 4294967296.0;
        c->thresh[a] = t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)t;
        c->alias[a] = b;
        p[b] = (p[b] + p[a]) - 1.0;
        if (p[b] < 1.0) small[ns++] = b; else large[nl++] = b;
    }
    while (nl) { uint32_t b = large[--nl]; c->thresh[b] = 0xFFFFFFFFu; c->alias[b] = b; }
    while (ns) { uint32_t a = small[--ns]; c->thresh[a] = 0xFFFFFFFFu; c->alias[a] = a; }
    free(p); free(small); free(large);
    return 0;
}

static inline uint32_t sample_rank(const corpus *c, uint64_t x) {
    uint32_t col = (uint32_t)(((x >> 32) * (uint64_t)c->vocab) >> 32);
    return ((uint32_t)x <= c->thresh[col]) ? col : c->alias[col];
}

/* ---- block writers ----------------------------------------------------- */

/* Fill exactly `rem` (>= 2) bytes with well-formed lines made of vocabulary words. */
static uint8_t *fill_exact(const corpus *c, uint8_t *p, size_t rem) {
    while (rem > 0) {
        size_t line = rem < 99 ? rem : 99;
        if (rem - line == 1) line -= 1; /* never leave a 1-byte line (it would be blank) */
        size_t content = line - 1;
        while (content > 0) {
            size_t w = content < c->fill_max ? content : c->fill_max;
            if (content - w == 1) w -= 1; /* room for " x" or nothing */
            uint32_t r = c->by_len[w];
            memcpy(p, c->words[r], w); p += w; content -= w;
            if (content > 0) { *p++ = ' '; content--; }
        }
        *p++ = '\n';
        rem -= line;
    }
    return p;
}

static void zipf_block(const corpus *c, uint64_t b, uint8_t *dst, size_t nbytes) {
    rng_t g = {mix64(c->text_seed) ^ mix64(b * 0x632BE59BD9B4E019ull + 1), 0};
    uint8_t *p = dst, *end = dst + nbytes;
    size_t col = 0, width = 0;
    while ((size_t)(end - p) > 64) {
        uint64_t x = rng_next(&g);
        if (col == 0) width = 40 + (size_t)((x >> 52) % 59); /* 40..98 characters before the newline */
        uint32_t r = sample_rank(c, x);
        size_t len = c->lens[r];
        if (col && col + 1 + len > width) { *p++ = '\n'; col = 0; width = 40 + (size_t)((x >> 52) % 59); }
        if (col) { *p++ = ' '; col++; }
        memcpy(p, c->words[r], len); p += len; col += len;
    }
    if (col) *p++ = '\n';
    fill_exact(c, p, (size_t)(end - p));
}

static const char B36[] = "abcdefghijklmnopqrstuvwxyz0123456789";
static const char B62[] = "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789";

static void adv_block(const corpus *c, uint64_t b, uint64_t block_off, uint8_t *dst, size_t nbytes) {
    rng_t g = {mix64(c->text_seed ^ 0xADull) ^ mix64(b * 0x632BE59BD9B4E019ull + 7), 0};
    uint8_t *p = dst, *end = dst + nbytes;
    /* one very long word per block: log-uniform 100 B .. 1 MiB, ~5 % of bytes on average */
    uint64_t x0 = rng_next(&g);
    double u = (double)(x0 >> 11) / 9007199254740992.0;
    size_t llen = (size_t)(100.0 * pow(10485.76, u * u)); /* skewed toward the short end */
    if (llen > nbytes / 3) llen = nbytes / 3;
    size_t lpos = nbytes / 2 - llen / 2;
    for (int i = 0; i < c->n_straddle; i++) {
        uint64_t s = c->straddle[i];
        if (s > block_off + 200 && s + 200 < block_off + nbytes) {
            if (llen < 400) llen = 400;
            lpos = (size_t)(s - block_off) - llen / 2;
            if ((size_t)(s - block_off) < llen / 2 + 100) lpos = 100;
            if (lpos + llen + 100 > nbytes) llen = nbytes - 100 - lpos;
        }
    }
    if (nbytes < 4096) llen = 0;
    size_t col = 0, width = 0; int long_done = (llen == 0);
    uint64_t uniq = b << 24;
    while ((size_t)(end - p) > 2200 + (long_done ? 0 : llen)) {
        uint64_t x = rng_next(&g);
        if (col == 0) width = 200 + (size_t)((x >> 50) % 1800);
        if (!long_done && (size_t)(p - dst) + 1 >= lpos) {
            if (col) { *p++ = ' '; }
            uint64_t y = 0;
            for (size_t k = 0; k < llen; k++) { if ((k & 7) == 0) y = mix64(x + k); p[k] = (uint8_t)B36[(y & 0xFF) % 36]; y >>= 8; }
            p += llen; col += llen + 1; long_done = 1;
            continue;
        }
        size_t len; uint8_t tok[16];
        if ((x & 0xFF) < 51) { /* ~20 % of tokens are 1 byte: ~5 % of bytes */
            tok[0] = (uint8_t)B62[(x >> 8) % 62]; len = 1;
        } else {
            len = 8 + (size_t)((x >> 8) % 9);
            uint64_t y = mix64(uniq++), y2 = mix64(y);
            for (size_t k = 0; k < len; k++) {
                uint64_t *src = k < 10 ? &y : &y2;
                tok[k] = (uint8_t)B36[*src % 36]; *src /= 36;
            }
        }
        if (col && col + 1 + len > width) { *p++ = '\n'; col = 0; width = 200 + (size_t)((x >> 50) % 1800); }
        if (col) { *p++ = ' '; col++; }
        memcpy(p, tok, len); p += len; col += len;
    }
    if (col) *p++ = '\n';
    /* filler: lines of one-character words, exact length */
    size_t rem = (size_t)(end - p);
    while (rem > 0) {
        size_t line = rem < 1000 ? rem : 1000;
        if (rem - line == 1) line -= 1;
        size_t content = line - 1;
        while (content > 0) {
            size_t w = content >= 3 ? 1 : content; /* "x y" ... ends with a 1- or 2-char word */
            if (content - w == 1) w = content;
            for (size_t k = 0; k < w; k++) *p++ = 'z';
            content -= w;
            if (content > 0) { *p++ = ' '; content--; }
        }
        *p++ = '\n';
        rem -= line;
    }
}

/* ---- public API ---------------------------------------------------------- */

corpus *corpus_create(int kind, uint32_t vocab, double s, uint64_t vocab_seed, uint64_t text_seed) {
    corpus *c = calloc(1, sizeof(corpus));
    if (!c) return NULL;
    c->kind = kind; c->vocab = vocab; c->s = s; c->vocab_seed = vocab_seed; c->text_seed = text_seed;
    if (kind == 0) {
        if (vocab < 64 || build_vocab(c) != 0 || build_alias(c) != 0) { free(c); return NULL; }
    }
    return c;
}

void corpus_destroy(corpus *c) {
    if (!c) return;
    free(c->words); free(c->lens); free(c->alias); free(c->thresh); free(c);
}

size_t corpus_block_bytes(const corpus *c) { return c->kind == 0 ? ZIPF_BLOCK : ADV_BLOCK; }

int corpus_add_straddle(corpus *c, uint64_t offset) {
    if (c->n_straddle >= 16) return -1;
    c->straddle[c->n_straddle++] = offset;
    return 0;
}

const uint8_t *corpus_word(const corpus *c, uint32_t rank, uint32_t *len) {
    if (c->kind != 0 || rank >= c->vocab) return NULL;
    *len = c->lens[rank];
    return c->words[rank];
}

/* The tables the Zipf block writer reads, for the device port of it (tools/corpus_gpu.cu). */
int corpus_tables(const corpus *c, const uint8_t **words, const uint8_t **lens, const uint32_t **by_len,
                  uint64_t *fill_max, const uint32_t **alias, const uint32_t **thresh, uint32_t *vocab,
                  uint64_t *text_seed) {
    if (c->kind != 0) return -1;
    *words = &c->words[0][0]; *lens = c->lens; *by_len = c->by_len; *fill_max = c->fill_max;
    *alias = c->alias; *thresh = c->thresh; *vocab = c->vocab; *text_seed = c->text_seed;
    return 0;
}

typedef struct { const corpus *c; uint8_t *dst; uint64_t off, total; size_t n; uint64_t b0, b1; int tid, nt; } fill_job;

static void *fill_main(void *arg) {
    fill_job *j = arg;
    size_t B = corpus_block_bytes(j->c);
    for (uint64_t b = j->b0 + (uint64_t)j->tid; b < j->b1; b += (uint64_t)j->nt) {
        uint64_t lo = b * B, hi = lo + B > j->total ? j->total : lo + B;
        uint8_t *d = j->dst + (lo - j->off);
        if (j->c->kind == 0) zipf_block(j->c, b, d, (size_t)(hi - lo));
        else adv_block(j->c, b, lo, d, (size_t)(hi - lo));
    }
    return NULL;
}

/* Writes bytes [off, off+n) of the `total`-byte corpus into dst.  off must be a
 * multiple of the block size and off+n must be total or a multiple of it. */
int corpus_fill(const corpus *c, uint8_t *dst, uint64_t off, size_t n, uint64_t total, int nthreads) {
    size_t B = corpus_block_bytes(c);
    if (off % B) return -1;
    if (off + n != total && (off + n) % B) return -2;
    if (off + n > total) return -3;
    if (total % B == 1 || (total % B && total % B < 2)) return -4;
    if (c->kind == 1 && total % B && total % B < 4096) return -5;
    if (n == 0) return 0;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    uint64_t b0 = off / B, b1 = (off + n + B - 1) / B;
    pthread_t th[256]; fill_job jobs[256];
    for (int t = 0; t < nthreads; t++) {
        fill_job j = {c, dst, off, total, n, b0, b1, t, nthreads};
        jobs[t] = j;
        pthread_create(&th[t], NULL, fill_main, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    return 0;
}
