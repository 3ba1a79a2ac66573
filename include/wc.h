/*
 * wc.h -- C ABI of libwc.so, a B200-native (sm_100a) word-count engine that is a
 * drop-in for the tokenize -> count path of zimisoho/cuda-mapreduce.
 *
 * What it replaces in the reference (all in /root/reference/main.cu):
 *   - the host splitter                         main.cu:174-206
 *   - mapper / mapKernel / cudaMap              main.cu:37-54, 109-117, 124-127
 *   - compare / reducer / reduceKernel / cudaReduce   main.cu:57-108, 119-123, 129-132
 *   - the driver runMapReduce()                 main.cu:133-162
 * The engine reads raw text bytes and produces the rows the reference prints
 * (main.cu:210-218), bit for bit, with the reference's fixed capacities
 * (main.cu:9-15) lifted.  Semantics: SURVEY.md Appendix A.
 *
 * Conventions: plain pointers and sizes, no C++ or torch types.  Every call
 * returns 0 (WC_OK) or a negative wc_status; nothing prints or exits.  The
 * engine owns all device and pinned memory it allocates; a wc_result owns the
 * host memory its accessors point into.  One engine is used by one host
 * thread at a time.  Counts and offsets are 64-bit.
 */
#ifndef WC_H
#define WC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define WC_API __attribute__((visibility("default")))
#else
#define WC_API
#endif

typedef struct wc_engine wc_engine;
typedef struct wc_result wc_result;

typedef enum wc_status {
    WC_OK = 0,
    WC_ERR_INVALID = -1,    /* bad argument / call order */
    WC_ERR_CUDA = -2,       /* a CUDA runtime call failed (wc_last_error has the text) */
    WC_ERR_NOMEM = -3,      /* host or device allocation failed */
    WC_ERR_TABLE_FULL = -4, /* a device table or the long-word arena could not grow further */
    WC_ERR_UNALIGNED = -5,  /* device buffer not 16-byte aligned */
    WC_ERR_NCCL = -6,
    WC_ERR_TOO_LARGE = -7   /* size beyond what this entry point supports */
} wc_status;

/* Reduce semantics. */
#define WC_MODE_REF 0   /* main.cu's reducer: first match by PREFIX in insertion order (main.cu:57-67,
                           81-89), empty words credited to row 0 (main.cu:81,103)  */
#define WC_MODE_EXACT 1 /* plain equality counts of non-empty words */

/* WC_FLAG_LANE_SERIAL (wc_config.flags): count forward launches of the hash path with the alternative
 * tokenizer kernel count_kernel2 (csrc/wc_count2.cuh: every warp tokenizes, no consumer warps; words that
 * miss the CTA's table are listed per tile and settled 32 at a time).  Same results; on the measured B200
 * it is slower than the default kernel on config 2 (1.86 ms vs 1.47 ms per 1e9 bytes) -- both end up bound
 * by shared-memory wavefronts, see DESIGN.md section 7 -- so it is off by default. */
#define WC_FLAG_LANE_SERIAL 2u

/* Aggregation path (wc_config.path). */
#define WC_PATH_AUTO 0 /* hash tables, unless most words are distinct: the first job of an engine is probed (the head of
                          its first piece is counted, the counters read, the tables reset), later jobs follow the ratio
                          of the job before them; the sort path is taken above WC_AUTO_SORT_RATIO (environment).  Its
                          default is 2.0 = never: on one B200 the hash path won at every ratio measured, up to 0.80
                          (profiles/r2_path_crossover.txt).  wc_stats.path_used says which one ran. */
#define WC_PATH_HASH 1 /* shared-memory table spilling to a global open-addressing table */
#define WC_PATH_SORT 2 /* radix sort of (hash, offset, length) + segmented reduce */

/* wc_config.flags */
#define WC_FLAG_STRICT_REF 1u /* read lines the way main.cu does: in pieces of at most 99 bytes (fgets(szLine, 100),
                                 main.cu:175,179).  On a line longer than that the word that straddles a piece edge
                                 loses its head, 0x0D / 0x00 hide only the rest of the piece, and a one-byte piece
                                 (line length = 1 mod 99) stops the input (main.cu:185).  Without the flag a line is
                                 one piece however long it is; the two agree on every line of at
                                 most 99 bytes. */

/* wc_config.mapper: what a word is mapped to before it is counted.  The reference's mapper()
 * (main.cu:37-54) emits (word, 1) for the word as it is: WC_MAP_IDENTITY.  Other mappers are
 * byte-wise, length-preserving transforms applied inside the kernel wherever a word's bytes are read;
 * they are compiled-in policies (csrc/wc_device.cuh, "mappers"), this field selects one.  The reducer
 * side of the skeleton is wc_config.mode. */
#define WC_MAP_IDENTITY 0u
#define WC_MAP_LOWER_ASCII 1u /* 'A'..'Z' counted as 'a'..'z' (a case-insensitive word count) */

/* wc_config.job: which map/reduce pair runs over the input.  The reference's skeleton takes a user mapper()
 * (main.cu:37) and reducer() (main.cu:69); the word count is the one pair it ships.
 *   WC_JOB_WORDCOUNT     map: word -> (word, 1); reduce: count per word, merged by main.cu's prefix rule
 *                        (WC_MODE_REF) or plain equality (WC_MODE_EXACT).  Everything above describes it.
 *   WC_JOB_LINE_LENGTHS  map: every line (what fgets delivers, main.cu:179; the final one may lack its
 *                        line feed) -> (its length in bytes without the line feed, 1); reduce: count per
 *                        length.  Same engine calls (wc_feed / wc_count_device pieces must end at line ends,
 *                        as for the word count); wc_finish returns one row per length that occurs, the
 *                        key rendered in decimal, rows in increasing length, total = number of lines.
 *                        mode, path, mapper and flags do not apply (no STOP rule: every line counts). */
#define WC_JOB_WORDCOUNT 0u
#define WC_JOB_LINE_LENGTHS 1u

typedef struct wc_config {
    uint32_t struct_size;      /* = sizeof(wc_config) */
    int32_t device;            /* CUDA device ordinal */
    int32_t mode;              /* WC_MODE_* */
    int32_t path;              /* WC_PATH_* */
    uint64_t chunk_bytes;      /* wc_feed() streaming chunk (0: 64 MiB) */
    uint32_t short_table_log2; /* global table slots for words of 1..16 bytes (0: 2^21; at most 32) */
    uint32_t long_table_log2;  /* global table slots for words of 17+ bytes (0: 2^18) */
    uint64_t arena_bytes;      /* device byte store for the 17+ byte words (0: 256 MiB) */
    uint64_t base_offset;      /* offset of this engine's first byte in the whole input (shards) */
    uint32_t flags;            /* WC_FLAG_* */
    uint32_t mapper;           /* WC_MAP_* */
    uint32_t job;              /* WC_JOB_* */
    uint32_t reserved;         /* 0 */
} wc_config;

/* Counters of one finished job; the same numbers the oracle reports. */
typedef struct wc_stats {
    uint64_t bytes;          /* input bytes consumed */
    uint64_t n_flush;        /* words the splitter emitted, empty ones included (main.cu:190-194) */
    uint64_t n_empty;        /* of which empty */
    uint64_t n_lost_empty;   /* empties before the first non-empty word (lost, main.cu:81,103) */
    uint64_t stop;           /* offset at which input processing stopped (main.cu:185), == bytes if none */
    uint64_t n_distinct;     /* distinct non-empty words counted */
    uint64_t n_long_words;   /* words of 9+ bytes (they bypass the shared-memory table) */
    uint64_t n_spilled;      /* short words that bypassed the shared-memory table */
    uint64_t n_slow_tiles;   /* 1 KiB tiles that held a control byte and took the exact line-state path */
    uint64_t table_grows;    /* times a device table was re-hashed into a larger one */
    double device_ms;        /* device time of this job's kernels (CUDA events on the engine's stream) */
    double count_kernel_ms;  /* of which the tokenizer/count kernel alone */
    uint64_t path_used;      /* WC_PATH_HASH or WC_PATH_SORT: what wc_config.path = WC_PATH_AUTO chose for this job */
    double distinct_ratio;   /* distinct words / non-empty words of this job (what WC_PATH_AUTO goes by) */
    uint64_t tokenizing_warps; /* of the count kernel's 28 warps per CTA, how many tokenized in this job's last launch (the
                                * others insert into the device tables): 20, or 18 once the engine has seen text that sends
                                * 22-60 % of its words there; 28 with WC_FLAG_LANE_SERIAL */
} wc_stats;

WC_API void wc_config_default(wc_config *cfg);

/* Replaces the allocation half of runMapReduce (main.cu:143-145,153-154). */
WC_API int wc_create(wc_engine **out, const wc_config *cfg);
WC_API void wc_destroy(wc_engine *e);

/* Forget all counts; keeps the allocations.  Start of a new job. */
WC_API int wc_reset(wc_engine *e);

/* Feed the next `n` consecutive bytes of the input from HOST memory.  Replaces
 * the fgets loop + cudaMemcpy H2D (main.cu:176-206, 147): bytes go through a
 * pinned, double-buffered cudaMemcpyAsync pipeline overlapped with the
 * tokenizer kernel.  `bytes` may be pageable or pinned; it is free on return. */
WC_API int wc_feed(wc_engine *e, const void *bytes, size_t n);

/* Count `n` bytes that are already in device memory (16-byte aligned, same
 * device).  The piece must start at a line start of the input; every piece but
 * the last must end with '\n'.  Asynchronous on the engine's stream; the
 * buffer must stay valid until wc_sync()/wc_finish().  Bytes 8..15 of longer
 * words are fetched in aligned 8-byte units, so up to 7 bytes past `n` (to the
 * next multiple of 8) may be read, never used: the allocation must cover them,
 * as every cudaMalloc'd buffer does. */
WC_API int wc_count_device(wc_engine *e, const void *dev_bytes, size_t n, int last_piece);

/* No more input will come: counts the unterminated last line, if any (its final word is dropped
 * like in main.cu:198-203).  wc_finish and wc_shard_stats_get call it themselves. */
WC_API int wc_end_input(wc_engine *e);

/* Completes the device-side aggregation of what has been fed so far.  On the hash path the tables
 * are always aggregated and this does nothing; on the sort path (WC_PATH_SORT) it runs the radix sort
 * and the run reduce and leaves (word, count, first) triples in device memory.  wc_finish calls it. */
WC_API int wc_aggregate(wc_engine *e);

/* Wait for everything queued so far. */
WC_API int wc_sync(wc_engine *e);

/* End of input: drains the pipeline, applies the reduce semantics (mode) and
 * returns the rows.  Replaces cudaReduce + the D2H copies (main.cu:155-158). */
WC_API int wc_finish(wc_engine *e, wc_result **out);

/* Rows are in the reference's output order (insertion order, main.cu:212). */
WC_API size_t wc_result_rows(const wc_result *r);
WC_API int wc_result_row(const wc_result *r, size_t i, const char **word, size_t *len, int64_t *count);
WC_API uint64_t wc_result_row_first(const wc_result *r, size_t i); /* offset that flushed its first occurrence */
WC_API int64_t wc_result_total(const wc_result *r);              /* "Total Count", main.cu:214,218 */
WC_API int wc_result_stats(const wc_result *r, wc_stats *out);
/* A digest of all rows for results too large to compare row by row (config 5: 2.9e8 rows):
 * out[0] = sum over rows i of mix(fnv1a64(word_i)) * count_i, out[1] = sum of mix(fnv1a64(word_i) + i) * (first_i + 1),
 * both modulo 2^64 -- the first says "the same multiset of (word, count)", the second "in the same order, with
 * the same first offsets".  The oracle computes the same two numbers (orc_result_checksum). */
WC_API int wc_result_checksum(const wc_result *r, uint64_t out[2]);
WC_API void wc_result_free(wc_result *r);

/* Writes what main.cu prints after the echo (main.cu:210-218): separator, rows
 * "word\tcount\n", separator, "Total Count:N\n".  Returns bytes needed. */
WC_API size_t wc_result_render(const wc_result *r, char *buf, size_t cap);

WC_API const char *wc_strerror(int status);
WC_API const char *wc_last_error(const wc_engine *e);

/* Pinned host memory for zero-staging wc_feed(). */
WC_API int wc_host_alloc(void **p, size_t n);
WC_API int wc_host_free(void *p);

/* Device time since the last wc_reset(), measured with CUDA events on the engine's stream (call
 * after wc_sync/wc_finish): all kernels of the job, the tokenizer/count kernel alone, and how many
 * kernels were launched. */
WC_API int wc_device_ms(wc_engine *e, double *ms, double *count_kernel_ms, uint64_t *kernel_launches);

/* Bytes the last wc_finish() copied from the device to the host (counters + the used table entries
 * and their two sort orders; the tables themselves stay on the device).  No reference counterpart:
 * main.cu:158 copies its ten output rows unconditionally. */
WC_API uint64_t wc_finish_d2h_bytes(const wc_engine *e);

/* The CUDA stream (cudaStream_t) the engine launches on, for callers that
 * bracket work with their own events. */
WC_API void *wc_stream(wc_engine *e);

/* ---- multi-GPU: byte-range shards + key-owner exchange ---------------------------
 *
 * Nothing to replace in the reference: it drives one GPU (main.cu:133-162).  One engine per GPU counts
 * one byte range of the input (ranges are cut at line starts; wc_config.base_offset = where the range
 * begins, so that first-occurrence offsets stay global).  Afterwards:
 *   1. wc_shard_stats_get on every engine; wc_shard_stats_reduce over all of them (any host);
 *      an engine whose range begins at or after the job's STOP calls wc_clear_tables.
 *   2. wc_export_counts / wc_export_words: the word table as 32-byte records {key0, key1, count,
 *      first} grouped by owner = wc_key_owner(key, n_parts); the caller runs the all-to-all (NCCL).
 *   3. every engine: wc_clear_tables, then wc_import_words of what it received -> each distinct word
 *      is reduced on exactly one GPU.
 *   4. gather: owners export again (n_parts = 1) and the gathering engine imports; words of 17+ bytes
 *      travel once, as wc_export_long / wc_import_long.
 *   5. wc_shard_stats_set(job counters) + wc_finish on the gathering engine.
 * The transport is the caller's: bench.py uses torch.distributed (NCCL over NVLink). */
typedef struct wc_shard_stats {
    uint64_t base, bytes;      /* the shard: [base, base + bytes) of the input */
    uint64_t n_flush, n_empty, lost_empty;
    uint64_t p0;               /* offset that flushed the shard's first non-empty word (if p0_found) */
    uint64_t stop;             /* absolute STOP offset seen in this shard, ~0 if none */
    uint32_t p0_found, reserved;
} wc_shard_stats;
#define WC_WORD_RECORD_BYTES 32
#define WC_LONG_RECORD_BYTES 32
/* Where this engine's byte range begins in the whole input (wc_config.base_offset), for the next
 * job of an engine that is reused for another shard.  Only right after wc_create / wc_reset. */
WC_API int wc_set_base_offset(wc_engine *e, uint64_t base_offset);
WC_API int wc_shard_stats_get(wc_engine *e, wc_shard_stats *out);
WC_API int wc_shard_stats_reduce(const wc_shard_stats *shards, size_t n, wc_shard_stats *job);
WC_API int wc_shard_stats_set(wc_engine *e, const wc_shard_stats *job);
WC_API int wc_clear_tables(wc_engine *e);
WC_API int wc_export_counts(wc_engine *e, uint32_t n_parts, uint64_t *word_counts, uint64_t *long_entries,
                            uint64_t *long_blob_bytes);
WC_API int wc_export_words(wc_engine *e, uint32_t n_parts, const uint64_t *part_offsets, void *dev_records);
WC_API int wc_export_long(wc_engine *e, void *dev_meta, void *dev_blob);
WC_API int wc_import_words(wc_engine *e, const void *dev_records, uint64_t n_records);
WC_API int wc_import_long(wc_engine *e, const void *dev_meta, uint64_t n_entries, const void *dev_blob);
WC_API uint32_t wc_key_owner(uint64_t k0, uint64_t k1, uint32_t n_parts);

/* ---- peer gather: steps 2-5 above in one kernel per sender, over NVLink peer memory ------------------
 *
 * No reference counterpart (one GPU, main.cu:133-162).  For GPUs of ONE node.  The gathering ("root")
 * engine owns a landing area in its GPU's memory with one slot per rank; every other engine maps it
 * (CUDA IPC between processes, or directly inside one process) and wc_peer_push() scans its tables
 * once, storing each used entry -- and, last, a header with its shard counters -- straight into its
 * slot on the root GPU.  No staging buffer and no size exchange; the stores cross NVLink while the
 * scan is still running.  After a barrier between the ranks (the caller's: any stream-ordered
 * collective behind the push, e.g. a one-word NCCL all-reduce), wc_peer_pull() on the root reduces
 * the shard counters, adds the entries of every rank that lies before the job's STOP to its own
 * tables and installs the job counters; wc_finish() then gives the rows.
 * Slots are sized by the caller (a slot of words_per_rank = the sender's table size can never
 * overflow); a sender whose tables overflowed says so in its header and wc_peer_pull() fails with
 * WC_ERR_TOO_LARGE -- fall back to the exchange above.  dev_flag (optional, device memory, 4 bytes)
 * receives 1 if this sender has such a problem, for callers that fold it into their barrier. */
#define WC_PEER_HANDLE_BYTES 64
WC_API int wc_peer_root_create(wc_engine *root, uint32_t n_ranks, uint32_t root_rank, uint64_t words_per_rank,
                               uint64_t long_per_rank, uint64_t blob_bytes_per_rank, void *handle_out);
WC_API int wc_peer_open(wc_engine *e, uint32_t my_rank, uint32_t n_ranks, uint32_t root_rank, uint64_t words_per_rank,
                        uint64_t long_per_rank, uint64_t blob_bytes_per_rank, const void *handle);
WC_API int wc_peer_attach(wc_engine *e, uint32_t my_rank, wc_engine *root); /* same process */
WC_API int wc_peer_push(wc_engine *e, void *dev_flag);                      /* asynchronous, on the engine's stream */
WC_API int wc_peer_pull(wc_engine *root, wc_shard_stats *job_out);
/* The same without the caller's barrier and without a host round trip, for ranks that each have a GPU of
 * their own: queues, on the root's stream, a one-block kernel that waits until every sender's header of
 * this job has arrived in the root GPU's memory and reduces the shard counters there, then the import.
 * Nothing is returned to the host: the job's counters are in the engine (wc_finish reports them) and a
 * sender that overflowed, or whose push never arrived, makes wc_finish fail with WC_ERR_TOO_LARGE.
 * Landing areas are double buffered and a sender two jobs ahead of the root waits for it, so jobs can be
 * queued back to back with no synchronisation between the ranks at all.  Do NOT use it when a sender
 * shares the root's GPU (the waiting kernel could then keep the sender's kernel from running): there,
 * order the push before wc_peer_pull() yourself. */
WC_API int wc_peer_pull_device(wc_engine *root);

/* ---- wc_count_device_static: counting without any allocation -------------------------------
 *
 * No counterpart in the reference (whose runMapReduce cudaMallocs and cudaFrees on every call,
 * main.cu:143-161).  It exists for callers that cannot allocate device memory at all: programs
 * that gave the whole device heap to their own pool, and hot paths where the device-wide stall of
 * cudaMalloc/cudaFree is not acceptable.  It runs the same tokenizer/count kernel as
 * wc_count_device(), but the kernel's working state sits in a fixed array inside the library image
 * (about 106 KiB, resident on every GPU that loads libwc.so until the process exits) and the caller
 * provides the text and every result buffer.  The call makes no cudaMalloc/cudaFree and no other
 * allocating call, on the device or on the host.
 *
 *   dev_text          n bytes of text in device memory, 16-byte aligned; the WHOLE input (it is
 *                     treated as starting at a line start and ending the input).
 *   dev_short_table   4 << short_log2 uint64, 16-byte aligned: keys (2 per slot) | counts | ~first-offset
 *                     of words of 1..16 bytes (a key is the word itself, zero padded to 16 bytes).
 *   dev_long_table    3 << long_log2 uint64: refs | counts | ~first-offset of words of 17+ bytes; a
 *                     ref is (24-bit tag << 40) | unit index into dev_arena.
 *   dev_arena         arena_bytes of storage for the 17+ byte words: unit[i] = length in bytes,
 *                     unit[i+1...] = the bytes, zero padded to 8.
 *   dev_counters      WC_STATIC_COUNTERS uint64 written at the end: n_flush, n_empty, stop (~0 if
 *                     none), p0 (~0 if none), lost_empty, p0_found, err (0 = ok, else a table or the
 *                     arena was too small), short slots used, long slots used, arena units used,
 *                     long words, spilled words.
 *   stream            a cudaStream_t (NULL = the default stream); everything is asynchronous on it.
 *
 * The tables are cleared by the call itself, so the results describe this text alone.  Limits:
 * n <= WC_STATIC_MAX_BYTES and a GPU of at most 160 SMs; anything larger is refused with
 * WC_ERR_TOO_LARGE before any work is queued (wc_static_last_error() says which).
 *
 * The first call on each device looks up the array's device address and raises the kernel's
 * shared-memory limit; make that first call before starting any stream capture.  All calls on a
 * device share the one array and it keeps whatever the previous call left, so the call resets it
 * first -- which also means two calls whose GPU work overlaps (different streams, different
 * threads) corrupt each other: run them one after the other. */
#define WC_STATIC_MAX_BYTES (1ull << 36)
#define WC_STATIC_COUNTERS 12
WC_API int wc_count_device_static(int device, const void *dev_text, size_t n, uint64_t *dev_short_table,
                                  uint32_t short_log2, uint64_t *dev_long_table, uint32_t long_log2,
                                  uint64_t *dev_arena, uint64_t arena_bytes, uint64_t *dev_counters,
                                  void *stream);
WC_API const char *wc_static_last_error(void);

/* Test hook (no GPU needed): runs the host half of the reduce -- the order-free form of
 * main.cu:57-108 -- on a caller-supplied table of exact counts.  `words` is the concatenation of
 * the n distinct words. */
WC_API int wc_debug_finalize(const char *words, const uint32_t *lens, const int64_t *counts,
                             const uint64_t *firsts, size_t n, int mode, uint64_t n_empty,
                             uint64_t n_lost_empty, wc_result **out);

/* Test hooks (no GPU needed): the tokenizer's SWAR byte classification (delimiters of main.cu:188,
 * line feeds, the 0x0D / 0x00 bytes that end a line's scan, main.cu:187,195) on 16 bytes, bit p of
 * each mask = byte p; and the shared-memory table slot of a packed key. */
WC_API int wc_debug_classify16(const uint8_t *bytes16, uint32_t *delim_mask, uint32_t *lf_mask,
                               uint32_t *kill_mask);
WC_API uint32_t wc_debug_smem_slot(uint64_t key);
WC_API uint32_t wc_debug_smem_entries(void);
WC_API int wc_debug_counters(wc_engine *e, uint64_t *out8); /* cycle counters of -DWC_PROFILE builds */

#ifdef __cplusplus
}
#endif
#endif
