// wc_device.cuh -- sm_100a tokenizer + counting kernels of libwc.so.
//
// One pass over raw text bytes replaces the reference's host splitter
// (/root/reference/main.cu:174-206), mapper (main.cu:37-54) and reducer
// (main.cu:69-108): delimiters are found with SWAR byte tests, word boundaries
// with a warp scan, every word's key is taken from a shared-memory copy of the
// tile and counted in a per-CTA shared-memory hash table that spills to global
// open-addressing tables.  No (word, 1) pair is ever written to HBM.
//
// Semantics implemented here: SURVEY.md Appendix A.2 (byte-parallel statement
// of the splitter) -- checked against oracle/ in tests/.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace wc {

typedef unsigned long long u64;
typedef unsigned int u32;

#ifndef WC_WARPS
#define WC_WARPS 28
#endif
#ifndef WC_CONSUMERS
#define WC_CONSUMERS 8
#endif
#ifndef WC_DRAIN
#define WC_DRAIN 64
#endif
// further build knobs (experiments): staging records per producer warp, ring cells, consumer sleep (ns)
#ifndef WC_STAGE
#define WC_STAGE 64
#endif
#ifndef WC_RING
#define WC_RING 1024
#endif
#ifndef WC_SLEEP0
#define WC_SLEEP0 400
#endif
#ifndef WC_SLEEP1
#define WC_SLEEP1 1200
#endif
constexpr int kSub = 512;                  // bytes per sub-tile: 32 lanes x 16 B
constexpr int kTile = 1024;                // bytes per warp-tile: two sub-tiles
constexpr int kWarps = WC_WARPS;           // warps per CTA
constexpr int kConsumers = WC_CONSUMERS;   // of which only insert deferred words into the global tables
constexpr int kProducers = kWarps - kConsumers; // the others tokenize and count; the text is split over these
// The split is a launch parameter (Launch::producers, Launch::smem_entries): kProducers / kConsumers are the
// default one (config 2: 17 % of the words go to the consumers), kProducersHeavy the one for text that sends
// more of its words to the global tables (config 3: 18 + 10 warps runs 13 % faster there, 5 % slower on config 2).
constexpr int kProducersHeavy = kProducers - 2;
constexpr int kThreads = kWarps * 32;
constexpr int kBufBytes = 1072;            // one tile buffer: 16 B tail of the previous tile + tile + slack
constexpr int kRingBytes = 2 * kBufBytes;  // per producer warp: two buffers, the next tile lands by cp.async
constexpr int kDlCap = 512;                // per producer warp: delimiter list entries
constexpr int kDlBytes = kDlCap * 2;
constexpr int kDrainBatch = WC_DRAIN;      // deferred words a consumer warp takes at a time (WC_DRAIN / 32 per lane)
constexpr u32 kDefRing = WC_RING;             // per CTA: ring of deferred-word records (16 bytes each)
constexpr u64 kNone = ~0ull;
constexpr u32 kMaxWordProbes = 128;        // longer probe chains in the global word table mean: grow it

#ifndef WC_EXP_UNROLL
#define WC_EXP_UNROLL 1
#endif
constexpr int kRoundUnroll = WC_EXP_UNROLL;      // word rounds unrolled by this much (experiment)
constexpr u32 kStageCap = WC_STAGE;              // per producer warp: deferred records gathered before they go to the ring
constexpr int kStageBytes = kStageCap * 16;
constexpr size_t kWarpSmemBytes = kRingBytes + kDlBytes + kStageBytes;
constexpr size_t kDefRingBytes = (size_t)kDefRing * 16;
constexpr u32 smem_entries_for(int producers) { // the per-CTA table gets what the producers' areas leave
    return (u32)((232448 - 256 - kDefRingBytes - (size_t)producers * kWarpSmemBytes) / 16) & ~31u;
}
constexpr u32 kSmemEntries = smem_entries_for(kProducers);
constexpr size_t smem_total_for(int producers) {
    return (size_t)smem_entries_for(producers) * 16 + kDefRingBytes + (size_t)producers * kWarpSmemBytes;
}
// Splits a launch may ask for (the engine uses kProducers and kProducersHeavy; WC_PRODUCERS can force any of them
// for an experiment).  One dynamic shared-memory size serves them all: the table takes what the producers leave.
constexpr int kMinProducers = kWarps / 2, kMaxProducers = kWarps - 4;
constexpr size_t smem_total_max() {
    size_t m = 0;
    for (int p = kMinProducers; p <= kMaxProducers; p++) m = smem_total_for(p) > m ? smem_total_for(p) : m;
    return m;
}
constexpr size_t kSmemTotal = smem_total_max();
static_assert(kProducers <= kMaxProducers && kProducersHeavy >= kMinProducers, "the engine's splits must be launchable");
static_assert(kWarpSmemBytes % 16 == 0, "per-warp shared memory must keep 16-byte alignment");
static_assert(kSmemTotal <= 232448 - 64, "more shared memory than an sm_100 CTA can have");
static_assert(kConsumers >= 1 && kProducersHeavy >= 1 && kDrainBatch % 32 == 0 && kDrainBatch <= 128, "bad build knobs");
static_assert((kDefRing & (kDefRing - 1)) == 0 && kDefRing >= 4 * kDrainBatch, "ring size");

enum : u32 { kErrWordFull = 1u, kErrLongFull = 2u, kErrArenaFull = 4u, kErrWordTooLong = 8u, kErrRingTimeout = 16u,
             kErrEmitFull = 32u, kErrPeer = 64u };

// Device-resident job state (one per engine).
struct Status {
    u64 n_flush;      // words emitted, empty ones included (two's complement adds for roll-back)
    u64 n_empty;
    u64 stop;         // min offset at which the reference stops reading (main.cu:185); kNone if none
    u64 w_entries;    // claimed slots, word table (1..16 bytes)
    u64 x_entries;    // claimed slots, long table (17+ bytes)
    u64 n_long;       // words of 9+ bytes (they bypass the shared-memory table)
    u64 n_spill;      // words of 1..8 bytes the shared-memory table had no room for
    u64 n_slow_tiles; // tiles that took the exact line-state path (control bytes)
    u64 arena_used;   // 8-byte units handed out (unit 0 is never used)
    u64 p0;           // offset that flushed the first non-empty word
    u64 lost_empty;   // empty words before p0 (main.cu:81,103: lost)
    u64 ovf_count;    // records appended to the overflow log (may exceed its capacity: then err is set)
    u32 p0_found;
    u32 err;
    u64 job_bytes;    // wc_peer_pull_device: the counters above describe the whole job (job_set), of this many bytes
    u64 job_set;
    u64 own_words;    // with job_set: the non-empty words this engine counted itself (note_consumer_load wants its own shard's)
    u64 dbg[8];       // WC_PROFILE builds: consumer cycles (wait, claim, global), batches; producer ring-full cycles, total
};

struct Tables {
    // words of 1..16 bytes: the key is the word itself, zero padded to 16 (exact by value).  Three
    // parallel arrays on purpose: the key sectors are read-only once a word is in, so the many readers
    // of a frequent word's key never queue behind the atomics on its count (as one 32-byte slot per
    // word the same kernel ran 1.6x slower: every key read waited for the read-modify-writes on its sector).
    ulonglong2 *w_keys;
    u64 *w_counts, *w_finv; // finv = ~first_offset, so 0 == unset and atomicMax keeps the min
    u32 w_log2;
    // words of 17+ bytes: slot = (24-bit hash tag << 40) | arena unit index; bytes compared in the arena
    u64 *x_refs, *x_counts, *x_finv;
    u32 x_log2;
    u64 *arena;       // unit[i] = length, unit[i+1..] = bytes, zero padded to 8
    u64 arena_units;
    // Words the word table could not take (too full) are not lost: they go, as {key0, key1, count,
    // first}, to this log, and the host replays it after re-hashing the table into a larger one.
    u64 *ovf_log;     // 4 words per record
    u64 ovf_cap;      // records
    Status *st;
};

struct WarpSummary { // per warp, in input order: what finding P0 needs
    u64 p0;            // offset that flushed this warp's first non-empty word, kNone if none
    u64 empties_before; // empty words this warp emitted before that
    u64 empties_total;
};

// Sort path (WC_PATH_SORT): instead of counting in hash tables, every word of 1..16 bytes is written
// out as a record -- 64-bit hash (the sort key), the word itself (16 bytes, zero padded) and the offset
// that flushed it -- to be radix-sorted and run-length reduced at the end of the job.
struct EmitBuf {
    u64 *cursor;       // records written so far (device counter); nullptr: hash path
    u64 *hash;         // [cap] sort keys, filled in by wc_aggregate
    u32 *idx;          // [cap] record index (the sort's payload), filled in by wc_aggregate
    ulonglong2 *key;   // [cap] the word, zero padded to 16 bytes
    u64 *end;          // [cap] absolute offset of the delimiter that flushed it
    u64 cap;
};

struct Launch {
    const uint8_t *buf; // 16-byte aligned, starts at a line start
    u64 n;
    u64 base;           // offset of buf[0] in the whole input
    u32 tiles_per_warp;
    u32 producers;      // count_kernel: tokenizing warps per CTA (kProducers or kProducersHeavy); the rest are consumers
    u32 smem_entries;   // count_kernel: entries of the CTA's shared-memory table = smem_entries_for(producers)
    int last_piece;     // the input ends with this piece
    int negate;         // roll-back launch: subtract everything at or after Status::stop
    int strict;         // read lines in pieces of kLineCap bytes like the reference's fgets does (main.cu:175,179)
    WarpSummary *ws;
    u32 n_ws;
    EmitBuf em;
};

// ---------------------------------------------------------------- helpers

__device__ __forceinline__ u32 lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ u64 mix64(u64 x) {
    x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    return x;
}

// Slot hash of the CTA's shared-memory table: multiplicative, used through __umulhi (its high bits
// pick the slot).  Two multiply-adds; on the config-2 vocabulary it misses 0.8 % more often than a
// full avalanche mix would, which costs less than the mix did.
__host__ __device__ __forceinline__ u32 hash32(u32 lo, u32 hi) { return lo * 0x9E3779B1u + hi * 0x85EBCA77u; }
// The same with a first multiplier that differs from CTA to CTA: two frequent words that share a slot of
// the CTA's table (one of them then takes the slow path every time) do so in one CTA, not in all of them.
__device__ __forceinline__ u32 hash32_of_cta(u32 lo, u32 hi) {
#ifdef WC1_CTA_HASH
    return lo * (0x9E3779B1u + 2u * (blockIdx.x * 0x632BE5ABu)) + hi * 0x85EBCA77u;
#else
    return hash32(lo, hi);
#endif
}

// Flags (0x80 per byte) -> 16-bit mask, bit p = byte p of the 16 bytes (w0 lowest).
// Two words share one multiply: (f1 | f0>>4) * (2^21+2^14+2^7+1) gathers the 8 flag bits
// into the product's top byte without any carry (all partial-product bits are distinct).
__host__ __device__ __forceinline__ u32 movemask16(u32 f0, u32 f1, u32 f2, u32 f3) {
    u32 p0 = (f1 | (f0 >> 4)) * 0x00204081u;
    u32 p1 = (f3 | (f2 >> 4)) * 0x00204081u;
    return (p0 >> 24) | ((p1 >> 16) & 0xFF00u);
}

struct ByteFlags { u32 delim, lf, kill; }; // 0x80 in each matching byte

// delim: 0x20 / 0x0A / 0x0D (main.cu:188); lf: 0x0A; kill: 0x0D or 0x00 (main.cu:187,195)
__host__ __device__ __forceinline__ ByteFlags classify4(u32 x) {
    u32 y = x & 0x7F7F7F7Fu;
    u32 a = (y ^ 0x20202020u) + 0x7F7F7F7Fu; // bit 7 clear <=> low 7 bits == 0x20
    u32 b = (y ^ 0x0A0A0A0Au) + 0x7F7F7F7Fu;
    u32 c = (y ^ 0x0D0D0D0Du) + 0x7F7F7F7Fu;
    u32 z = y + 0x7F7F7F7Fu;                 // bit 7 clear <=> low 7 bits == 0
    ByteFlags f;
    f.delim = ~((a & b & c) | x) & 0x80808080u;
    f.lf = ~(b | x) & 0x80808080u;
    f.kill = ~((c & z) | x) & 0x80808080u;
    return f;
}

__device__ __forceinline__ u32 warp_incl_scan(u32 v) {
    const u32 lane = lane_id();
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        u32 t = __shfl_up_sync(0xFFFFFFFFu, v, d);
        if (lane >= (u32)d) v += t;
    }
    return v;
}

// `nbytes` (1..8) bytes starting at byte offset `off` of an 8-byte-aligned global buffer, zero
// padded.  Only aligned 8-byte units that hold a requested byte are read.
__device__ __forceinline__ u64 load8_unaligned(const uint8_t *buf, u64 off, u32 nbytes = 8) {
    const u64 *p = reinterpret_cast<const u64 *>(buf + (off & ~7ull));
    const u32 o = (u32)(off & 7);
    u64 v = __ldg(p) >> (o * 8);
    if (o + nbytes > 8) v |= __ldg(p + 1) << (64 - o * 8);
    if (nbytes < 8) v &= (1ull << (8 * nbytes)) - 1;
    return v;
}

// The same in two steps, so that several reads can be in flight before the first is used.
struct Raw8 { u64 a, b; u32 o, nbytes; };
__device__ __forceinline__ Raw8 load8_issue(const uint8_t *buf, u64 off, u32 nbytes) {
    const u64 *p = reinterpret_cast<const u64 *>(buf + (off & ~7ull));
    Raw8 r;
    r.o = (u32)(off & 7); r.nbytes = nbytes;
    r.a = __ldg(p);
    r.b = (r.o + nbytes > 8) ? __ldg(p + 1) : 0ull;
    return r;
}
__device__ __forceinline__ u64 load8_finish(const Raw8 &r) {
    u64 v = r.a >> (r.o * 8);
    if (r.o) v |= r.b << (64 - r.o * 8);
    if (r.nbytes < 8) v &= (1ull << (8 * r.nbytes)) - 1;
    return v;
}

__device__ __forceinline__ u64 warp_sum64(u64 v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
    return v;
}

// ---- atomics with the address space spelled out.  Inside non-inlined device functions the compiler
// only sees generic pointers: it then emits returning generic atomics (a memory round trip each) and
// a run-time dispatch for shared memory.  These are what the code means.
__device__ __forceinline__ void red_add_u64(u64 *p, u64 v) { // global, fire and forget
    asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_max_u64(u64 *p, u64 v) {
    asm volatile("red.global.max.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_or_u32(u32 *p, u32 v) {
    asm volatile("red.global.or.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ u64 atom_add_u64(u64 *p, u64 v) { // global, returns the old value
    u64 old;
    asm volatile("atom.global.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
    return old;
}
__device__ __forceinline__ u64 atom_cas_u64(u64 *p, u64 cmp, u64 val) {
    u64 old;
    asm volatile("atom.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
    return old;
}
__device__ __forceinline__ u32 smem_addr(const void *p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ u64 atoms_cas_u64(void *p, u64 cmp, u64 val) { // shared
    u64 old;
    asm volatile("atom.shared.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "r"(smem_addr(p)), "l"(cmp), "l"(val) : "memory");
    return old;
}
__device__ __forceinline__ void reds_add_u32(void *p, u32 v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(smem_addr(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void reds_min_u32(void *p, u32 v) {
    asm volatile("red.shared.min.u32 [%0], %1;" ::"r"(smem_addr(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_u32(void *p, u32 v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(smem_addr(p)), "r"(v) : "memory");
}
__device__ __forceinline__ u32 lds_volatile_u32(const void *p) {
    u32 v;
    asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_addr(p)) : "memory");
    return v;
}

// ------------------------------------------------------- global word table (1..16 bytes)

__host__ __device__ __forceinline__ u32 hash_key128(u32 a, u32 b, u32 c, u32 d) {
    u32 h = a * 0x9E3779B1u;
    h ^= b * 0x85EBCA77u;
    h ^= c * 0xC2B2AE3Du;
    h ^= d * 0x27D4EB2Fu;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 13;
    h *= 0x297A2D39u;
    h ^= h >> 16;
    return h;
}

__device__ __forceinline__ ulonglong2 cas128(ulonglong2 *addr, ulonglong2 cmp, ulonglong2 val) {
    ulonglong2 old;
    asm volatile("{\n\t.reg .b128 c, v, o;\n\t"
                 "mov.b128 c, {%3, %4};\n\t"
                 "mov.b128 v, {%5, %6};\n\t"
                 "atom.global.cas.b128 o, [%2], c, v;\n\t"
                 "mov.b128 {%0, %1}, o;\n\t}"
                 : "=l"(old.x), "=l"(old.y)
                 : "l"(addr), "l"(cmp.x), "l"(cmp.y), "l"(val.x), "l"(val.y)
                 : "memory");
    return old;
}

// The global word table can be probed a BUCKET of kBucket slots at a time (all keys of the bucket are
// read at once).  The shipping build uses 1: the table is kept below half full, so a second probe
// is rare, and a 4-slot bucket measured 40 % slower (its 16 key registers per lookup went through
// the stack).  WC_BUCKET is a build knob for experiments.
#ifndef WC_BUCKET
#define WC_BUCKET 1
#endif
constexpr u32 kBucket = WC_BUCKET; // 1, 2 or 4
constexpr u32 kBucketLog2 = kBucket == 4 ? 2 : kBucket == 2 ? 1 : 0;
static_assert(kBucket == 1u << kBucketLog2, "bucket size");
struct Bucket { ulonglong2 k[kBucket]; };

// Home bucket (its first slot) of a word of 1..16 bytes: a function of its first 8 bytes and its
// length only, so that it is known before bytes 8..15 have been fetched.
__device__ __forceinline__ u32 word_home(const Tables &T, u64 k0, u32 len) {
    u32 h = (u32)k0 * 0x9E3779B1u + (u32)(k0 >> 32) * 0x85EBCA77u + len * 0xC2B2AE3Du;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du; // (the high bits pick the bucket)
    return (h >> (32 + kBucketLog2 - T.w_log2)) * kBucket;
}

__device__ __forceinline__ Bucket load_bucket(const Tables &T, u32 slot) {
    Bucket b;
#pragma unroll
    for (u32 i = 0; i < kBucket; i++) b.k[i] = __ldcg(&T.w_keys[slot + i]);
    return b;
}

__device__ __forceinline__ u32 key_len8(u64 k0) { return 8u - ((u32)__clzll(k0) >> 3); } // bytes used, k0 != 0

// Adds `delta` occurrences of a word of 1..16 bytes (k0 = bytes 0..7, k1 = bytes 8..15, zero padded)
// to the global table, starting at the bucket at `slot` whose keys the caller has already loaded.
// Returns 1 if it claimed a new slot.
__device__ __forceinline__ u32 word_insert_at(const Tables &T, u64 k0, u64 k1, u64 delta, u64 first_abs, u32 slot,
                                              Bucket b, u64 finv_seen = 0) {
    // finv_seen: the home slot's ~first as the caller read it (0: not read).  It only ever grows, so if it
    // already covers this offset the update below is not needed (one atomic less on the word's sector).
    const u32 mask = (u32)((1ull << T.w_log2) - 1);
    const u32 slot_home = slot;
    u32 is_new = 0;
    for (u32 probes = 0;; ++probes) {
        u32 found = kBucket;
#pragma unroll
        for (u32 i = 0; i < kBucket; i++)
            if (found == kBucket && b.k[i].x == k0 && b.k[i].y == k1) found = i;
        if (found == kBucket) {
#pragma unroll
            for (u32 i = 0; i < kBucket; i++) {
                if (found != kBucket || b.k[i].x != 0) continue; // a used slot never has k.x == 0
                const ulonglong2 old = cas128(&T.w_keys[slot + i], make_ulonglong2(0, 0), make_ulonglong2(k0, k1));
                if (old.x == 0) { is_new = 1; found = i; }
                else if (old.x == k0 && old.y == k1) found = i;
                else b.k[i] = old; // somebody else's word got there first
            }
        }
        if (found != kBucket) { slot += found; break; }
        slot = (slot + kBucket) & mask;
        if (probes >= kMaxWordProbes / kBucket || probes * kBucket >= mask) { // too full here: defer to the host
            const u64 at = atom_add_u64(&T.st->ovf_count, 1ull);
            if (at < T.ovf_cap) {
                T.ovf_log[4 * at] = k0; T.ovf_log[4 * at + 1] = k1;
                T.ovf_log[4 * at + 2] = delta; T.ovf_log[4 * at + 3] = first_abs;
            } else red_or_u32(&T.st->err, kErrWordFull);
            return 0;
        }
        b = load_bucket(T, slot);
    }
    red_add_u64(&T.w_counts[slot], delta); // both fire-and-forget
    if (slot != slot_home || ~first_abs > finv_seen) red_max_u64(&T.w_finv[slot], ~first_abs);
    return is_new;
}

__device__ __forceinline__ u32 word_insert(const Tables &T, u64 k0, u64 k1, u32 len, u64 delta, u64 first_abs) {
    const u32 slot = word_home(T, k0, len);
    return word_insert_at(T, k0, k1, delta, first_abs, slot, load_bucket(T, slot));
}

// -------------------------------------------------------- mappers
// The reference is nominally a MapReduce skeleton: its mapper() (main.cu:37-54) turns every word into
// the pair (word, 1) and a maintainer may swap it.  Here the mapper is a policy of the kernel: a
// byte-wise, length-preserving transform of the word's bytes that is applied wherever they are read
// (key extraction from the tile, bytes 8..15 fetched by the consumers, hashing / comparing / storing
// the 17+ byte words).  It must map 0x00 to 0x00 and must not produce or consume delimiters.  Adding
// one: a struct with map4 / map8, an entry in wc.h's WC_MAP_* and in launch_count_kernel() (wc_engine.cu).
struct MapIdentity { // main.cu's own mapper: the word as it is
    __device__ __forceinline__ static u32 map4(u32 x) { return x; }
    __device__ __forceinline__ static u64 map8(u64 x) { return x; }
};
struct MapLowerAscii { // 'A'..'Z' -> 'a'..'z': a case-insensitive word count
    __device__ __forceinline__ static u32 map4(u32 x) {
        const u32 h = x & 0x7F7F7F7Fu;
        const u32 up = ~x & 0x80808080u & ((h + 0x3F3F3F3Fu) ^ (h + 0x25252525u)); // 0x80 where 'A' <= byte <= 'Z'
        return x | (up >> 2);
    }
    __device__ __forceinline__ static u64 map8(u64 x) {
        return (u64)map4((u32)x) | ((u64)map4((u32)(x >> 32)) << 32);
    }
};

// -------------------------------------------------------- global long table (17+ bytes)

// Very long words (the adversarial corpus has some of 1 MiB) are handled by one lane, so their loops
// keep 16 independent 8-byte loads in flight per step: they are bound by load latency, not by arithmetic.
template <class Map>
__device__ __forceinline__ u64 long_hash(const uint8_t *buf, u64 start, u64 len) {
    u64 h = 0x9E3779B97F4A7C15ull ^ (len * 0xFF51AFD7ED558CCDull);
    u64 i = 0;
    for (; i + 128 <= len; i += 128) {
        u64 v[16];
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = Map::map8(load8_unaligned(buf, start + i + 8 * j));
#pragma unroll
        for (int j = 0; j < 16; j++) { h = (h ^ v[j]) * 0x9FB21C651E98DF25ull; h ^= h >> 32; }
    }
    for (; i + 8 <= len; i += 8) {
        h = (h ^ Map::map8(load8_unaligned(buf, start + i))) * 0x9FB21C651E98DF25ull;
        h ^= h >> 32;
    }
    if (i < len) {
        u64 v = Map::map8(load8_unaligned(buf, start + i, (u32)(len - i)));
        h = (h ^ v) * 0x9FB21C651E98DF25ull;
        h ^= h >> 32;
    }
    return mix64(h);
}

template <class Map>
__device__ __forceinline__ bool long_equal(const Tables &T, u64 unit, const uint8_t *buf, u64 start, u64 len) {
    if (__ldcg(&T.arena[unit]) != len) return false;
    u64 i = 0;
    for (; i + 64 <= len; i += 64) {
        u64 a[8], b[8];
#pragma unroll
        for (int j = 0; j < 8; j++) { a[j] = __ldcg(&T.arena[unit + 1 + (i >> 3) + j]); b[j] = Map::map8(load8_unaligned(buf, start + i + 8 * j)); }
        u64 diff = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) diff |= a[j] ^ b[j];
        if (diff) return false;
    }
    for (; i + 8 <= len; i += 8)
        if (__ldcg(&T.arena[unit + 1 + (i >> 3)]) != Map::map8(load8_unaligned(buf, start + i))) return false;
    if (i < len) {
        u64 v = Map::map8(load8_unaligned(buf, start + i, (u32)(len - i)));
        if (__ldcg(&T.arena[unit + 1 + (i >> 3)]) != v) return false;
    }
    return true;
}

// A word of 17+ bytes at buf[start, start+len).  Never waits on another thread: a new word is
// first written to the arena, then published with one CAS of the slot word.  Returns 1 if new.
template <class Map>
__device__ __noinline__ u32 long_insert(const Tables *Tp, const uint8_t *buf, u64 start, u64 len, u64 delta,
                                        u64 first_abs) {
    const Tables &T = *Tp;
    const u64 h = long_hash<Map>(buf, start, len);
    const u64 mask = (1ull << T.x_log2) - 1;
    const u64 tag = h >> 40;
    u64 slot = h & mask;
    u64 my_ref = 0;
    u32 is_new = 0;
    for (u32 probes = 0;; ++probes) {
        u64 r = __ldcg(&T.x_refs[slot]);
        if (r == 0) {
            if (my_ref == 0) {
                u64 units = 1 + ((len + 7) >> 3);
                u64 idx = atom_add_u64(&T.st->arena_used, units);
                if (idx + units > T.arena_units) { red_or_u32(&T.st->err, kErrArenaFull); return 0; }
                T.arena[idx] = len;
                u64 i = 0;
                for (; i + 128 <= len; i += 128) {
                    u64 v[16];
#pragma unroll
                    for (int j = 0; j < 16; j++) v[j] = Map::map8(load8_unaligned(buf, start + i + 8 * j));
#pragma unroll
                    for (int j = 0; j < 16; j++) T.arena[idx + 1 + (i >> 3) + j] = v[j];
                }
                for (; i + 8 <= len; i += 8) T.arena[idx + 1 + (i >> 3)] = Map::map8(load8_unaligned(buf, start + i));
                if (i < len) T.arena[idx + 1 + (i >> 3)] = Map::map8(load8_unaligned(buf, start + i, (u32)(len - i)));
                __threadfence();
                my_ref = (tag << 40) | idx;
            }
            r = atom_cas_u64(&T.x_refs[slot], 0ull, my_ref);
            if (r == 0) { is_new = 1; break; }
        }
        if ((r >> 40) == tag && long_equal<Map>(T, r & ((1ull << 40) - 1), buf, start, len)) break;
        slot = (slot + 1) & mask;
        if (probes > 4096 || probes >= mask) { red_or_u32(&T.st->err, kErrLongFull); return 0; }
    }
    red_add_u64(&T.x_counts[slot], delta);
    red_max_u64(&T.x_finv[slot], ~first_abs);
    return is_new;
}

// ------------------------------------------------------------ main kernel

struct WarpCarry {
    long long last_delim; // buffer-relative position of the last delimiter byte seen, -1 if none
    bool killed;          // a 0x0D / 0x00 since the last 0x0A: rest of the line is not split (main.cu:187,196)
    bool last_is_lf;      // previous byte was 0x0A (or this is the start of the piece)
    long long line_start; // buffer-relative position of the first byte of the line that holds pos
};

// State at buffer offset `pos` (> 0), found by scanning backwards to the previous line start.
__device__ __forceinline__ WarpCarry warp_lookback(const uint8_t *buf, u64 pos) {
    const u32 lane = lane_id();
    WarpCarry c;
    c.last_delim = -1; c.killed = false; c.line_start = 0;
    c.last_is_lf = buf[pos - 1] == 0x0A;
    bool have_delim = false, seen_kill = false;
    for (long long hi = (long long)pos; hi > 0; hi -= 32) { // window [hi-32, hi), lane 0 nearest
        long long p = hi - 1 - (long long)lane;
        u32 ch = p >= 0 ? buf[p] : 0x78u;
        u32 bd = __ballot_sync(0xFFFFFFFFu, p >= 0 && (ch == 0x20 || ch == 0x0A || ch == 0x0D));
        u32 bl = __ballot_sync(0xFFFFFFFFu, p >= 0 && ch == 0x0A);
        u32 bk = __ballot_sync(0xFFFFFFFFu, p >= 0 && (ch == 0x0D || ch == 0x00));
        if (!have_delim && bd) { c.last_delim = hi - 1 - (long long)(__ffs(bd) - 1); have_delim = true; }
        if (bl) { // nearest line feed: kills between it and pos count
            u32 nearer = (1u << (__ffs(bl) - 1)) - 1;
            c.killed = seen_kill || (bk & nearer);
            c.line_start = hi - (long long)(__ffs(bl) - 1); // the byte after that line feed
            return c;
        }
        seen_kill = seen_kill || bk;
    }
    c.killed = seen_kill;
    return c;
}

// 16 bytes at buffer offset `off` (a multiple of 16); bytes at or past n read as 'x' (a word byte).
__device__ __forceinline__ uint4 load16(const uint8_t *buf, u64 n, u64 off) {
    if (off + 16 <= n) return __ldcs(reinterpret_cast<const uint4 *>(buf + off));
    u32 w[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        u32 x = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const u64 i = off + (u64)(j * 4 + b);
            x |= (i < n ? (u32)buf[i] : 0x78u) << (8 * b);
        }
        w[j] = x;
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// Fast classification of 4 bytes: cand = byte <= 0x20, lf = 0x0A, ex = a byte <= 0x20 that is
// neither 0x20 nor 0x0A (0x0D, 0x00, tab, other controls).  In text without such bytes cand is
// exactly the delimiter set of main.cu:188; a tile that has one takes the exact path below.
struct FastFlags { u32 cand, lf, ex; };
__host__ __device__ __forceinline__ FastFlags classify_fast4(u32 x) {
    const u32 y = x & 0x7F7F7F7Fu;
    const u32 c = y + 0x5F5F5F5Fu; // bit 7 set <=> low 7 bits >= 0x21
    FastFlags f;
    f.cand = ~(c | x) & 0x80808080u;
    const u32 b = (y ^ 0x0A0A0A0Au) + 0x7F7F7F7Fu;
    f.lf = ~(b | x) & 0x80808080u;
    f.ex = f.cand & ~(x << 2) & ~f.lf; // among bytes <= 0x20 only 0x20 has bit 5
    return f;
}

// Exact line state of one 512-byte sub-tile: a delimiter flushes only if no
// 0x0D / 0x00 precedes it on its line.  Taken by tiles that hold a control byte: CRLF files, NULs, tabs.
struct ExactOut { u32 dm, live, lfm, killed_out; };
__device__ __noinline__ ExactOut exact_sub(uint4 v, u32 killed_in, const uint8_t *buf, u64 n, u64 base,
                                           u64 sub_off, u64 *stop_ptr, int negate) {
    const u32 lane = lane_id();
    const ByteFlags f0 = classify4(v.x), f1 = classify4(v.y), f2 = classify4(v.z), f3 = classify4(v.w);
    ExactOut o;
    o.dm = movemask16(f0.delim, f1.delim, f2.delim, f3.delim);
    o.lfm = movemask16(f0.lf, f1.lf, f2.lf, f3.lf);
    const u32 km = movemask16(f0.kill, f1.kill, f2.kill, f3.kill);
    // per-lane transfer function of "killed": s -> reset ? val : (s | val)
    u32 reset = o.lfm != 0;
    u32 val = reset ? ((km >> (32 - __clz(o.lfm))) != 0) : (km != 0);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 r2 = __shfl_up_sync(0xFFFFFFFFu, reset, d);
        const u32 v2 = __shfl_up_sync(0xFFFFFFFFu, val, d);
        if (lane >= (u32)d) { // compose: earlier (r2, v2) then mine
            val = reset ? val : (val | v2);
            reset = reset | r2;
        }
    }
    const u32 r_in = __shfl_up_sync(0xFFFFFFFFu, reset, 1), v_in = __shfl_up_sync(0xFFFFFFFFu, val, 1);
    bool s = killed_in != 0;
    if (lane > 0) s = r_in ? (v_in != 0) : (s || v_in);
    const u32 r_all = __shfl_sync(0xFFFFFFFFu, reset, 31), v_all = __shfl_sync(0xFFFFFFFFu, val, 31);
    o.killed_out = r_all ? (v_all != 0) : (killed_in || v_all);
    o.live = 0;
#pragma unroll
    for (int p = 0; p < 16; p++) {
        const u32 bit = 1u << p;
        if ((o.dm & bit) && !s) o.live |= bit;
        if (km & bit) s = true;
        if (o.lfm & bit) s = false;
    }
    // A NUL at line offset 0 or 1 makes strlen() < 2: the reference stops (main.cu:185).
    if (!negate) {
        u32 nm = km & ~o.dm; // NUL bytes (0x0D is in dm)
        while (nm) {
            const u32 p = __ffs(nm) - 1; nm &= nm - 1;
            const u64 i = sub_off + lane * 16 + p;
            if (i >= n) break;
            const bool at0 = i == 0 || buf[i - 1] == 0x0A;
            const bool at1 = !at0 && (i == 1 || buf[i - 2] == 0x0A);
            if (at0) atomicMin(stop_ptr, base + i);
            else if (at1) atomicMin(stop_ptr, base + i - 1);
        }
    }
    return o;
}

// ---- strict-ref mode: the reference reads with fgets(szLine, 100) (main.cu:175,179),
// i.e. a line longer than 99 bytes arrives in PIECES of 99 bytes, and everything main.cu does per line
// it really does per piece: the pending word is dropped at a piece's end (so a word that straddles the
// edge loses its head), 0x0D / 0x00 hide only the rest of the piece, and a piece of one byte stops the
// input (main.cu:185).  A "piece start" is a line start or a byte at line offset 99, 198, ...
constexpr u32 kLineCap = 99;

// Any 0x0D / 0x00 in buf[lo, hi)?  (hi - lo < 128; whole warp)
__device__ __forceinline__ bool warp_any_kill(const uint8_t *buf, long long lo, long long hi) {
    bool any = false;
    for (long long p = lo + lane_id(); __any_sync(0xFFFFFFFFu, p < hi); p += 32) {
        const u32 ch = p < hi ? buf[p] : 0x78u;
        any = any || __any_sync(0xFFFFFFFFu, ch == 0x0D || ch == 0x00);
    }
    return any;
}

// Exact masks of one lane's 32 bytes when pieces matter.  starts: bit p = byte p starts a line or a piece;
// carry_start: the byte before this lane's first byte starts one.
struct StrictOut { u32 dm, live, lfm, killed_out; };
__device__ __noinline__ StrictOut strict_exact(uint4 va, uint4 vb, u32 starts, u32 carry_start, u32 killed_in, u64 abs_pos,
                                               u64 *stop_ptr, int negate) {
    const u32 lane = lane_id();
    StrictOut o;
    u32 km;
    {
        const ByteFlags a0 = classify4(va.x), a1 = classify4(va.y), a2 = classify4(va.z), a3 = classify4(va.w);
        const ByteFlags b0 = classify4(vb.x), b1 = classify4(vb.y), b2 = classify4(vb.z), b3 = classify4(vb.w);
        o.dm = movemask16(a0.delim, a1.delim, a2.delim, a3.delim) | (movemask16(b0.delim, b1.delim, b2.delim, b3.delim) << 16);
        o.lfm = movemask16(a0.lf, a1.lf, a2.lf, a3.lf) | (movemask16(b0.lf, b1.lf, b2.lf, b3.lf) << 16);
        km = movemask16(a0.kill, a1.kill, a2.kill, a3.kill) | (movemask16(b0.kill, b1.kill, b2.kill, b3.kill) << 16);
    }
    // per-lane transfer function of "killed": s -> reset ? val : (s | val); a piece start resets it
    u32 reset = starts != 0;
    u32 val = reset ? ((km >> (31 - __clz(starts))) != 0) : (km != 0);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 r2 = __shfl_up_sync(0xFFFFFFFFu, reset, d);
        const u32 v2 = __shfl_up_sync(0xFFFFFFFFu, val, d);
        if (lane >= (u32)d) { val = reset ? val : (val | v2); reset = reset | r2; }
    }
    const u32 r_in = __shfl_up_sync(0xFFFFFFFFu, reset, 1), v_in = __shfl_up_sync(0xFFFFFFFFu, val, 1);
    bool s = killed_in != 0;
    if (lane > 0) s = r_in ? (v_in != 0) : (s || v_in);
    const u32 r_all = __shfl_sync(0xFFFFFFFFu, reset, 31), v_all = __shfl_sync(0xFFFFFFFFu, val, 31);
    o.killed_out = r_all ? (v_all != 0) : (killed_in || v_all);
    o.live = 0;
#pragma unroll 4
    for (int p = 0; p < 32; p++) {
        const u32 bit = 1u << p;
        if (starts & bit) s = false;
        if ((o.dm & bit) && !s) o.live |= bit;
        if (km & bit) s = true;
    }
    // A NUL at offset 0 or 1 of a piece makes strlen() < 2: the reference stops at that piece (main.cu:185).
    if (!negate) {
        const u32 nm = km & ~o.dm;
        const u32 at0 = nm & starts, at1 = nm & ((starts << 1) | (carry_start & 1u)) & ~starts;
        if (at0) atomicMin(stop_ptr, abs_pos + (__ffs(at0) - 1));
        if (at1) atomicMin(stop_ptr, abs_pos + (__ffs(at1) - 1) - 1);
    }
    return o;
}

// Deferred words: words of 1..8 bytes that missed in the CTA's table, and every word of 9+ bytes.
// The warps that tokenize (producers) only append a 16-byte record (first 8 bytes, end offset
// relative to the CTA's range, length) to a ring in shared memory; the CTA's consumer warps take the
// records out and do the slow part, so that no tokenizing warp ever waits for global memory:
//   1..8 bytes  : claim the word's shared-memory slot if it is still free (this is how the CTA's
//                 table fills up), else count it in the global word table;
//   9..16 bytes : fetch bytes 8..15 from the text, count in the global word table;
//   17+ bytes   : the arena-backed long table.
// Ring protocol (a bounded multi-producer multi-consumer queue with a sequence number per cell):
// a producer warp gathers its records in a staging area of its own and reserves ring positions with
// one atomicAdd on `tail` per 33..64 of them (flush_stage); position p belongs to
// cell p % kDefRing in lap p / kDefRing.  A cell's last word says whose turn it is:
//     lap                       (bit 31 clear)  the cell is free and waits for the writer of that lap
//     1<<31 | lap<<16 | t<<15 | len15  (bit 31 set)  the record of that lap is in the cell
// so a writer that is a whole lap ahead waits for its turn instead of overwriting, and the record
// is published by the same single 16-byte store that carries it.  A consumer warp reserves
// kDrainBatch positions with an atomicAdd on `head`, waits until they are published (or until every
// producer has finished and the position is past `tail`), copies the records out and hands the cells
// to the next lap.  The oldest unconsumed position can always be written and read, so nobody waits
// in a circle; both sides carry a watchdog that raises an error flag instead of hanging.
// len15 saturates; records of 17+ byte words carry the full length in their first word.  t: the
// word's shared-memory slot already belongs to another word.
struct DefCtl { u32 tail, head, done, abort_flag; };

__device__ __forceinline__ u32 ring_lap(u32 pos) { return (pos / kDefRing) & 0x7FFFu; }
// bit 15: the word's shared-memory slot was seen occupied by another word (no point in trying to claim it)
__device__ __forceinline__ u32 ring_full_word(u32 pos, u32 len, bool slot_taken) {
    return 0x80000000u | (ring_lap(pos) << 16) | (slot_taken ? 0x8000u : 0u) | (len > 0x7FFFu ? 0x7FFFu : len);
}

__device__ __forceinline__ u32 atoms_add(u32 *p, u32 v) { // one lane's shared-memory atomicAdd, as written
    u32 old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(smem_addr(p)), "r"(v) : "memory");
    return old;
}

__device__ __forceinline__ uint4 lds128_volatile(const void *p) { // one 16-byte shared-memory read, never cached in registers
    uint4 v;
    asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_addr(p)) : "memory");
    return v;
}

// Moves a producer warp's n staged records (n <= kStageCap) to the CTA's ring: one reservation for
// all of them, then each lane waits for its cell's turn (normally free already) and writes the
// record with the position's lap number in its last word.
__device__ __forceinline__ void flush_stage(const uint4 *stage, u32 n, uint4 *ring, DefCtl *ctl, Status *st) {
    const u32 lane = lane_id();
    u32 pos0 = 0;
    if (lane == 0) pos0 = atoms_add(&ctl->tail, n);
    pos0 = __shfl_sync(0xFFFFFFFFu, pos0, 0);
#ifdef WC_PROFILE
    const long long pw0 = clock64();
    bool waited = false;
#endif
    for (u32 i0 = 0; i0 < n; i0 += 32) {
        const u32 i = i0 + lane;
        const bool mine = i < n;
        const u32 pos = pos0 + i, cell = pos & (kDefRing - 1);
        uint4 r = make_uint4(0, 0, 0, 0);
        if (mine) r = stage[i];
        u32 spins = 0;
        while (__any_sync(0xFFFFFFFFu, mine && lds_volatile_u32(&ring[cell].w) != ring_lap(pos))) {
            if (__any_sync(0xFFFFFFFFu, ++spins > (1u << 25) || lds_volatile_u32(&ctl->abort_flag) != 0)) {
                sts_u32(&ctl->abort_flag, 1); atomicOr(&st->err, kErrRingTimeout); break;
            }
#ifdef WC_PROFILE
            waited = true;
#endif
            __nanosleep(64); // the ring is full: leave the issue slots to the consumers
        }
        { // staged w = length | slot_taken << 31  ->  the ring's format (a 17+ byte word carries its length in x)
            const u32 len = r.w & 0x7FFFFFFFu;
            if (len > 16) r.x = len;
            r.w = ring_full_word(pos, len, (r.w >> 31) != 0);
        }
#ifndef WC_EXP_NODRAIN
        if (mine) ring[cell] = r;
#endif
    }
#ifdef WC_PROFILE
    if (lane == 0 && waited) atomicAdd(&st->dbg[4], (u64)(clock64() - pw0));
#endif
    __syncwarp(); // the staged records have been read: the area may be overwritten
}

template <class Map>
#ifdef WC_EXP_NOINLINE
__device__ __noinline__
#else
__device__ __forceinline__
#endif
void consume_deferred(uint4 *ring, DefCtl *ctl, const Launch *Lp, const Tables *Tp,
                                              u64 cta_base, uint4 *tab) {
    const u32 lane = lane_id();
    const Tables &T = *Tp;
    const u64 delta = Lp->negate ? ~0ull : 1ull;
    const u32 n_producers = Lp->producers, n_entries = Lp->smem_entries;
    constexpr int R = kDrainBatch / 32;
    u32 n_wnew = 0, n_xnew = 0, n_spill = 0, n_long = 0;
#ifdef WC_PROFILE
    long long pc_wait = 0, pc_claim = 0, pc_glob = 0, pc_batches = 0, pc_t = clock64();
#endif
    for (;;) {
        u32 b = 0;
        if (lane == 0) b = atoms_add(&ctl->head, (u32)kDrainBatch);
        b = __shfl_sync(0xFFFFFFFFu, b, 0);
        uint4 rec[R];
        u32 live = 0; // bit j: record j was taken
        u32 spins = 0;
        bool finished = false;
        // Cells fill roughly in order: watch the batch's last cell only (one broadcast read per poll) and
        // look at all of them once that one is in, or once every producer has finished.
#ifndef WC_EXP_NOPREWAIT
        __syncwarp();
        for (const u32 last = b + kDrainBatch - 1;;) {
            if ((lds_volatile_u32(&ring[last & (kDefRing - 1)].w) >> 16) == (0x8000u | ring_lap(last))) break;
            if (lds_volatile_u32(&ctl->done) == n_producers || ++spins > (1u << 23)) break;
            __nanosleep(spins < 2 ? WC_SLEEP0 : WC_SLEEP1); // a batch takes a few microseconds to fill: do not burn issue slots polling
        }
#endif
        u32 polls = spins > (1u << 23) ? spins : 0; // (counted per pass of this loop, which the vote below keeps in step)
        polls = __shfl_sync(0xFFFFFFFFu, polls, 0);
        for (;;) { // take the batch
            const bool all_done = lds_volatile_u32(&ctl->done) == n_producers; // read before tail: tail is final once this holds
            const u32 tail = lds_volatile_u32(&ctl->tail);
            bool ok = true;
#pragma unroll
            for (int j = 0; j < R; j++) {
                if (live & (1u << j)) continue;
                const u32 pos = b + lane + 32 * j;
                const uint4 r = lds128_volatile(&ring[pos & (kDefRing - 1)]);
                if ((r.w >> 16) == (0x8000u | ring_lap(pos))) { // this position's record is in
                    rec[j] = r; live |= 1u << j;
                    sts_u32(&ring[pos & (kDefRing - 1)].w, (ring_lap(pos) + 1) & 0x7FFFu); // the cell goes to the next lap
                } else if (!(all_done && (int)(tail - pos) <= 0)) ok = false; // reserved but not yet written
            }
            if (__all_sync(0xFFFFFFFFu, ok)) {
                // lane 0's view decides for the warp: lanes that read `done` at different moments must
                // not disagree about leaving the loop (the next __shfl_sync would wait for ever)
                finished = __shfl_sync(0xFFFFFFFFu, (u32)(all_done && (int)(tail - b) <= 0), 0) != 0;
                break;
            }
            if (++polls > (1u << 23)) { sts_u32(&ctl->abort_flag, 1); red_or_u32(&T.st->err, kErrRingTimeout); break; }
            __nanosleep(200);
        }
#ifdef WC_PROFILE
        { long long t = clock64(); pc_wait += t - pc_t; pc_t = t; pc_batches++; }
#endif
        // Decode, then issue every global read of the batch (the home bucket of each 1..16 byte word
        // and bytes 8..15 of the longer ones) BEFORE touching the shared-memory table: the claims below
        // run while those reads are in flight.  A word whose claim succeeds wasted one read.
        u32 slot[R]; Bucket key[R]; Raw8 raw[R]; u64 fin[R];
        u32 todo = 0; // bit j: record j goes to the global word table
#pragma unroll
        for (int j = 0; j < R; j++) {
            slot[j] = 0; raw[j] = Raw8{0, 0, 0, 8}; fin[j] = 0;
            if (!(live & (1u << j))) { rec[j] = make_uint4(0, 0, 0, 0); continue; }
            const bool slot_taken = (rec[j].w & 0x8000u) != 0;
            rec[j].w &= 0x7FFFu;                 // the word's length, saturated ...
            if (rec[j].w > 16) rec[j].w = rec[j].x; // ... a 17+ byte word carries its full length instead of its first bytes
            const u32 len = rec[j].w;
            n_long += len > 8;
            if (len <= 16) {
                const u64 k0 = (u64)rec[j].x | ((u64)rec[j].y << 32);
                slot[j] = word_home(T, k0, len);
                key[j] = load_bucket(T, slot[j]);
#ifdef WC1_FINV_HINT
                fin[j] = __ldcg(&T.w_finv[slot[j]]);
#endif
                if (len > 8) raw[j] = load8_issue(Lp->buf, cta_base + rec[j].z - len + 8, len - 8);
                if (len > 8 || slot_taken) todo |= 1u << j;
            }
        }
#pragma unroll
        for (int j = 0; j < R; j++) { // the slot looked free: claim it (this is how the CTA's table fills up)
            const u32 len = rec[j].w;
            if (len == 0 || len > 8) continue;
            if (todo & (1u << j)) { n_spill++; continue; }
            const u64 k0 = (u64)rec[j].x | ((u64)rec[j].y << 32);
            const u32 sidx = __umulhi(hash32_of_cta(rec[j].x, rec[j].y), n_entries);
            const u64 old = atoms_cas_u64(&tab[sidx], 0ull, k0);
            if (old == 0 || old == k0) {
                u32 *cell = reinterpret_cast<u32 *>(&tab[sidx]);
                reds_add_u32(cell + 2, (u32)delta);
                reds_min_u32(cell + 3, rec[j].z);
            } else { todo |= 1u << j; n_spill++; }
        }
#ifdef WC_PROFILE
        { long long t = clock64(); pc_claim += t - pc_t; pc_t = t; }
#endif
#pragma unroll
        for (int j = 0; j < R; j++) {
            const u64 k0 = (u64)rec[j].x | ((u64)rec[j].y << 32);
            const u64 end = cta_base + rec[j].z; // buffer-relative offset of the delimiter
            if (todo & (1u << j))
                n_wnew += word_insert_at(T, k0, rec[j].w > 8 ? Map::map8(load8_finish(raw[j])) : 0ull, delta, Lp->base + end, slot[j], key[j], fin[j]);
        }
        // 17+ byte words last and on their own: a function call between the loads above and their use
        // would push every loaded key through the stack.
        u32 has_long = 0;
#pragma unroll
        for (int j = 0; j < R; j++) has_long |= rec[j].w > 16 ? 1u << j : 0u;
        if (__any_sync(0xFFFFFFFFu, has_long != 0)) {
#pragma unroll 1
            for (int j = 0; j < R; j++) {
                uint4 r = rec[0];
#pragma unroll
                for (int q = 1; q < R; q++) if (q == j) r = rec[q];
                const u64 end = cta_base + r.z;
                if (r.w > 16) n_xnew += long_insert<Map>(Tp, Lp->buf, end - r.w, r.w, delta, Lp->base + end);
            }
        }
#ifdef WC_PROFILE
        { long long t = clock64(); pc_glob += t - pc_t; pc_t = t; }
#endif
        if (__any_sync(0xFFFFFFFFu, finished || lds_volatile_u32(&ctl->abort_flag) != 0)) break;
    }
#ifdef WC_PROFILE
    if (lane == 0) {
        atomicAdd(&T.st->dbg[0], (u64)pc_wait); atomicAdd(&T.st->dbg[1], (u64)pc_claim);
        atomicAdd(&T.st->dbg[2], (u64)pc_glob); atomicAdd(&T.st->dbg[3], (u64)pc_batches);
    }
#endif
    n_wnew = __reduce_add_sync(0xFFFFFFFFu, n_wnew);
    n_xnew = __reduce_add_sync(0xFFFFFFFFu, n_xnew);
    n_spill = __reduce_add_sync(0xFFFFFFFFu, n_spill);
    n_long = __reduce_add_sync(0xFFFFFFFFu, n_long);
    if (lane == 0) {
        if (n_long && !Lp->negate) red_add_u64(&T.st->n_long, (u64)n_long);
        if (n_wnew) red_add_u64(&T.st->w_entries, (u64)n_wnew);
        if (n_xnew) red_add_u64(&T.st->x_entries, (u64)n_xnew);
        if (n_spill && !Lp->negate) red_add_u64(&T.st->n_spill, (u64)n_spill);
    }
}


// a16 / b16: this lane's 16-bit masks of bytes [16*lane, +16) of a tile's first and second 512 bytes.
// Returns the 32-bit mask of bytes [32*lane, +32) of the tile.
__device__ __forceinline__ u32 halves_to_lanes(u32 a16, u32 b16) {
    const u32 lane = lane_id();
    const u32 both = (a16 & 0xFFFFu) | (b16 << 16);
    const u32 x0 = __shfl_sync(0xFFFFFFFFu, both, (2 * lane) & 31), x1 = __shfl_sync(0xFFFFFFFFu, both, (2 * lane + 1) & 31);
    const u32 sh = lane < 16 ? 0u : 16u;
    return ((x0 >> sh) & 0xFFFFu) | (((x1 >> sh) & 0xFFFFu) << 16);
}

// One lane's two 16-byte pieces of tile `tile` into its buffer: asynchronous copies when the
// bytes are all inside the input, a padded synchronous load at the very end of it.
__device__ __forceinline__ void fetch_tile(const Launch &L, u64 tile, unsigned char *buf, u32 lane) {
    const u64 offA = tile * kTile + lane * 16, offB = offA + kSub;
    unsigned char *dA = buf + 16 + lane * 16, *dB = dA + kSub;
    if (offB + 16 <= L.n) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(dA)), "l"(L.buf + offA) : "memory");
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(dB)), "l"(L.buf + offB) : "memory");
    } else {
        *reinterpret_cast<uint4 *>(dA) = load16(L.buf, L.n, offA);
        *reinterpret_cast<uint4 *>(dB) = load16(L.buf, L.n, offB);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

extern __shared__ __align__(16) unsigned char wc_smem[];

template <class Map>
__global__ void __launch_bounds__(kThreads, 1)
count_kernel(const __grid_constant__ Launch L, const __grid_constant__ Tables T) {
    uint4 *tab = reinterpret_cast<uint4 *>(wc_smem);
    const u32 n_producers = L.producers, n_entries = L.smem_entries;
    uint4 *def_ring = reinterpret_cast<uint4 *>(wc_smem + (size_t)n_entries * 16);
    unsigned char *warp_all = wc_smem + (size_t)n_entries * 16 + kDefRingBytes;
    __shared__ u32 s_first_ws;
    __shared__ u64 s_stop;
    __shared__ DefCtl s_ctl;

    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    Status *st = T.st;

    // ---- which part of the piece does this launch cover?  (One thread reads STOP so that the
    // whole CTA takes the same branch even while other CTAs are still lowering it.)
    if (tid == 0) s_stop = *reinterpret_cast<volatile u64 *>(&st->stop);
    __syncthreads();
    const u64 stop_seen = s_stop;
    u64 range_lo = 0; // only flush points at buffer offsets >= range_lo are counted
    if (L.negate) {
        // Roll-back launch.  CTA 0 first folds the forward launch's per-warp summaries into the
        // job-wide P0 / lost-empties state; then every CTA subtracts what lies at or after STOP.
        if (blockIdx.x == 0 && L.ws != nullptr && st->p0_found == 0) {
            if (tid == 0) s_first_ws = 0xFFFFFFFFu;
            __syncthreads();
            for (u32 i = tid; i < L.n_ws; i += kThreads)
                if (L.ws[i].p0 != kNone) atomicMin(&s_first_ws, i);
            __syncthreads();
            const u32 first = s_first_ws;
            u64 lost = 0;
            for (u32 i = tid; i < L.n_ws && i < first; i += kThreads) lost += L.ws[i].empties_total;
            if (first != 0xFFFFFFFFu && tid == 0) lost += L.ws[first].empties_before;
            if (lost) atomicAdd(&st->lost_empty, lost);
            __syncthreads();
            if (tid == 0 && first != 0xFFFFFFFFu) { st->p0 = L.ws[first].p0; st->p0_found = 1; }
        }
        // Nothing to undo unless STOP lies inside this piece (a piece after STOP was never counted).
        const u64 stop = stop_seen;
        if (stop == kNone || stop >= L.base + L.n || stop < L.base) return;
        range_lo = stop - L.base;
    } else {
        const u64 stop = stop_seen;
        if (stop != kNone && stop < L.base) { // the reference stopped reading before this piece
            if (L.ws != nullptr)
                for (u32 i = blockIdx.x * kWarps + warp; lane == 0 && i < L.n_ws; i += gridDim.x * kWarps) {
                    L.ws[i].p0 = kNone; L.ws[i].empties_before = 0; L.ws[i].empties_total = 0;
                }
            return;
        }
    }
    const u32 delta32 = L.negate ? 0xFFFFFFFFu : 1u;

    for (u32 i = tid; i < n_entries; i += kThreads) tab[i] = make_uint4(0, 0, 0, 0xFFFFFFFFu);
    for (u32 i = tid; i < kDefRing; i += kThreads) def_ring[i] = make_uint4(0, 0, 0, 0);
    if (tid == 0) { s_ctl.tail = 0; s_ctl.head = 0; s_ctl.done = 0; s_ctl.abort_flag = 0; }
    __syncthreads();

    const u64 cta_base = (u64)blockIdx.x * n_producers * L.tiles_per_warp * kTile; // buffer-relative
    if (warp >= n_producers) { // ---- consumer warps: everything that needs global memory
        consume_deferred<Map>(def_ring, &s_ctl, &L, &T, cta_base, tab);
    } else { // ---- producer warps: tokenize and count

    // per-warp shared memory: [two tile buffers] [delimiter list]
    unsigned char *wbase = warp_all + warp * kWarpSmemBytes;
    unsigned short *dl = reinterpret_cast<unsigned short *>(wbase + kRingBytes);
    uint4 *stage = reinterpret_cast<uint4 *>(wbase + kRingBytes + kDlBytes);
    u32 n_staged = 0; // warp-uniform

    const u32 tiles_total = (u32)((L.n + kTile - 1) / kTile);
    const u32 gwarp = blockIdx.x * n_producers + warp;
    u32 t0 = gwarp * L.tiles_per_warp < tiles_total ? gwarp * L.tiles_per_warp : tiles_total;
    const u32 t1 = t0 + L.tiles_per_warp < tiles_total ? t0 + L.tiles_per_warp : tiles_total;
    if (range_lo > 0 && t0 < t1) { // roll-back: start at the tile that holds range_lo
        const u64 tl = range_lo / kTile;
        if (tl >= t1) t0 = t1; else if (tl > t0) t0 = (u32)tl;
    }
    const u32 lt_mask = (1u << lane) - 1;

    u32 acc_flush = 0, acc_empty = 0, acc_slow = 0;  // per lane; a warp's range is < 4 GiB
    bool seen_nonempty = L.negate != 0;              // roll-back launches do not look for P0
    u32 p0_rel = 0, my_empties_before = 0;           // P0 relative to cta_base

    if (t0 < t1) {
        // carry from tile to tile: position of the last delimiter (relative to the NEXT tile's start,
        // so always negative), the line state, and whether the previous byte was a line feed
        int carry_prev; bool killed, last_is_lf;
        u32 lmod = 0; // strict-ref: line offset of the tile's first byte, modulo the reference's line buffer
        if (t0 == 0) { carry_prev = -1; killed = false; last_is_lf = true; }
        else {
            const WarpCarry c = warp_lookback(L.buf, (u64)t0 * kTile);
            const long long cr = c.last_delim - (long long)t0 * kTile;
            carry_prev = cr < -(1ll << 30) ? -(1 << 30) : (int)cr;
            killed = c.killed; last_is_lf = c.last_is_lf;
            if (L.strict) { // where in its line's 99-byte pieces does this range begin?
                const long long pos = (long long)t0 * kTile;
                lmod = (u32)((pos - c.line_start) % (long long)kLineCap);
                const long long c0 = pos - (long long)lmod; // the last piece start at or before pos
                if (c0 > c.line_start) { // it lies inside the line: the piece, not the line, bounds what came before
                    killed = lmod ? warp_any_kill(L.buf, c0, pos) : false;
                    if (c.last_delim < c0 - 1) carry_prev = -(int)lmod - 1;
                }
            }
            // the bytes just before the range, where a word ending in the first tile may start
            if (lane == 0)
                *reinterpret_cast<uint4 *>(wbase + (t0 & 1) * kBufBytes) = load16(L.buf, L.n, (u64)t0 * kTile - 16);
        }
        fetch_tile(L, t0, wbase + (t0 & 1) * kBufBytes, lane);
        const u32 last_tile = (u32)((L.n - 1) / kTile); // holds the piece's last byte
        for (u32 t = t0; t < t1; ++t) {
            unsigned char *buf = wbase + (t & 1) * kBufBytes;
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            if (t + 1 < t1) { // next tile into the other buffer, with this tile's last 16 bytes in front of it
                unsigned char *nbuf = wbase + ((t + 1) & 1) * kBufBytes;
                if (lane == 31) *reinterpret_cast<uint4 *>(nbuf) = *reinterpret_cast<const uint4 *>(buf + kTile);
                fetch_tile(L, t + 1, nbuf, lane);
            }
            const u32 *ring32 = reinterpret_cast<const u32 *>(buf);
            // Each lane owns 32 consecutive bytes of the tile: one 32-bit mask per property.
            const uint4 va = *reinterpret_cast<const uint4 *>(buf + 16 + lane * 32);
            const uint4 vb = *reinterpret_cast<const uint4 *>(buf + 32 + lane * 32);
            const u64 T0 = (u64)t * kTile;          // buffer-relative offset of this tile
            const u32 relT = (u32)(T0 - cta_base);

            // strict-ref: a piece start at the tile's first byte acts like a delimiter just before the tile
            if (L.strict && lmod == 0 && !last_is_lf) carry_prev = -1;
            u32 cs_mid = 0, lmod_next = 0; // strict-ref: this lane's piece start that is not a line start (one bit at most)

            // ---- classify 2 x 16 bytes
            u32 dm, lf, live;
            {
                // (one half at a time: keeps the number of live registers down)
                u32 ex;
                {
                    const FastFlags a0 = classify_fast4(va.x), a1 = classify_fast4(va.y), a2 = classify_fast4(va.z),
                                    a3 = classify_fast4(va.w);
                    dm = movemask16(a0.cand, a1.cand, a2.cand, a3.cand);
                    lf = movemask16(a0.lf, a1.lf, a2.lf, a3.lf);
                    ex = a0.ex | a1.ex | a2.ex | a3.ex;
                }
                asm volatile("" ::: "memory");
                {
                    const FastFlags b0 = classify_fast4(vb.x), b1 = classify_fast4(vb.y), b2 = classify_fast4(vb.z),
                                    b3 = classify_fast4(vb.w);
                    dm |= movemask16(b0.cand, b1.cand, b2.cand, b3.cand) << 16;
                    lf |= movemask16(b0.lf, b1.lf, b2.lf, b3.lf) << 16;
                    ex |= b0.ex | b1.ex | b2.ex | b3.ex;
                }
                live = dm;
                if (L.strict) {
                    // Line offset (mod 99) of this lane's first byte: from the nearest line feed in the
                    // lanes before it, else from the carry.  A lane holds at most one piece start that is
                    // not a line start: at p* = (99 - offset) mod 99, if that is before its first line feed.
                    const int mylast = lf ? (int)(lane * 32 + 31 - __clz(lf)) : -1;
                    int incl = mylast;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const int tv = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                        if (lane >= (u32)d && tv > incl) incl = tv;
                    }
                    int qprev = __shfl_up_sync(0xFFFFFFFFu, incl, 1);
                    if (lane == 0) qprev = -1;
                    const u32 off_mod = qprev >= 0 ? (u32)((int)(lane * 32) - (qprev + 1)) % kLineCap : (lmod + lane * 32) % kLineCap;
                    const bool line_start = qprev >= 0 ? (qprev + 1 == (int)(lane * 32)) : (lane == 0 && last_is_lf);
                    const u32 pstar = off_mod == 0 ? 0u : kLineCap - off_mod;
                    const u32 first_lf = lf ? (u32)__ffs(lf) - 1 : 32u;
                    if (pstar < 32 && pstar <= first_lf && !(pstar == 0 && line_start)) cs_mid = 1u << pstar;
                    const int last_lf_tile = __shfl_sync(0xFFFFFFFFu, incl, 31);
                    lmod_next = last_lf_tile >= 0 ? (u32)(kTile - 1 - last_lf_tile) % kLineCap : (lmod + kTile) % kLineCap;
                    if (__any_sync(0xFFFFFFFFu, ex != 0 || cs_mid != 0) || killed) {
                        acc_slow += lane == 0;
                        const u32 up_lf = __shfl_up_sync(0xFFFFFFFFu, lf >> 31, 1);
                        const u32 starts = (lf << 1) | (lane == 0 ? (last_is_lf ? 1u : 0u) : up_lf) | cs_mid;
                        const u32 up_st = __shfl_up_sync(0xFFFFFFFFu, starts >> 31, 1);
                        const u32 carry_start = lane == 0 ? (lmod == 1 ? 1u : 0u) : up_st;
                        const StrictOut so = strict_exact(va, vb, starts, carry_start, killed, L.base + T0 + lane * 32,
                                                          &st->stop, L.negate);
                        dm = so.dm; live = so.live; lf = so.lfm; killed = so.killed_out != 0;
                    }
                } else if (__any_sync(0xFFFFFFFFu, ex != 0) || killed) {
                    // The exact path sees the tile as two 512-byte halves of 16 bytes per lane; its masks
                    // are handed back to the lanes that own those bytes here.
                    acc_slow += lane == 0;
                    const uint4 xa = *reinterpret_cast<const uint4 *>(buf + 16 + lane * 16);
                    const uint4 xb = *reinterpret_cast<const uint4 *>(buf + 16 + kSub + lane * 16);
                    const ExactOut oa = exact_sub(xa, killed, L.buf, L.n, L.base, T0, &st->stop, L.negate);
                    const ExactOut ob = exact_sub(xb, oa.killed_out, L.buf, L.n, L.base, T0 + kSub, &st->stop, L.negate);
                    dm = halves_to_lanes(oa.dm, ob.dm); live = halves_to_lanes(oa.live, ob.live);
                    lf = halves_to_lanes(oa.lfm, ob.lfm);
                    killed = ob.killed_out != 0;
                }
            }

            // ---- what the byte before this lane's 32 bytes was: bit 0 line feed, bit 1 delimiter
            u32 prev;
            {
                const u32 mine = (lf >> 31) | ((dm >> 31) << 1);
                const u32 up = __shfl_up_sync(0xFFFFFFFFu, mine, 1);
                prev = lane == 0 ? ((last_is_lf ? 1u : 0u) | (carry_prev == -1 ? 2u : 0u)) : up;
                last_is_lf = (__shfl_sync(0xFFFFFFFFu, mine, 31) & 1u) != 0;
            }
            // ---- STOP: a line that is just "\n" (main.cu:185), or a 1-byte tail at the very end
            if (!L.negate) {
                const u32 start = (lf << 1) | (prev & 1u) | cs_mid; // bit p: byte p starts a line (strict-ref: or a piece)
                const u32 blank = lf & start;
                if (blank) atomicMin(&st->stop, L.base + T0 + lane * 32 + (__ffs(blank) - 1));
                if (t == last_tile && L.last_piece) {
                    const u32 q = (u32)(L.n - 1 - T0); // position of the last byte in this tile
                    if ((q >> 5) == lane && ((start >> (q & 31)) & 1)) atomicMin(&st->stop, L.base + L.n - 1);
                }
            }

            // Roll-back launches only see flush points at or after range_lo.
            if (range_lo > T0) {
                const u64 lo = T0 + lane * 32;
                if (range_lo >= lo + 32) live = 0;
                else if (range_lo > lo) live &= ~((1u << (u32)(range_lo - lo)) - 1);
            }

            // ---- flush counts straight from the masks: a flush right after a delimiter is an empty word
            const u32 after = (dm << 1) | (prev >> 1) | cs_mid; // (strict-ref: a flush at a piece start is an empty word too)
            const u32 em = live & after;
            acc_flush += __popc(live);
            acc_empty += __popc(em);
            if (!seen_nonempty) { // this warp's first non-empty word, and the empty ones before it
                const u32 ne = live & ~after;
                const u32 bn = __ballot_sync(0xFFFFFFFFu, ne != 0);
                u32 x = __popc(em);
                if (bn) {
                    const u32 fl = __ffs(bn) - 1;
                    const u32 pf = __shfl_sync(0xFFFFFFFFu, ne ? __ffs(ne) - 1 : 0, fl);
                    if (lane == fl) x = __popc(em & ((1u << pf) - 1));
                    if (lane > fl) x = 0;
                    p0_rel = relT + fl * 32 + pf;
                    seen_nonempty = true;
                }
                my_empties_before += __reduce_add_sync(0xFFFFFFFFu, x);
            }

            // ---- position-ordered list of every delimiter.  strict-ref: a piece start at byte c that does
            // not follow a delimiter adds an entry at c - 1 that never flushes: the word before it is
            // dropped there and the next one starts at c (main.cu:183: nIndex = 0 for every fgets).
            u32 lm = dm;
            if (L.strict) {
                const u32 dn = __shfl_down_sync(0xFFFFFFFFu, cs_mid & 1u, 1);
                lm |= ((cs_mid >> 1) | ((lane < 31 ? dn : 0u) << 31)) & ~dm;
            }
            const u32 cnt = __popc(lm);
            const u32 incl = warp_incl_scan(cnt);
            const u32 tot = __shfl_sync(0xFFFFFFFFu, incl, 31);
            const u32 excl = incl - cnt;

            // The list holds at most kDlCap entries: a tile dense with delimiters is done one half (16
            // lanes) at a time.  `first_prev` is the position (tile relative, negative if it lies before
            // the tile) of the delimiter that precedes list entry 0.
            int first_prev = carry_prev;
            const bool dense = tot > (u32)kDlCap;
            const u32 half0 = __shfl_sync(0xFFFFFFFFu, excl, 16); // delimiters in the first 512 bytes
            for (int ph = 0; ph < (dense ? 2 : 1); ph++) {
                if (!dense || (ph == 0) == (lane < 16)) {
                    u32 m = lm;
                    unsigned short *o = dl + excl - ((dense && ph) ? half0 : 0u);
                    while (m) {
                        const u32 p = __ffs(m) - 1; m &= m - 1;
                        *o++ = (unsigned short)((lane * 32 + p) | (((live >> p) & 1u) << 15));
                    }
                }
                __syncwarp();
                const u32 total = dense ? (ph ? tot - half0 : half0) : tot;

                // ---- one word per lane: a hit in the CTA's table is counted here, everything else
                // (misses, words of 9+ bytes) is queued for drain_deferred()
#ifdef WC_EXP_NOROUNDS
                for (u32 k0 = 0; k0 < total && relT == 0xFFFFFFF1u; k0 += 32) {
#else
#pragma unroll kRoundUnroll
                for (u32 k0 = 0; k0 < total; k0 += 32) {
#endif
                    const u32 k = k0 + lane;
                    const u32 d = k < total ? (u32)dl[k] : 0u;
                    const int e = (int)(d & 0x3FFu);
                    const int prev = k ? (int)(dl[k - 1] & 0x3FFu) : first_prev; // the delimiter before this one
                    const int len = e - prev - 1;
                    const bool word = (d >> 15) && len > 0;       // a flush of a non-empty word
                    const int rlen = len > 16 ? 16 : (len < 1 ? 1 : len); // bytes looked at (keeps reads in bounds)

                    // key = the word's first 8 bytes, from the warp's shared-memory copy of the tile
                    const u32 idx = (u32)(16 + e - rlen);          // >= 0: reaches into the previous tile's tail
                    const u32 *r = ring32 + (idx >> 2);
                    const u32 w0 = r[0], w1 = r[1], w2 = r[2];
                    const u64 keep = ~0ull >> (8u * (8u - (u32)(rlen > 8 ? 8 : rlen))); // the first min(len, 8) bytes
                    const u32 sh = (idx & 3) * 8;
                    const u32 klo = Map::map4(__funnelshift_r(w0, w1, sh) & (u32)keep),
                              khi = Map::map4(__funnelshift_r(w1, w2, sh) & (u32)(keep >> 32));

                    if (L.em.cursor != nullptr) { // ---- sort path: write the word out instead of counting it
                        const bool emit = word && len <= 16 && !L.negate; // (what lies after STOP is dropped at the end)
                        const u32 be = __ballot_sync(0xFFFFFFFFu, emit);
                        if (be) {
                            u64 at = 0;
                            if (lane == 0) at = atomicAdd(L.em.cursor, (u64)__popc(be));
                            at = __shfl_sync(0xFFFFFFFFu, at, 0) + __popc(be & lt_mask);
                            if (emit && at < L.em.cap) {
                                u32 k1lo = 0, k1hi = 0;
                                if (len > 8) { // bytes 8..15
                                    const u64 keep1 = ~0ull >> (8u * (16u - (u32)len));
                                    k1lo = Map::map4(__funnelshift_r(w2, r[3], sh) & (u32)keep1);
                                    k1hi = Map::map4(__funnelshift_r(r[3], r[4], sh) & (u32)(keep1 >> 32));
                                }
                                const u64 k0 = (u64)klo | ((u64)khi << 32), k1 = (u64)k1lo | ((u64)k1hi << 32);
                                L.em.key[at] = make_ulonglong2(k0, k1); // (sort keys are derived from these at the end)
                                L.em.end[at] = L.base + T0 + (u64)e;
                            } else if (emit) atomicOr(&st->err, kErrEmitFull);
                        }
                        const bool dlong = word && len > 16; // the few longer words still go to the long table
                        const u32 bl = __ballot_sync(0xFFFFFFFFu, dlong);
                        if (bl) {
                            u32 pos = 0;
                            if (lane == 0) pos = atoms_add(&s_ctl.tail, (u32)__popc(bl));
                            pos = __shfl_sync(0xFFFFFFFFu, pos, 0) + __popc(bl & lt_mask);
                            const u32 cell = pos & (kDefRing - 1);
                            u32 spins = 0;
                            while (__any_sync(0xFFFFFFFFu, dlong && reinterpret_cast<volatile uint4 *>(def_ring)[cell].w != ring_lap(pos))) {
                                if (__any_sync(0xFFFFFFFFu, ++spins > (1u << 25))) { s_ctl.abort_flag = 1; atomicOr(&st->err, kErrRingTimeout); break; }
                                __nanosleep(64);
                            }
                            if (dlong) def_ring[cell] = make_uint4((u32)len, khi, relT + (u32)e, ring_full_word(pos, (u32)len, false));
                        }
                        continue;
                    }

                    // single-probe shared-memory table
                    const u32 slot = __umulhi(hash32_of_cta(klo, khi), n_entries);
                    const uint4 ent = tab[slot];
                    const bool hit = word && len <= 8 && ent.x == klo && ent.y == khi;
                    const u32 rel = relT + (u32)e;
                    u32 *cell = reinterpret_cast<u32 *>(&tab[slot]);
#ifndef WC_EXP_NOATOM
                    if (hit) atomicAdd(cell + 2, delta32);
                    if (hit && rel < ent.w) atomicMin(cell + 3, rel);
#else
                    if (hit && rel == 0xFFFFFFF1u) atomicAdd(cell + 2, delta32);
#endif

#ifdef WC_EXP_NODEFER
                    const bool defer = false;
#else
                    const bool defer = word && !hit;
#endif
                    const u32 bq = __ballot_sync(0xFFFFFFFFu, defer);
                    if (bq) { // stage the misses and the 9+ byte words; full stages go to the consumer warps' ring
                        if (defer) // w: length, bit 31 = the word's shared-memory slot belongs to another word
                            stage[n_staged + __popc(bq & lt_mask)] =
                                make_uint4(klo, khi, rel, (u32)len | ((ent.x | ent.y) != 0 ? 0x80000000u : 0u));
                        n_staged += __popc(bq);
                        if (n_staged > kStageCap - 32) { // the next round might not fit
                            __syncwarp();
                            flush_stage(stage, n_staged, def_ring, &s_ctl, st);
                            n_staged = 0;
                        }
                    }
                }
                if (total > 0) first_prev = (int)(dl[total - 1] & 0x3FFu);
                __syncwarp();
            }
            // the next tile sees this tile's last delimiter at a negative position
            carry_prev = first_prev - kTile < -(1 << 30) ? -(1 << 30) : first_prev - kTile;
            lmod = lmod_next;
        }
    }
    // ---- per-warp results
    const u64 w_flush = warp_sum64(acc_flush), w_empty = warp_sum64(acc_empty);
    if (lane == 0) {
        if (w_flush) atomicAdd(&st->n_flush, L.negate ? (u64)(-(long long)w_flush) : w_flush);
        if (w_empty) atomicAdd(&st->n_empty, L.negate ? (u64)(-(long long)w_empty) : w_empty);
        if (!L.negate) {
            if (acc_slow) atomicAdd(&st->n_slow_tiles, (u64)acc_slow);
            if (L.ws != nullptr && gwarp < L.n_ws) {
                L.ws[gwarp].p0 = (seen_nonempty && t0 < t1) ? L.base + cta_base + p0_rel : kNone;
                L.ws[gwarp].empties_before = my_empties_before;
                L.ws[gwarp].empties_total = w_empty;
            }
        }
    }

    __syncwarp();
    if (n_staged) flush_stage(stage, n_staged, def_ring, &s_ctl, st);
    if (lane == 0) atomicAdd(&s_ctl.done, 1u); // this producer will append nothing more
    } // producer warps

    // ---- flush the CTA's table into the global one
    __syncthreads();
    u32 n_wnew = 0;
    for (u32 i = tid; i < n_entries; i += kThreads) {
        const uint4 ent = tab[i];
        if ((ent.x | ent.y) != 0 && ent.z != 0) {
            const u64 key = (u64)ent.x | ((u64)ent.y << 32);
            n_wnew += word_insert(T, key, 0ull, key_len8(key), (u64)(long long)(int)ent.z, L.base + cta_base + ent.w);
        }
    }
    const u32 w_wnew = __reduce_add_sync(0xFFFFFFFFu, n_wnew);
    if (lane == 0 && w_wnew) atomicAdd(&st->w_entries, (u64)w_wnew);
}

} // namespace wc
