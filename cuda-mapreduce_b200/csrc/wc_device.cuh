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

constexpr int kTile = 512;                 // bytes per warp-tile: 32 lanes x 16 B
constexpr int kWarps = 16;                 // warps per CTA
constexpr int kThreads = kWarps * 32;
constexpr u32 kSmemEntries = 11264;        // 16-byte entries of the per-CTA table (176 KiB)
constexpr u32 kSmemLimit = 9011;           // stop claiming new slots at 80 % load
constexpr int kSmemMaxProbe = 12;
constexpr int kRingBytes = 1024;           // per warp: previous + current tile
constexpr int kLongQueue = 64;             // per warp: pending long words
constexpr u64 kNone = ~0ull;

constexpr size_t kSmemTableBytes = (size_t)kSmemEntries * 16;
constexpr size_t kSmemRingBytes = (size_t)kWarps * kRingBytes;
constexpr size_t kSmemDlistBytes = (size_t)kWarps * kTile * 2;
constexpr size_t kSmemQueueBytes = (size_t)kWarps * kLongQueue * 12;
constexpr size_t kSmemTotal = kSmemTableBytes + kSmemRingBytes + kSmemDlistBytes + kSmemQueueBytes;

enum : u32 { kErrShortFull = 1u, kErrLongFull = 2u, kErrArenaFull = 4u, kErrWordTooLong = 8u };

// Device-resident job state (one per engine).
struct Status {
    u64 n_flush;      // words emitted, empty ones included (two's complement adds for roll-back)
    u64 n_empty;
    u64 stop;         // min offset at which the reference stops reading (main.cu:185); kNone if none
    u64 s_entries;    // claimed slots, short table
    u64 l_entries;    // claimed slots, long table
    u64 n_long;       // words > 8 bytes
    u64 n_spill;      // short words that went straight to the global table
    u64 n_slow_tiles; // tiles that took the CR / NUL line-state path
    u64 arena_used;   // 8-byte units handed out (unit 0 is never used)
    u64 p0;           // offset that flushed the first non-empty word
    u64 lost_empty;   // empty words before p0 (main.cu:81,103: lost)
    u32 p0_found;
    u32 err;
};

struct Tables {
    // words of 1..8 bytes: the key is the word itself, zero padded (exact by value)
    u64 *s_keys, *s_counts, *s_finv; // finv = ~first_offset, so 0 == unset and atomicMax keeps the min
    u32 s_log2;
    // words of 9+ bytes: slot = (24-bit hash tag << 40) | arena unit index; bytes compared in the arena
    u64 *l_refs, *l_counts, *l_finv;
    u32 l_log2;
    u64 *arena;       // unit[i] = length, unit[i+1..] = bytes, zero padded to 8
    u64 arena_units;
    Status *st;
};

struct WarpSummary { // per warp, in input order: what finding P0 needs
    u64 p0;            // offset that flushed this warp's first non-empty word, kNone if none
    u64 empties_before; // empty words this warp emitted before that
    u64 empties_total;
};

struct Launch {
    const uint8_t *buf; // 16-byte aligned, starts at a line start
    u64 n;
    u64 base;           // offset of buf[0] in the whole input
    u32 tiles_per_warp;
    int last_piece;     // the input ends with this piece
    int negate;         // roll-back launch: subtract everything at or after Status::stop
    WarpSummary *ws;
    u32 n_ws;
};

// ---------------------------------------------------------------- helpers

__device__ __forceinline__ u32 lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ u64 mix64(u64 x) {
    x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    return x;
}

__host__ __device__ __forceinline__ u32 hash32(u32 lo, u32 hi) {
    u32 h = lo * 0x9E3779B1u;
    h ^= hi * 0x85EBCA77u;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 13;
    return h;
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

__device__ __forceinline__ u64 warp_sum64(u64 v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
    return v;
}

// ------------------------------------------------------- global short table

struct ThreadCounters { u32 s_new, l_new, n_long, n_spill; };

__device__ __forceinline__ void gs_insert(const Tables &T, u64 key, u64 delta, u64 first_abs,
                                          ThreadCounters &tc) {
    const u64 mask = (1ull << T.s_log2) - 1;
    u64 slot = mix64(key) >> (64 - T.s_log2);
    for (u32 probes = 0;; ++probes) {
        u64 k = __ldcg(&T.s_keys[slot]);
        if (k == 0) {
            k = atomicCAS(&T.s_keys[slot], 0ull, key);
            if (k == 0) { tc.s_new++; k = key; }
        }
        if (k == key) break;
        slot = (slot + 1) & mask;
        if (probes > 4096 || probes >= mask) { atomicOr(&T.st->err, kErrShortFull); return; }
    }
    atomicAdd(&T.s_counts[slot], delta);
    u64 inv = ~first_abs;
    if (__ldcg(&T.s_finv[slot]) < inv) atomicMax(&T.s_finv[slot], inv);
}

// -------------------------------------------------------- global long table

__device__ __forceinline__ u64 long_hash(const uint8_t *buf, u64 start, u64 len) {
    u64 h = 0x9E3779B97F4A7C15ull ^ (len * 0xFF51AFD7ED558CCDull);
    u64 i = 0;
    for (; i + 8 <= len; i += 8) {
        h = (h ^ load8_unaligned(buf, start + i)) * 0x9FB21C651E98DF25ull;
        h ^= h >> 32;
    }
    if (i < len) {
        u64 v = load8_unaligned(buf, start + i, (u32)(len - i));
        h = (h ^ v) * 0x9FB21C651E98DF25ull;
        h ^= h >> 32;
    }
    return mix64(h);
}

__device__ __forceinline__ bool long_equal(const Tables &T, u64 unit, const uint8_t *buf, u64 start, u64 len) {
    if (__ldcg(&T.arena[unit]) != len) return false;
    u64 i = 0;
    for (; i + 8 <= len; i += 8)
        if (__ldcg(&T.arena[unit + 1 + (i >> 3)]) != load8_unaligned(buf, start + i)) return false;
    if (i < len) {
        u64 v = load8_unaligned(buf, start + i, (u32)(len - i));
        if (__ldcg(&T.arena[unit + 1 + (i >> 3)]) != v) return false;
    }
    return true;
}

// A word of 9+ bytes at buf[start, start+len).  Never waits on another thread: a new word is
// first written to the arena, then published with one CAS of the slot word.
__device__ __noinline__ void long_insert(const Tables &T, const uint8_t *buf, u64 start, u64 len,
                                         u64 delta, u64 first_abs, ThreadCounters &tc) {
    const u64 h = long_hash(buf, start, len);
    const u64 mask = (1ull << T.l_log2) - 1;
    const u64 tag = h >> 40;
    u64 slot = h & mask;
    u64 my_ref = 0;
    for (u32 probes = 0;; ++probes) {
        u64 r = __ldcg(&T.l_refs[slot]);
        if (r == 0) {
            if (my_ref == 0) {
                u64 units = 1 + ((len + 7) >> 3);
                u64 idx = atomicAdd(&T.st->arena_used, units);
                if (idx + units > T.arena_units) { atomicOr(&T.st->err, kErrArenaFull); return; }
                T.arena[idx] = len;
                u64 i = 0;
                for (; i + 8 <= len; i += 8) T.arena[idx + 1 + (i >> 3)] = load8_unaligned(buf, start + i);
                if (i < len) T.arena[idx + 1 + (i >> 3)] = load8_unaligned(buf, start + i, (u32)(len - i));
                __threadfence();
                my_ref = (tag << 40) | idx;
            }
            r = atomicCAS(&T.l_refs[slot], 0ull, my_ref);
            if (r == 0) { tc.l_new++; break; }
        }
        if ((r >> 40) == tag && long_equal(T, r & ((1ull << 40) - 1), buf, start, len)) break;
        slot = (slot + 1) & mask;
        if (probes > 4096 || probes >= mask) { atomicOr(&T.st->err, kErrLongFull); return; }
    }
    atomicAdd(&T.l_counts[slot], delta);
    u64 inv = ~first_abs;
    if (__ldcg(&T.l_finv[slot]) < inv) atomicMax(&T.l_finv[slot], inv);
}

// ------------------------------------------------------------ main kernel

struct WarpCarry {
    long long last_delim; // buffer-relative position of the last delimiter byte seen, -1 if none
    bool killed;          // a 0x0D / 0x00 since the last 0x0A: rest of the line is not split (main.cu:187,196)
    bool last_is_lf;      // previous byte was 0x0A (or this is the start of the piece)
};

// State at buffer offset `pos` (> 0), found by scanning backwards to the previous line start.
__device__ __forceinline__ WarpCarry warp_lookback(const uint8_t *buf, u64 pos) {
    const u32 lane = lane_id();
    WarpCarry c;
    c.last_delim = -1; c.killed = false;
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
            return c;
        }
        seen_kill = seen_kill || bk;
    }
    c.killed = seen_kill;
    return c;
}

__device__ __forceinline__ uint4 load_tile(const Launch &L, u64 tile, u32 lane) {
    const u64 off = tile * kTile + (u64)lane * 16;
    if (off + 16 <= L.n) return __ldcs(reinterpret_cast<const uint4 *>(L.buf + off));
    u32 w[4] = {0x78787878u, 0x78787878u, 0x78787878u, 0x78787878u}; // past the end: plain word bytes
    for (u32 i = 0; i < 16 && off + i < L.n; i++) {
        u32 sh = (i & 3) * 8;
        w[i >> 2] = (w[i >> 2] & ~(0xFFu << sh)) | ((u32)L.buf[off + i] << sh);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

extern __shared__ __align__(16) unsigned char wc_smem[];

__global__ void __launch_bounds__(kThreads, 1) count_kernel(const Launch L, const Tables T) {
    uint4 *tab = reinterpret_cast<uint4 *>(wc_smem);
    u64 *ring_all = reinterpret_cast<u64 *>(wc_smem + kSmemTableBytes);
    unsigned short *dl_all = reinterpret_cast<unsigned short *>(wc_smem + kSmemTableBytes + kSmemRingBytes);
    unsigned char *q_all = wc_smem + kSmemTableBytes + kSmemRingBytes + kSmemDlistBytes;
    __shared__ u32 s_entries;
    __shared__ u32 s_first_ws;
    __shared__ u64 s_stop;

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
    const u64 delta64 = L.negate ? ~0ull : 1ull;

    for (u32 i = tid; i < kSmemEntries; i += kThreads) tab[i] = make_uint4(0, 0, 0, 0xFFFFFFFFu);
    if (tid == 0) s_entries = 0;
    __syncthreads();

    u64 *ring = ring_all + warp * (kRingBytes / 8);
    unsigned short *dl = dl_all + warp * kTile;
    u64 *q_start = reinterpret_cast<u64 *>(q_all + warp * (kLongQueue * 12));
    u32 *q_len = reinterpret_cast<u32 *>(q_all + warp * (kLongQueue * 12) + kLongQueue * 8);

    const u64 tiles_total = (L.n + kTile - 1) / kTile;
    const u64 gwarp = (u64)blockIdx.x * kWarps + warp;
    u64 t0 = gwarp * L.tiles_per_warp;
    u64 t1 = t0 + L.tiles_per_warp < tiles_total ? t0 + L.tiles_per_warp : tiles_total;
    const u64 cta_base = (u64)blockIdx.x * kWarps * L.tiles_per_warp * kTile; // buffer-relative
    if (range_lo > 0 && t0 < t1) { // roll-back: start at the tile that holds range_lo
        u64 tl = range_lo / kTile;
        if (tl >= t1) t0 = t1; else if (tl > t0) t0 = tl;
    }

    ThreadCounters tc = {0, 0, 0, 0};
    u64 acc_flush = 0, acc_empty = 0, acc_slow = 0;
    u32 q_count = 0;
    bool seen_nonempty = false;
    u64 my_p0 = kNone, my_empties_before = 0;

    // Drain up to 32 queued long words, one per lane.
    auto drain_long = [&](u32 take) {
        if (lane < take) {
            u64 s = q_start[lane]; u64 l = q_len[lane];
            long_insert(T, L.buf, s, l, delta64, L.base + s + l, tc);
        }
        __syncwarp();
        u64 ms = 0; u32 ml = 0;
        if (lane + take < q_count) { ms = q_start[lane + take]; ml = q_len[lane + take]; }
        __syncwarp();
        if (lane + take < q_count) { q_start[lane] = ms; q_len[lane] = ml; }
        q_count -= take;
        __syncwarp();
    };

    if (t0 < t1) {
        WarpCarry carry;
        if (t0 == 0) { carry.last_delim = -1; carry.killed = false; carry.last_is_lf = true; }
        else {
            carry = warp_lookback(L.buf, t0 * kTile);
            reinterpret_cast<uint4 *>(ring)[((t0 - 1) & 1) * 32 + lane] = load_tile(L, t0 - 1, lane);
        }
        uint4 vnext = load_tile(L, t0, lane);
        for (u64 t = t0; t < t1; ++t) {
            const uint4 v = vnext;
            if (t + 1 < t1) vnext = load_tile(L, t + 1, lane);
            const u64 T0 = t * kTile; // buffer-relative offset of this tile
            reinterpret_cast<uint4 *>(ring)[(t & 1) * 32 + lane] = v;

            // ---- classify 16 bytes
            ByteFlags f0 = classify4(v.x), f1 = classify4(v.y), f2 = classify4(v.z), f3 = classify4(v.w);
            u32 dm = movemask16(f0.delim, f1.delim, f2.delim, f3.delim);
            u32 lfm = movemask16(f0.lf, f1.lf, f2.lf, f3.lf);
            const bool special = (f0.kill | f1.kill | f2.kill | f3.kill) != 0;
            u32 live = dm; // delimiters that flush a word
            const u32 prev_lf_in = __shfl_up_sync(0xFFFFFFFFu, lfm >> 15, 1);
            const u32 prev_lf = lane == 0 ? (carry.last_is_lf ? 1u : 0u) : prev_lf_in;

            if (__any_sync(0xFFFFFFFFu, special) || carry.killed) {
                // ---- exact line state: a delimiter flushes only if no 0x0D / 0x00
                // precedes it on its line.  Rare: CRLF files, stray NULs.
                acc_slow += lane == 0;
                const u32 km = movemask16(f0.kill, f1.kill, f2.kill, f3.kill);
                // per-lane transfer function of "killed": s -> reset ? val : (s | val)
                u32 reset = lfm != 0;
                u32 val = reset ? ((km >> (32 - __clz(lfm))) != 0) : (km != 0);
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    u32 r2 = __shfl_up_sync(0xFFFFFFFFu, reset, d);
                    u32 v2 = __shfl_up_sync(0xFFFFFFFFu, val, d);
                    if (lane >= (u32)d) { // compose: earlier (r2, v2) then mine
                        val = reset ? val : (val | v2);
                        reset = reset | r2;
                    }
                }
                // state entering this lane = transfer of lanes < lane applied to the carry
                u32 r_in = __shfl_up_sync(0xFFFFFFFFu, reset, 1), v_in = __shfl_up_sync(0xFFFFFFFFu, val, 1);
                bool s = carry.killed;
                if (lane > 0) s = r_in ? (v_in != 0) : (s || v_in);
                const u32 r_all = __shfl_sync(0xFFFFFFFFu, reset, 31), v_all = __shfl_sync(0xFFFFFFFFu, val, 31);
                const bool killed_out = r_all ? (v_all != 0) : (carry.killed || v_all);
                live = 0;
#pragma unroll
                for (int p = 0; p < 16; p++) {
                    const u32 bit = 1u << p;
                    if ((dm & bit) && !s) live |= bit;
                    if (km & bit) s = true;
                    if (lfm & bit) s = false;
                }
                // A NUL at line offset 0 or 1 makes strlen() < 2: the reference stops (main.cu:185).
                if (!L.negate) {
                    u32 nm = km & ~dm; // NUL bytes (0x0D is in dm)
                    while (nm) {
                        const u32 p = __ffs(nm) - 1; nm &= nm - 1;
                        const u64 i = T0 + lane * 16 + p;
                        if (i >= L.n) break;
                        const bool at0 = i == 0 || L.buf[i - 1] == 0x0A;
                        const bool at1 = !at0 && (i == 1 || L.buf[i - 2] == 0x0A);
                        if (at0) atomicMin(&st->stop, L.base + i);
                        else if (at1) atomicMin(&st->stop, L.base + i - 1);
                    }
                }
                carry.killed = killed_out;
            }

            // ---- STOP: a line that is just "\n" (main.cu:185), or a 1-byte tail at the very end
            if (!L.negate) {
                const u32 at_line_start = ((lfm << 1) | prev_lf) & 0xFFFFu; // bit p: byte p starts a line
                const u32 blank = lfm & at_line_start;
                if (blank) atomicMin(&st->stop, L.base + T0 + lane * 16 + (__ffs(blank) - 1));
                if (L.last_piece && L.n - 1 >= T0 + lane * 16 && L.n - 1 < T0 + lane * 16 + 16) {
                    const u32 p = (u32)(L.n - 1 - (T0 + lane * 16));
                    if ((at_line_start >> p) & 1) atomicMin(&st->stop, L.base + L.n - 1);
                }
            }
            carry.last_is_lf = __shfl_sync(0xFFFFFFFFu, lfm >> 15, 31) != 0;

            // Roll-back launches only see flush points at or after range_lo.
            if (range_lo > T0) {
                const u64 lane_lo = T0 + lane * 16;
                if (range_lo >= lane_lo + 16) live = 0;
                else if (range_lo > lane_lo) live &= ~((1u << (u32)(range_lo - lane_lo)) - 1);
            }

            // ---- list every delimiter of the tile in position order
            const u32 cnt = __popc(dm);
            const u32 incl = warp_incl_scan(cnt);
            const u32 total = __shfl_sync(0xFFFFFFFFu, incl, 31);
            {
                u32 idx = incl - cnt, m = dm;
                while (m) {
                    const u32 p = __ffs(m) - 1; m &= m - 1;
                    dl[idx++] = (unsigned short)((lane * 16 + p) | (((live >> p) & 1u) << 15));
                }
            }
            acc_flush += __popc(live);
            __syncwarp();

            // ---- one word per lane
            const long long carry_rel = carry.last_delim - (long long)T0; // <= -1
            for (u32 k0 = 0; k0 < total; k0 += 32) {
                const u32 k = k0 + lane;
                const bool valid = k < total;
                const u32 d = valid ? dl[k] : 0;
                const int e = (int)(d & 0x1FFu);
                const bool flushes = valid && (d >> 15);
                const u32 d_up = __shfl_up_sync(0xFFFFFFFFu, d, 1);
                long long prev;
                if (lane > 0) prev = (long long)(d_up & 0x1FFu);
                else prev = k > 0 ? (long long)(dl[k - 1] & 0x1FFu) : carry_rel;
                const long long len = (long long)e - prev - 1;
                const bool nonempty = flushes && len > 0;
                const bool empty = flushes && len == 0;
                acc_empty += empty;

                if (!seen_nonempty && !L.negate) { // still looking for this warp's first non-empty word
                    const u32 bn = __ballot_sync(0xFFFFFFFFu, nonempty);
                    const u32 be = __ballot_sync(0xFFFFFFFFu, empty);
                    if (bn) {
                        const u32 fl = __ffs(bn) - 1;
                        my_empties_before += __popc(be & ((1u << fl) - 1));
                        my_p0 = L.base + T0 + (u64)__shfl_sync(0xFFFFFFFFu, e, fl);
                        seen_nonempty = true;
                    } else my_empties_before += __popc(be);
                }

                const bool is_long = nonempty && len > 8;
                if (nonempty && !is_long) {
                    // ---- key = the word's bytes, from the shared-memory copy of this and the previous tile
                    const int s = e - (int)len; // >= -8
                    const u32 ridx = (u32)(((int)((t & 1) << 9) + s) & (kRingBytes - 1));
                    const u32 sh = (ridx & 7) * 8;
                    u64 key = ring[ridx >> 3] >> sh;
                    if ((ridx & 7) + (u32)len > 8) key |= ring[((ridx >> 3) + 1) & (kRingBytes / 8 - 1)] << (64 - sh);
                    if (len < 8) key &= (1ull << (8 * (u32)len)) - 1;
                    const u32 klo = (u32)key, khi = (u32)(key >> 32);
                    const u32 rel = (u32)(T0 + (u64)e - cta_base);

                    u32 slot = __umulhi(hash32(klo, khi), kSmemEntries);
                    bool done = false;
#pragma unroll 1
                    for (int probe = 0; probe < kSmemMaxProbe; ++probe) {
                        uint4 ent = tab[slot];
                        if (ent.x == klo && ent.y == khi) done = true;
                        else if ((ent.x | ent.y) == 0) {
                            if (*reinterpret_cast<volatile u32 *>(&s_entries) >= kSmemLimit) break;
                            const u64 old = atomicCAS(reinterpret_cast<u64 *>(&tab[slot]), 0ull, key);
                            if (old == 0) { atomicAdd(&s_entries, 1u); done = true; }
                            else if (old == key) done = true;
                            ent.w = 0xFFFFFFFFu;
                        }
                        if (done) {
                            u32 *cell = reinterpret_cast<u32 *>(&tab[slot]);
                            atomicAdd(cell + 2, delta32);
                            if (rel < ent.w) atomicMin(cell + 3, rel);
                            break;
                        }
                        slot = slot + 1 == kSmemEntries ? 0 : slot + 1;
                    }
                    if (!done) { // table full or a long probe chain: count in the global table
                        tc.n_spill++;
                        gs_insert(T, key, delta64, L.base + T0 + (u64)e, tc);
                    }
                }

                // ---- words of 9+ bytes are queued and handled 32 at a time
                const u32 bl = __ballot_sync(0xFFFFFFFFu, is_long);
                if (bl) {
                    if (is_long) {
                        const u32 qi = q_count + __popc(bl & ((1u << lane) - 1));
                        const long long start = (long long)T0 + prev + 1;
                        q_start[qi] = (u64)start;
                        if (len > 0x7FFFFFFFll) { atomicOr(&st->err, kErrWordTooLong); q_len[qi] = 9; }
                        else q_len[qi] = (u32)len;
                        tc.n_long++;
                    }
                    q_count += __popc(bl);
                    __syncwarp();
                    if (q_count >= 32) drain_long(32);
                }
            }
            if (total > 0) carry.last_delim = (long long)T0 + (long long)(dl[total - 1] & 0x1FFu);
            __syncwarp();
        }
        if (q_count > 0) drain_long(q_count);
    }

    // ---- per-warp results
    acc_flush = warp_sum64(acc_flush);
    acc_empty = warp_sum64(acc_empty);
    const u32 w_snew = __reduce_add_sync(0xFFFFFFFFu, tc.s_new), w_lnew = __reduce_add_sync(0xFFFFFFFFu, tc.l_new);
    const u32 w_long = __reduce_add_sync(0xFFFFFFFFu, tc.n_long), w_spill = __reduce_add_sync(0xFFFFFFFFu, tc.n_spill);
    if (lane == 0) {
        if (acc_flush) atomicAdd(&st->n_flush, L.negate ? (u64)(-(long long)acc_flush) : acc_flush);
        if (acc_empty) atomicAdd(&st->n_empty, L.negate ? (u64)(-(long long)acc_empty) : acc_empty);
        if (w_snew) atomicAdd(&st->s_entries, (u64)w_snew);
        if (w_lnew) atomicAdd(&st->l_entries, (u64)w_lnew);
        if (!L.negate) {
            if (w_long) atomicAdd(&st->n_long, (u64)w_long);
            if (w_spill) atomicAdd(&st->n_spill, (u64)w_spill);
            if (acc_slow) atomicAdd(&st->n_slow_tiles, acc_slow);
            if (L.ws != nullptr && gwarp < L.n_ws) {
                L.ws[gwarp].p0 = my_p0;
                L.ws[gwarp].empties_before = my_empties_before;
                L.ws[gwarp].empties_total = acc_empty;
            }
        }
    }

    // ---- flush the CTA's table into the global one
    __syncthreads();
    ThreadCounters fc = {0, 0, 0, 0};
    for (u32 i = tid; i < kSmemEntries; i += kThreads) {
        const uint4 ent = tab[i];
        if ((ent.x | ent.y) != 0 && ent.z != 0) {
            const u64 key = (u64)ent.x | ((u64)ent.y << 32);
            gs_insert(T, key, (u64)(long long)(int)ent.z, L.base + cta_base + ent.w, fc);
        }
    }
    const u32 f_new = __reduce_add_sync(0xFFFFFFFFu, fc.s_new);
    if (lane == 0 && f_new) atomicAdd(&st->s_entries, (u64)f_new);
}

// ------------------------------------------------------------ table growth

__global__ void rehash_short_kernel(const u64 *keys, const u64 *counts, const u64 *finv, u64 n_old, Tables T) {
    ThreadCounters tc = {0, 0, 0, 0};
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n_old; i += (u64)gridDim.x * blockDim.x) {
        const u64 k = keys[i];
        if (k == 0) continue;
        gs_insert(T, k, counts[i], ~finv[i], tc);
    }
}

__global__ void rehash_long_kernel(const u64 *refs, const u64 *counts, const u64 *finv, u64 n_old, Tables T) {
    const u64 mask = (1ull << T.l_log2) - 1;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n_old; i += (u64)gridDim.x * blockDim.x) {
        const u64 r = refs[i];
        if (r == 0) continue;
        const u64 unit = r & ((1ull << 40) - 1);
        const u64 len = T.arena[unit];
        // same hash as long_hash(), over the arena copy (8-byte aligned, zero padded)
        u64 h = 0x9E3779B97F4A7C15ull ^ (len * 0xFF51AFD7ED558CCDull);
        for (u64 j = 0; j < ((len + 7) >> 3); j++) { h = (h ^ T.arena[unit + 1 + j]) * 0x9FB21C651E98DF25ull; h ^= h >> 32; }
        h = mix64(h);
        u64 slot = h & mask;
        for (;;) { // distinct words only: claim the first free slot
            if (atomicCAS(&T.l_refs[slot], 0ull, r) == 0) break;
            slot = (slot + 1) & mask;
        }
        T.l_counts[slot] = counts[i];
        T.l_finv[slot] = finv[i];
    }
}

} // namespace wc
