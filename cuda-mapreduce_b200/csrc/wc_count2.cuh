// wc_count2.cuh -- the forward tokenize+count kernel of the hash path (sm_100a).
//
// Replaces, in one pass over raw bytes, the reference's host splitter (/root/reference/main.cu:174-206),
// its mapper (main.cu:37-54, mapKernel main.cu:109-117) and the counting half of its reducer
// (main.cu:69-108).  Same results as wc::count_kernel (wc_device.cuh), which stays the kernel of the
// roll-back launches, of strict-ref mode and of the sort path; this one is built for the common case
// and for instruction count:
//
//   * every warp tokenizes (no consumer warps, no ring between warps): a warp streams its byte range
//     in 1 KiB tiles through a double-buffered cp.async copy in shared memory, each lane owns 32
//     consecutive bytes;
//   * no delimiter list: a lane walks the set bits of its own 32-bit mask of word ends; the word that
//     ends at a delimiter starts one byte after the next lower delimiter (one FLO on the lane's mask),
//     so there is no list to build, store and read back, and no scan -- the position of the last
//     delimiter before a lane's bytes comes from one ballot and one shuffle;
//   * the word's first 8 bytes come from the tile copy (3 x LDS.32 + 2 funnel shifts), the length mask
//     from a 9-entry table in shared memory, then one probe of the CTA's main table (8-byte keys):
//     a hit is one shared atomicAdd (+ a rare atomicMin of the first offset);
//   * everything else only sets a bit in a per-lane mask.  At the end of the tile the warp lists those
//     words ({end position, length}, 4 bytes each, one prefix sum over the lanes' counts) and settles them
//     32 at a time, one per lane (settle_tile): a word of 1..8 bytes takes its main-table slot if that is
//     still free; otherwise words of 1..16 bytes try the CTA's second table (16-byte keys, one probe) --
//     the home of frequent 9..16 byte words, which would otherwise hammer one address of the global
//     table; what is left (the long tail of rare words) is counted in the global word table with one
//     slot read and one fire-and-forget red.global.add per record.  The expensive part of a miss is thus
//     done at full lane occupancy, not inside the divergent word loop.
//
// Semantics: SURVEY.md Appendix A.2, identical to count_kernel's (the exact line-state path, STOP
// rules, P0 / lost-empties bookkeeping are the same code); parity is tested against oracle/ in tests/.
#pragma once
#include "wc_device.cuh"

namespace wc {

#ifndef WC2_WARPS
#define WC2_WARPS 28
#endif
#ifndef WC2_SECOND
#define WC2_SECOND 512   // entries of the CTA's second table (32 bytes each); 0: none
#endif
constexpr int k2Warps = WC2_WARPS;
constexpr int k2Threads = k2Warps * 32;
constexpr u32 k2ListCap = 256;                 // per warp: the tile's words that missed ({position, length}, 4 bytes each);
constexpr int k2ListBytes = k2ListCap * 4;     // half a tile never holds more words than that
constexpr int k2ScratchBytes = 64;             // the warp's rarely used counters
constexpr size_t k2WarpSmemBytes = kRingBytes + k2ListBytes + k2ScratchBytes;
constexpr size_t k2StaticSmem = 1024;          // (static __shared__ of the kernel, rounded up)
constexpr size_t k2KeepBytes = 128;            // the 9 length masks
constexpr u32 k2Second = WC2_SECOND;           // second table: 32-byte entries {key 16 B, count, first}, one probe
constexpr size_t k2SecondBytes = (size_t)k2Second * 32;
constexpr u32 k2Entries =
    (u32)((232448 - k2StaticSmem - k2KeepBytes - k2SecondBytes - k2Warps * k2WarpSmemBytes) / 16) & ~31u;
constexpr size_t k2TableBytes = (size_t)k2Entries * 16;
// layout: [main table][length masks][second table][per-warp areas]
constexpr size_t k2SmemTotal = k2TableBytes + k2KeepBytes + k2SecondBytes + (size_t)k2Warps * k2WarpSmemBytes;
static_assert(k2WarpSmemBytes % 16 == 0 && k2KeepBytes % 32 == 0, "shared-memory areas must keep their alignment");
static_assert(k2SmemTotal + k2StaticSmem <= 232448, "more shared memory than an sm_100 CTA can have");

enum : u32 { kScrWordNew = 0, kScrLongNew = 1, kScrSpill = 2, kScrLong = 3, kScrSlow = 4, kScrEmptiesBefore = 5, kScrP0 = 6 };
constexpr u64 kSecondLock = 0x100ull; // never a key: a word's first byte is not 0x00
constexpr u32 kListLenMax = (1u << 22) - 1; // a list entry is position | length << 10; longer words: the length is found by a scan

// Slot hash of the main table.  The first multiplier differs from CTA to CTA: two frequent words that
// share a slot -- one of them then goes to the slow path every time -- do so in one CTA, not in all.
__device__ __forceinline__ u32 hash32_cta(u32 lo, u32 hi, u32 mul) { return lo * mul + hi * 0x85EBCA77u; }
__device__ __forceinline__ u32 cta_hash_mul() { return 0x9E3779B1u + 2u * (blockIdx.x * 0x632BE5ABu); }

__device__ __forceinline__ u64 atoms_cas_u64_addr(u32 saddr, u64 cmp, u64 val) {
    u64 old;
    asm volatile("atom.shared.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "r"(saddr), "l"(cmp), "l"(val) : "memory");
    return old;
}
__device__ __forceinline__ void reds_add_addr(u32 saddr, u32 v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ void reds_min_addr(u32 saddr, u32 v) {
    asm volatile("red.shared.min.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ u64 lds_volatile_u64_addr(u32 saddr) {
    u64 v;
    asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(v) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_volatile_u64_addr(u32 saddr, u64 v) {
    asm volatile("st.volatile.shared.u64 [%0], %1;" ::"r"(saddr), "l"(v) : "memory");
}
__device__ __forceinline__ u32 lds_u32_addr(u32 saddr) {
    u32 v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
    return v;
}

// Bytes [off, off + 8) of the warp's tile copy (off relative to the tile's byte 0, may be negative down
// to -16), as a little-endian u64.
__device__ __forceinline__ u64 tile_bytes8(u32 tile_addr, int off) {
    const u32 a = tile_addr + (u32)off;
    const u32 w0 = lds_u32_addr(a & ~3u), w1 = lds_u32_addr((a & ~3u) + 4), w2 = lds_u32_addr((a & ~3u) + 8);
    const u32 sh = a << 3;
    return (u64)__funnelshift_r(w0, w1, sh) | ((u64)__funnelshift_r(w1, w2, sh) << 32);
}

// One entry of the second table: is the word (k0, k1) here, or can it move in?  The key is read with
// one 16-byte load; an entry that is being written shows the lock value and simply does not match.
__device__ __forceinline__ bool second_probe(u32 e, u64 k0, u64 k1) {
    uint4 k;
    asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(k.x), "=r"(k.y), "=r"(k.z), "=r"(k.w) : "r"(e) : "memory");
    const u64 e0 = (u64)k.x | ((u64)k.y << 32), e1 = (u64)k.z | ((u64)k.w << 32);
    if (e0 == k0 && e1 == k1) return true;
    if (e0 != 0) return false;
    const u64 old = atoms_cas_u64_addr(e, 0ull, kSecondLock);
    if (old != 0) return false; // (somebody else is moving in; if it is this very word, the next one finds it)
    sts_volatile_u64_addr(e + 8, k1); // bytes 8..15 first, then the first 8 bytes publish the entry
    __threadfence_block();
    sts_volatile_u64_addr(e, k0);
    return true;
}

// The next word of a lane in the word loop: takes the highest bit of `rem` (the lane's word ends still to
// do), finds where the word starts (one byte after the next lower delimiter of the lane, or after q_in),
// reads its first 8 bytes from the tile copy and computes its main-table slot.
struct WordProbe { u32 cb, klo, khi, ent_addr, rel; int len; bool word; };
template <class Map>
__device__ __forceinline__ WordProbe next_word(u32 &rem, u32 dm, int lane_pos, int q_in, u32 tile_addr, u32 keep_addr,
                                               u32 tab_addr, u32 hmul, u32 rel_tile) {
    WordProbe w;
    w.word = rem != 0;
    const u32 c = 31 - __clz(rem);                    // the word's delimiter (0xFFFFFFFF: none left)
    w.cb = 1u << (c & 31);
    rem &= ~w.cb;
    const u32 below = dm & (w.cb - 1);                // the lane's delimiters before it
    const int cur = lane_pos + (int)c;
    const int nxt = below ? lane_pos + 31 - (int)__clz(below) : q_in;
    w.len = cur - nxt - 1;                            // >= 1 for a word
    int s = nxt + 1;                                  // where the word starts (may lie before the tile)
    if (s < cur - 16) s = cur - 16;                   // (17+ bytes: the key is not used; keeps the reads in the buffer)
    // key = the word's first 8 bytes, from the warp's shared-memory copy of the tile
    const u32 a = tile_addr + (u32)s;
    u32 w0, w1, w2, mlo, mhi;
    asm("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(a & ~3u));
    asm("ld.shared.u32 %0, [%1+4];" : "=r"(w1) : "r"(a & ~3u));
    asm("ld.shared.u32 %0, [%1+8];" : "=r"(w2) : "r"(a & ~3u));
    const u32 mi = (u32)w.len < 8u ? (u32)w.len : 8u;
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(mlo), "=r"(mhi) : "r"(keep_addr + mi * 8));
    const u32 sh = a << 3; // the funnel shifts use the low 5 bits: (a & 3) * 8
    w.klo = Map::map4(__funnelshift_r(w0, w1, sh) & mlo);
    w.khi = Map::map4(__funnelshift_r(w1, w2, sh) & mhi);
    w.ent_addr = tab_addr + __umulhi(hash32_cta(w.klo, w.khi, hmul), k2Entries) * 16;
    w.rel = rel_tile + (u32)cur;
    return w;
}

// Settles the words of one tile that the word loop could not count: `total` list entries
// (position of the flushing delimiter in the tile | length << 10), 32 at a time, one per lane.
// The tile is still in shared memory (tile_addr = its byte 0, rel_tile = its offset from cta_base),
// so every byte of a word of up to 16 bytes is read from there.
template <class Map>
__device__ __noinline__ void settle_tile(u32 list_addr, u32 total, u32 tile_addr, u32 rel_tile, u64 tile_off,
                                         const Launch *Lp, const Tables *Tp, u32 tab_addr, u32 hmul) {
    const u32 lane = lane_id();
    const Tables &T = *Tp;
    const u32 keep_addr = tab_addr + (u32)k2TableBytes;
    const u32 scr = list_addr + k2ListBytes;
    u32 n_spill = 0, n_long = 0;
#pragma unroll 1
    for (u32 base = 0; base < total; base += 32) {
        const bool have = base + lane < total;
        u32 e = 0;
        if (have) e = lds_u32_addr(list_addr + (base + lane) * 4);
        const int p = (int)(e & 0x3FFu);
        u32 len = e >> 10;
        u64 k0 = 0, k1 = 0;
        if (have && len <= 16) { // the whole word, from the tile copy
            const int s = p - (int)len;
            const u32 m0 = len < 8u ? len : 8u, m1 = len - m0;
            k0 = Map::map8(tile_bytes8(tile_addr, s) & lds_volatile_u64_addr(keep_addr + m0 * 8));
            if (m1) k1 = Map::map8(tile_bytes8(tile_addr, s + 8) & lds_volatile_u64_addr(keep_addr + m1 * 8));
        }
        bool done = !have;
        const u32 rel = rel_tile + (u32)p;
        // A word of 1..8 bytes: its main-table slot may be free (this is how that table fills up: first
        // come, first served), or another warp may have put the word there since the word loop looked.
        if (!done && len <= 8) {
            const u32 ent_addr = tab_addr + __umulhi(hash32_cta((u32)k0, (u32)(k0 >> 32), hmul), k2Entries) * 16;
            u64 cur = lds_volatile_u64_addr(ent_addr);
            if (cur == 0) cur = atoms_cas_u64_addr(ent_addr, 0ull, k0);
            if (cur == 0 || cur == k0) {
                done = true;
                reds_add_addr(ent_addr + 8, 1u);
                reds_min_addr(ent_addr + 12, rel);
            }
        }
        // The second table: one probe, the whole word (up to 16 bytes) is the key.
        if (k2Second != 0 && !done && len <= 16) {
            const u32 h = ((u32)k0 * 0xC2B2AE3Du + (u32)(k0 >> 32) * 0x27D4EB2Fu) ^ ((u32)k1 * 0x165667B1u + (u32)(k1 >> 32) * 0x9E3779B1u);
            const u32 en = tab_addr + (u32)(k2TableBytes + k2KeepBytes) + __umulhi(h ^ (h >> 15), k2Second) * 32;
            if (second_probe(en, k0, k1)) {
                done = true;
                reds_add_addr(en + 16, 1u);
                reds_min_addr(en + 20, rel);
            }
        }
        const u64 end = tile_off + (u64)p; // buffer-relative offset of the delimiter that flushed the word
        const u64 first_abs = Lp->base + end;
#ifdef WC2_EXP_NOGLOBAL
        if (len == 0xFFFFFFF1u) { // (timing experiment: what the shared-memory tables do not take is dropped)
#else
        if (!done && len <= 16) { // the global word table: every read of the batch is issued before any is used
#endif
            const u32 slot = word_home(T, k0, len);
            const ulonglong2 key = __ldcg(&T.w_keys[slot]);
            const u64 finv = __ldcg(&T.w_finv[slot]);
            if (key.x == k0 && key.y == k1) { // the usual case: the word is in its home slot already
                red_add_u64(&T.w_counts[slot], 1ull);
                if (~first_abs > finv) red_max_u64(&T.w_finv[slot], ~first_abs);
            } else {
                Bucket b; b.k[0] = key;
#pragma unroll
                for (u32 i = 1; i < kBucket; i++) b.k[i] = __ldcg(&T.w_keys[slot + i]);
                if (word_insert_at(T, k0, k1, 1ull, first_abs, slot, b)) reds_add_addr(scr + 4 * kScrWordNew, 1u);
            }
        }
        // 17+ byte words on their own: a call in the middle of the loads above would serialise them
        if (__any_sync(0xFFFFFFFFu, !done && len > 16)) {
            if (!done && len > 16) {
                u64 wl = len;
                if (len >= kListLenMax) // (the list's length field saturates: find the start of such a word by scanning back)
                    while (wl < end && Lp->buf[end - wl - 1] != 0x20 && Lp->buf[end - wl - 1] != 0x0A && Lp->buf[end - wl - 1] != 0x0D) wl++;
                if (long_insert<Map>(Tp, Lp->buf, end - wl, wl, 1ull, first_abs)) reds_add_addr(scr + 4 * kScrLongNew, 1u);
            }
        }
        n_spill += !done && len <= 8; n_long += !done && len > 8;
    }
    // diagnostics: words counted in the global tables (1..8 bytes / 9+ bytes)
    n_spill = __reduce_add_sync(0xFFFFFFFFu, n_spill); n_long = __reduce_add_sync(0xFFFFFFFFu, n_long);
    if (lane == 0) {
        if (n_spill) reds_add_addr(scr + 4 * kScrSpill, n_spill);
        if (n_long) reds_add_addr(scr + 4 * kScrLong, n_long);
    }
}

template <class Map>
__global__ void __launch_bounds__(k2Threads, 1)
count_kernel2(const __grid_constant__ Launch L, const __grid_constant__ Tables T) {
    uint4 *tab = reinterpret_cast<uint4 *>(wc_smem);
    u64 *s_keep = reinterpret_cast<u64 *>(wc_smem + k2TableBytes); // s_keep[i] = the low i bytes set: the mask of a word's first min(len, 8) bytes
    uint4 *second = reinterpret_cast<uint4 *>(wc_smem + k2TableBytes + k2KeepBytes);
    unsigned char *warp_all = wc_smem + k2TableBytes + k2KeepBytes + k2SecondBytes;
    __shared__ u64 s_stop;

    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    Status *st = T.st;

    // (One thread reads STOP so that the whole CTA takes the same branch while other CTAs lower it.)
    if (tid == 0) s_stop = *reinterpret_cast<volatile u64 *>(&st->stop);
    if (tid < 9) s_keep[tid] = tid == 0 ? 0ull : ~0ull >> (8u * (8u - tid));
    __syncthreads();
    {
        const u64 stop = s_stop;
        if (stop != kNone && stop < L.base) { // the reference stopped reading before this piece (main.cu:185)
            if (L.ws != nullptr)
                for (u32 i = blockIdx.x * k2Warps + warp; lane == 0 && i < L.n_ws; i += gridDim.x * k2Warps) {
                    L.ws[i].p0 = kNone; L.ws[i].empties_before = 0; L.ws[i].empties_total = 0;
                }
            return;
        }
    }
    for (u32 i = tid; i < k2Entries; i += k2Threads) tab[i] = make_uint4(0, 0, 0, 0xFFFFFFFFu);
    for (u32 i = tid; i < k2Second * 2; i += k2Threads) second[i] = (i & 1) ? make_uint4(0, 0xFFFFFFFFu, 0, 0) : make_uint4(0, 0, 0, 0);

    const u64 cta_base = (u64)blockIdx.x * k2Warps * L.tiles_per_warp * kTile; // buffer-relative
    const u32 tab_addr = smem_addr(tab);
    const u32 keep_addr = tab_addr + (u32)k2TableBytes;
    u32 hmul = cta_hash_mul();
    asm volatile("" : "+r"(hmul)); // (kept in a register: recomputing it in the word loop costs three instructions per step)

    unsigned char *wbase = warp_all + warp * k2WarpSmemBytes;
    const u32 list_addr = smem_addr(wbase + kRingBytes);
    const u32 scr = list_addr + k2ListBytes;
    if (lane < k2ScratchBytes / 4) asm volatile("st.shared.u32 [%0], %1;" ::"r"(scr + lane * 4), "r"(0u) : "memory");
    __syncthreads();

    const u32 tiles_total = (u32)((L.n + kTile - 1) / kTile);
    const u32 gwarp = blockIdx.x * k2Warps + warp;
    const u32 t0 = gwarp * L.tiles_per_warp < tiles_total ? gwarp * L.tiles_per_warp : tiles_total;
    const u32 t1 = t0 + L.tiles_per_warp < tiles_total ? t0 + L.tiles_per_warp : tiles_total;
    const u32 lt_mask = (1u << lane) - 1;
    const int lane_pos = (int)(lane * 32);

    u32 acc_flush = 0, acc_empty = 0; // per lane; a warp's range is < 4 GiB
    bool seen_nonempty = false;

    if (t0 < t1) {
        // carry from tile to tile: position of the last delimiter (relative to the NEXT tile's start, so
        // always negative), the line state, and whether the previous byte was a line feed
        int carry_prev; bool killed, last_is_lf;
        if (t0 == 0) { carry_prev = -1; killed = false; last_is_lf = true; }
        else {
            const WarpCarry c = warp_lookback(L.buf, (u64)t0 * kTile);
            const long long cr = c.last_delim - (long long)t0 * kTile;
            carry_prev = cr < -(1ll << 30) ? -(1 << 30) : (int)cr;
            killed = c.killed; last_is_lf = c.last_is_lf;
            // the bytes just before the range, where a word ending in the first tile may start
            if (lane == 0)
                *reinterpret_cast<uint4 *>(wbase + (t0 & 1) * kBufBytes) = load16(L.buf, L.n, (u64)t0 * kTile - 16);
        }
        fetch_tile(L, t0, wbase + (t0 & 1) * kBufBytes, lane);
        const u32 last_tile = (u32)((L.n - 1) / kTile); // holds the piece's last byte
#pragma unroll 1
        for (u32 t = t0; t < t1; ++t) {
            unsigned char *buf = wbase + (t & 1) * kBufBytes;
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            if (t + 1 < t1) { // next tile into the other buffer, with this tile's last 16 bytes in front of it
                unsigned char *nbuf = wbase + ((t + 1) & 1) * kBufBytes;
                if (lane == 31) *reinterpret_cast<uint4 *>(nbuf) = *reinterpret_cast<const uint4 *>(buf + kTile);
                fetch_tile(L, t + 1, nbuf, lane);
            }
            const u64 T0 = (u64)t * kTile;          // buffer-relative offset of this tile
            const u32 relT = (u32)(T0 - cta_base);

            // ---- classify 2 x 16 bytes: one 32-bit mask per property
            u32 dm, lf, live;
            {
                u32 ex;
                {
                    const uint4 va = *reinterpret_cast<const uint4 *>(buf + 16 + lane * 32);
                    const FastFlags a0 = classify_fast4(va.x), a1 = classify_fast4(va.y), a2 = classify_fast4(va.z),
                                    a3 = classify_fast4(va.w);
                    dm = movemask16(a0.cand, a1.cand, a2.cand, a3.cand);
                    lf = movemask16(a0.lf, a1.lf, a2.lf, a3.lf);
                    ex = a0.ex | a1.ex | a2.ex | a3.ex;
                }
                {
                    const uint4 vb = *reinterpret_cast<const uint4 *>(buf + 32 + lane * 32);
                    const FastFlags b0 = classify_fast4(vb.x), b1 = classify_fast4(vb.y), b2 = classify_fast4(vb.z),
                                    b3 = classify_fast4(vb.w);
                    dm |= movemask16(b0.cand, b1.cand, b2.cand, b3.cand) << 16;
                    lf |= movemask16(b0.lf, b1.lf, b2.lf, b3.lf) << 16;
                    ex |= b0.ex | b1.ex | b2.ex | b3.ex;
                }
                live = dm;
                if (__any_sync(0xFFFFFFFFu, ex != 0) || killed) {
                    // a control byte in the tile (CRLF text, NUL, tab): the exact line state, SURVEY.md A.2.
                    // It sees the tile as two 512-byte halves of 16 bytes per lane.
                    if (lane == 0) reds_add_addr(scr + 4 * kScrSlow, 1u);
                    const uint4 xa = *reinterpret_cast<const uint4 *>(buf + 16 + lane * 16);
                    const uint4 xb = *reinterpret_cast<const uint4 *>(buf + 16 + kSub + lane * 16);
                    const ExactOut oa = exact_sub(xa, killed, L.buf, L.n, L.base, T0, &st->stop, 0);
                    const ExactOut ob = exact_sub(xb, oa.killed_out, L.buf, L.n, L.base, T0 + kSub, &st->stop, 0);
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
            {
                const u32 start = (lf << 1) | (prev & 1u); // bit p: byte p starts a line
                const u32 blank = lf & start;
                if (blank) atomicMin(&st->stop, L.base + T0 + lane * 32 + (__ffs(blank) - 1));
                if (t == last_tile && L.last_piece) {
                    const u32 q = (u32)(L.n - 1 - T0); // position of the last byte in this tile
                    if ((q >> 5) == lane && ((start >> (q & 31)) & 1)) atomicMin(&st->stop, L.base + L.n - 1);
                }
            }

            // ---- flush counts straight from the masks: a flush right after a delimiter is an empty word
            const u32 after = (dm << 1) | (prev >> 1);
            const u32 wm = live & ~after;            // bit p: the delimiter at byte p flushes a non-empty word
            acc_flush += __popc(live);
            acc_empty += __popc(live & after);
            if (!seen_nonempty) { // this warp's first non-empty word, and the empty ones before it
                const u32 em = live & after;
                const u32 bn = __ballot_sync(0xFFFFFFFFu, wm != 0);
                u32 x = __popc(em);
                if (bn) {
                    const u32 fl = __ffs(bn) - 1;
                    const u32 pf = __shfl_sync(0xFFFFFFFFu, wm ? __ffs(wm) - 1 : 0, fl);
                    if (lane == fl) x = __popc(em & ((1u << pf) - 1));
                    if (lane > fl) x = 0;
                    if (lane == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(scr + 4 * kScrP0), "r"(relT + fl * 32 + pf) : "memory");
                    seen_nonempty = true;
                }
                x = __reduce_add_sync(0xFFFFFFFFu, x);
                if (lane == 0 && x) reds_add_addr(scr + 4 * kScrEmptiesBefore, x);
            }

            // ---- the delimiter before this lane's first one: the last one of the nearest lane below that
            // has any, else the carry from the previous tiles (positions are tile relative)
            int q_in;
            {
                const int mylast = lane_pos + 31 - (int)__clz(dm);       // (garbage if dm == 0: never taken then)
                const u32 have = __ballot_sync(0xFFFFFFFFu, dm != 0);
                const u32 lower = have & lt_mask;
                const int got = __shfl_sync(0xFFFFFFFFu, mylast, (31 - (int)__clz(lower)) & 31);
                q_in = lower ? got : carry_prev;
                if (have) carry_prev = __shfl_sync(0xFFFFFFFFu, mylast, 31 - __clz(have)) - kTile;
                else carry_prev = carry_prev - kTile < -(1 << 30) ? -(1 << 30) : carry_prev - kTile;
            }

            // ---- two words per lane and step: the non-empty words that end in this lane's bytes, last one
            // first (two independent dependency chains per step: the warp is bound by latency, not by issue
            // slots).  The trip count comes from the largest number of words any lane holds: no vote needed.
            const u32 steps = (__reduce_max_sync(0xFFFFFFFFu, (u32)__popc(wm)) + 1) >> 1;
            const u32 tile_addr = smem_addr(buf) + 16; // shared-memory address of the tile's byte 0
            u32 rem = wm, missed = 0;                  // missed: the words this lane could not count (by delimiter bit)
#pragma unroll 1
            for (u32 it = 0; it < steps; ++it) {
                const WordProbe A = next_word<Map>(rem, dm, lane_pos, q_in, tile_addr, keep_addr, tab_addr, hmul, relT);
                const WordProbe B = next_word<Map>(rem, dm, lane_pos, q_in, tile_addr, keep_addr, tab_addr, hmul, relT);
                // single-probe main table: a hit is counted here ...
                uint4 ea, eb;
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(ea.x), "=r"(ea.y), "=r"(ea.z), "=r"(ea.w) : "r"(A.ent_addr) : "memory");
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(eb.x), "=r"(eb.y), "=r"(eb.z), "=r"(eb.w) : "r"(B.ent_addr) : "memory");
                const bool hit_a = A.word && A.len <= 8 && ea.x == A.klo && ea.y == A.khi;
                const bool hit_b = B.word && B.len <= 8 && eb.x == B.klo && eb.y == B.khi;
                if (hit_a) {
                    reds_add_addr(A.ent_addr + 8, 1u);
                    if (A.rel < ea.w) reds_min_addr(A.ent_addr + 12, A.rel);
                }
                if (hit_b) {
                    reds_add_addr(B.ent_addr + 8, 1u);
                    if (B.rel < eb.w) reds_min_addr(B.ent_addr + 12, B.rel);
                }
                // ... everything else (a miss, a free slot, a word of 9+ bytes) is left for the end of the tile
#ifdef WC2_EXP_NOSTAGE
                if (A.word && !hit_a && A.rel == 0xFFFFFFF1u) missed |= A.cb; // (timing experiment: misses are dropped)
#else
                if (A.word && !hit_a) missed |= A.cb;
                if (B.word && !hit_b) missed |= B.cb;
#endif
            }

            // ---- the words that missed: list them (one prefix sum over the lanes' counts), then settle them
            // one per lane.  A tile with more than k2ListCap of them is done one half (16 lanes) at a time.
            const u32 n_missed = __popc(missed);
            if (__any_sync(0xFFFFFFFFu, n_missed != 0)) {
                const u32 incl = warp_incl_scan(n_missed);
                const u32 tot = __shfl_sync(0xFFFFFFFFu, incl, 31), half0 = __shfl_sync(0xFFFFFFFFu, incl, 15);
                const bool dense = tot > k2ListCap;
                for (int ph = 0; ph < (dense ? 2 : 1); ph++) {
                    if (!dense || (ph == 0) == (lane < 16)) {
                        u32 at = list_addr + 4 * (incl - n_missed - ((dense && ph) ? half0 : 0u));
                        u32 m = missed;
                        while (m) {
                            const u32 c = 31 - __clz(m);
                            const u32 cb = 1u << c;
                            m &= ~cb;
                            const u32 below = dm & (cb - 1);
                            const int cur = lane_pos + (int)c;
                            const int nxt = below ? lane_pos + 31 - (int)__clz(below) : q_in;
                            const u32 len = (u32)(cur - nxt - 1);
                            asm volatile("st.shared.u32 [%0], %1;" ::"r"(at), "r"((u32)cur | ((len < kListLenMax ? len : kListLenMax) << 10)) : "memory");
                            at += 4;
                        }
                    }
                    __syncwarp();
                    settle_tile<Map>(list_addr, dense ? (ph ? tot - half0 : half0) : tot, tile_addr, relT, T0, &L, &T, tab_addr, hmul);
                    __syncwarp();
                }
            }
        }
    }
    // ---- per-warp results
    const u64 w_flush = warp_sum64(acc_flush), w_empty = warp_sum64(acc_empty);
    __syncwarp();
    if (lane == 0) {
        const u32 n_wnew = lds_u32_addr(scr + 4 * kScrWordNew), n_xnew = lds_u32_addr(scr + 4 * kScrLongNew),
                  n_spill = lds_u32_addr(scr + 4 * kScrSpill), n_long = lds_u32_addr(scr + 4 * kScrLong),
                  n_slow = lds_u32_addr(scr + 4 * kScrSlow);
        if (w_flush) atomicAdd(&st->n_flush, w_flush);
        if (w_empty) atomicAdd(&st->n_empty, w_empty);
        if (n_slow) atomicAdd(&st->n_slow_tiles, (u64)n_slow);
        if (n_wnew) red_add_u64(&st->w_entries, (u64)n_wnew);
        if (n_xnew) red_add_u64(&st->x_entries, (u64)n_xnew);
        if (n_spill) red_add_u64(&st->n_spill, (u64)n_spill);
        if (n_long) red_add_u64(&st->n_long, (u64)n_long);
        if (L.ws != nullptr && gwarp < L.n_ws) {
            L.ws[gwarp].p0 = (seen_nonempty && t0 < t1) ? L.base + cta_base + lds_u32_addr(scr + 4 * kScrP0) : kNone;
            L.ws[gwarp].empties_before = lds_u32_addr(scr + 4 * kScrEmptiesBefore);
            L.ws[gwarp].empties_total = w_empty;
        }
    }

    // ---- flush the CTA's tables into the global one
    __syncthreads();
    u32 made = 0;
    for (u32 i = tid; i < k2Entries; i += k2Threads) {
        const uint4 ent = tab[i];
        if ((ent.x | ent.y) != 0 && ent.z != 0) {
            const u64 key = (u64)ent.x | ((u64)ent.y << 32);
            made += word_insert(T, key, 0ull, key_len8(key), (u64)ent.z, L.base + cta_base + ent.w);
        }
    }
    for (u32 i = tid; i < k2Second; i += k2Threads) {
        const uint4 k = second[2 * i], v = second[2 * i + 1]; // {key}, {count, first, -, -}
        const u64 k0 = (u64)k.x | ((u64)k.y << 32), k1 = (u64)k.z | ((u64)k.w << 32);
        if (k0 != 0 && k0 != kSecondLock && v.x != 0)
            made += word_insert(T, k0, k1, k1 ? 8 + key_len8(k1) : key_len8(k0), (u64)v.x, L.base + cta_base + v.y);
    }
    made = __reduce_add_sync(0xFFFFFFFFu, made);
    if (lane == 0 && made) atomicAdd(&st->w_entries, (u64)made);
}

} // namespace wc
