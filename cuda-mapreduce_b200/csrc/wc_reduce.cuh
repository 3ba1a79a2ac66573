// wc_reduce.cuh -- the reduce semantics of /root/reference/main.cu:57-108 on the device (SURVEY.md
// 8(f) rank 1), for wc_finish.
//
// The reference's reducer walks the words in input order and credits each one to the FIRST output
// row of which it is a prefix (compare(), main.cu:57-67), appending a row when there is none.  In its
// order-free form: owner(w) = the distinct word with the smallest first-occurrence
// offset among the words that have w as a prefix; rows are the words that own themselves, in order of
// first occurrence; a row's count is the sum of the exact counts of what it owns.
//
// Input: the table entries as records (wc_shuffle.cuh), already ordered twice by radix sorts: by the
// word's first 16 bytes as a big-endian 128-bit number (`by_key`) and by first offset (`by_first`).
// In key order the words that have w as a prefix are one contiguous range that starts at w:
//   * a word of L <= 16 bytes: every key in [key(w), key(w) + 2^(8*(16-L))) -- the range's end is
//     one binary search away;
//   * a word of more than 16 bytes can only be a prefix of words that share its first 16 bytes: those
//     are compared byte by byte (they are few).
// owner(w) is then a range-minimum query over first offsets, answered by a two-level table of block
// minima (blocks of 256 and of 65536 entries) plus linear scans of the ragged ends: ranges are tiny
// for all but the shortest words.
#pragma once
#include "wc_shuffle.cuh"

namespace wc {

constexpr u32 kRmqBlock = 256;

struct ReduceArrays {
    // record index space: word record i -> i, long record j -> n_w_max + j
    const WordRecord *w; u64 n_w_max; const u64 *n_w_ptr;
    const LongRecord *x; u64 n_x_max; const u64 *n_x_ptr;
    const u64 *blob;
    const u32 *by_key, *by_first;   // n entries each
    u64 n;
    // sorted (key order) space, position s
    const u64 *s_hi;                // sorted high halves (output of the last radix sort)
    u64 *s_lo; u32 *s_len; u64 *s_first; u64 *s_count;
    u32 *pos_of_idx;                // record index -> s
    u32 *owner;                     // s -> s of its owner
    u64 *acc;                       // s -> sum of the counts it owns
    u64 *bmin1; u32 *bpos1; u64 *bmin2; u32 *bpos2;
    u64 *counters;                  // [0] valid entries, [1] rows, [2] sum of all counts, [3] bytes of the rows' words
};

__global__ void reduce_gather_kernel(ReduceArrays A, const u64 *lo_by_idx) {
    const u64 n_w = *A.n_w_ptr, n_x = *A.n_x_ptr;
    u32 valid = 0;
    u64 words = 0;
    for (u64 s = blockIdx.x * (u64)blockDim.x + threadIdx.x; s < A.n; s += (u64)gridDim.x * blockDim.x) {
        const u32 idx = A.by_key[s];
        A.pos_of_idx[idx] = (u32)s;
        A.s_lo[s] = lo_by_idx[idx];
        u32 len = 0; u64 first = ~0ull, cnt = 0;
        if (idx < A.n_w_max) {
            if (idx < n_w) {
                const WordRecord r = A.w[idx];
                cnt = r.count; first = r.first;
                len = r.k1 ? 8 + key_len8(r.k1) : (r.k0 ? key_len8(r.k0) : 0);
            }
        } else if (idx - A.n_w_max < n_x) {
            const LongRecord r = A.x[idx - A.n_w_max];
            cnt = r.count; first = r.first; len = (u32)r.len;
        }
        if (cnt == 0 || len == 0) { len = 0; first = ~0ull; cnt = 0; } // a slot without a record, or a rolled-back word
        valid += len != 0;
        words += cnt;
        A.s_len[s] = len; A.s_first[s] = first; A.s_count[s] = cnt; A.acc[s] = 0;
    }
    valid = __reduce_add_sync(0xFFFFFFFFu, valid);
    words = warp_sum64(words);
    if ((threadIdx.x & 31) == 0 && valid) { atomicAdd(&A.counters[0], (u64)valid); atomicAdd(&A.counters[2], words); }
}

// One block per kRmqBlock inputs: min value and where it is.  pos_in == nullptr: positions are the indices.
__global__ void reduce_blockmin_kernel(const u64 *v, const u32 *pos_in, u64 n, u64 *bmin, u32 *bpos) {
    __shared__ u64 sv[kRmqBlock];
    __shared__ u32 sp[kRmqBlock];
    for (u64 b = blockIdx.x; b * kRmqBlock < n; b += gridDim.x) {
        const u64 i = b * kRmqBlock + threadIdx.x;
        sv[threadIdx.x] = i < n ? v[i] : ~0ull;
        sp[threadIdx.x] = i < n ? (pos_in ? pos_in[i] : (u32)i) : 0xFFFFFFFFu;
        __syncthreads();
        for (u32 d = kRmqBlock / 2; d > 0; d >>= 1) {
            if (threadIdx.x < d && sv[threadIdx.x + d] < sv[threadIdx.x]) {
                sv[threadIdx.x] = sv[threadIdx.x + d]; sp[threadIdx.x] = sp[threadIdx.x + d];
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) { bmin[b] = sv[0]; bpos[b] = sp[0]; }
        __syncthreads();
    }
}

__device__ __forceinline__ bool key_less(u64 ahi, u64 alo, u64 bhi, u64 blo) { return ahi < bhi || (ahi == bhi && alo < blo); }

// first position in [0, n) whose key is >= (hi, lo)
__device__ __forceinline__ u64 key_lower_bound(const u64 *s_hi, const u64 *s_lo, u64 n, u64 hi, u64 lo) {
    u64 a = 0, b = n;
    while (a < b) {
        const u64 m = (a + b) >> 1;
        if (key_less(s_hi[m], s_lo[m], hi, lo)) a = m + 1; else b = m;
    }
    return a;
}

// argmin of s_first over [a, b), b > a; ties cannot happen between valid entries (offsets are distinct)
__device__ __forceinline__ u32 range_argmin(const ReduceArrays &A, u64 a, u64 b) {
    u64 best = ~0ull; u32 at = (u32)a;
    auto scan0 = [&](u64 lo, u64 hi) { for (u64 i = lo; i < hi; i++) { const u64 v = A.s_first[i]; if (v < best) { best = v; at = (u32)i; } } };
    auto scan1 = [&](u64 lo, u64 hi) { for (u64 i = lo; i < hi; i++) { const u64 v = A.bmin1[i]; if (v < best) { best = v; at = A.bpos1[i]; } } };
    auto scan2 = [&](u64 lo, u64 hi) { for (u64 i = lo; i < hi; i++) { const u64 v = A.bmin2[i]; if (v < best) { best = v; at = A.bpos2[i]; } } };
    if (b - a <= 2 * kRmqBlock) { scan0(a, b); return at; }
    const u64 a1 = (a + kRmqBlock - 1) / kRmqBlock, b1 = b / kRmqBlock; // whole level-1 blocks [a1, b1)
    scan0(a, a1 * kRmqBlock);
    scan0(b1 * kRmqBlock, b);
    if (b1 - a1 <= 2 * kRmqBlock) { scan1(a1, b1); return at; }
    const u64 a2 = (a1 + kRmqBlock - 1) / kRmqBlock, b2 = b1 / kRmqBlock; // whole level-2 blocks [a2, b2)
    scan1(a1, a2 * kRmqBlock);
    scan1(b2 * kRmqBlock, b1);
    scan2(a2, b2);
    return at;
}

__device__ __forceinline__ const u64 *long_bytes(const ReduceArrays &A, u64 s) { // of the long word at sorted position s
    const u32 idx = A.by_key[s];
    return A.blob + (A.x[idx - A.n_w_max].blob_off >> 3);
}

// is the long word at s (len bytes) a prefix of the long word at t (at least len bytes)?  Both share their first 16 bytes.
__device__ __forceinline__ bool long_is_prefix(const ReduceArrays &A, u64 s, u32 len, u64 t) {
    const u64 *a = long_bytes(A, s), *b = long_bytes(A, t);
    u32 i = 16;
    for (; i + 8 <= len; i += 8)
        if (a[i >> 3] != b[i >> 3]) return false;
    if (i < len) {
        const u64 m = (1ull << (8 * (len - i))) - 1;
        if ((a[i >> 3] ^ b[i >> 3]) & m) return false;
    }
    return true;
}

__global__ void reduce_owner_kernel(ReduceArrays A, int exact_mode) {
    for (u64 s0 = blockIdx.x * (u64)blockDim.x + (threadIdx.x & ~31u); s0 < A.n; s0 += (u64)gridDim.x * blockDim.x) {
        const u64 s = s0 + (threadIdx.x & 31);
        const bool in = s < A.n;
        const u32 len = in ? A.s_len[s] : 0;
        u32 own = (u32)s;
        if (len != 0 && !exact_mode) {
            const u64 hi = A.s_hi[s], lo = A.s_lo[s];
            // the first key past the words that start with this word's first min(len, 16) bytes
            u64 uh = hi, ul = lo;
            bool wrapped;
            if (len <= 8) { const u64 add = 1ull << (8 * (8 - len)); uh = hi + add; ul = 0; wrapped = uh < hi; }
            else { const u64 add = len >= 16 ? 1ull : 1ull << (8 * (16 - len)); ul = lo + add; wrapped = false;
                   if (ul < lo) { uh = hi + 1; wrapped = uh == 0; } }
            const u64 end = wrapped ? A.n : key_lower_bound(A.s_hi, A.s_lo, A.n, uh, ul);
            if (len <= 16) {
                // (a 16-byte word shares its key with longer words: start at the first entry of that key)
                const u64 start = len == 16 ? key_lower_bound(A.s_hi, A.s_lo, A.n, hi, lo) : s;
                own = range_argmin(A, start, end > s ? end : s + 1);
            } else { // same first 16 bytes: [g0, end); only longer words can continue this one
                const u64 g0 = key_lower_bound(A.s_hi, A.s_lo, A.n, hi, lo);
                u64 best = A.s_first[s];
                for (u64 t = g0; t < end; t++) {
                    if (t == s || A.s_len[t] < len || A.s_first[t] >= best) continue;
                    if (long_is_prefix(A, s, len, t)) { best = A.s_first[t]; own = (u32)t; }
                }
            }
        }
        if (in) A.owner[s] = own;
        // counts to the owners: lanes with the same owner add once
        const u64 cnt = (in && len != 0) ? A.s_count[s] : 0;
        const u32 key = (in && len != 0) ? own : 0xFFFFFFFFu;
        const u32 peers = __match_any_sync(0xFFFFFFFFu, key);
        u64 sum = 0;
        for (u32 m = peers; m; m &= m - 1) sum += __shfl_sync(peers, cnt, __ffs(m) - 1); // (same trip count within a peer group)
        if (key != 0xFFFFFFFFu && (u32)(__ffs(peers) - 1) == (threadIdx.x & 31)) atomicAdd(&A.acc[own], sum);
    }
}

// In order of first occurrence: is this entry a row?
__global__ void reduce_rowflags_kernel(ReduceArrays A, unsigned char *flags) {
    for (u64 r = blockIdx.x * (u64)blockDim.x + threadIdx.x; r < A.n; r += (u64)gridDim.x * blockDim.x) {
        const u32 s = A.pos_of_idx[A.by_first[r]];
        flags[r] = A.s_len[s] != 0 && A.owner[s] == s;
    }
}

// The result as the host wants it (layout of wc::Row in wc_finalize.h): rows in output order, their words packed
// back to back.  lens[r] feeds the prefix sum that places the words.
struct RowOut { u64 blob_off; u32 len; u32 pad; u64 count; u64 first; };

__global__ void reduce_rowlens_kernel(ReduceArrays A, const u32 *row_idx, const u64 *n_rows_ptr, u64 *lens) {
    const u64 n_rows = *n_rows_ptr;
    for (u64 r = blockIdx.x * (u64)blockDim.x + threadIdx.x; r < A.n; r += (u64)gridDim.x * blockDim.x)
        lens[r] = r < n_rows ? A.s_len[A.pos_of_idx[row_idx[r]]] : 0;
}

__global__ void reduce_emit_rows_kernel(ReduceArrays A, const u32 *row_idx, const u64 *n_rows_ptr, const u64 *offs, RowOut *rows,
                                        uint8_t *blob_out) {
    const u64 n_rows = *n_rows_ptr;
    for (u64 r = blockIdx.x * (u64)blockDim.x + threadIdx.x; r < n_rows; r += (u64)gridDim.x * blockDim.x) {
        const u32 idx = row_idx[r];
        const u32 s = A.pos_of_idx[idx];
        const u32 len = A.s_len[s];
        const u64 off = offs[r];
        rows[r] = RowOut{off, len, 0u, A.acc[s], A.s_first[s]};
        uint8_t *dst = blob_out + off;
        if (idx < A.n_w_max) {
            const WordRecord w = A.w[idx];
            for (u32 b = 0; b < len; b++) dst[b] = (uint8_t)((b < 8 ? w.k0 >> (8 * b) : w.k1 >> (8 * (b - 8))) & 0xFF);
        } else {
            const uint8_t *src = reinterpret_cast<const uint8_t *>(A.blob) + A.x[idx - A.n_w_max].blob_off;
            for (u32 b = 0; b < len; b++) dst[b] = src[b];
        }
        if (r == n_rows - 1) A.counters[3] = off + len;
    }
}

} // namespace wc
