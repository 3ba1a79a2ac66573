// wc_shuffle.cuh -- kernels of the multi-GPU exchange step.
//
// The reference has one GPU and no exchange of any kind (/root/reference/main.cu:133-162 is its whole
// driver); sharding is new here.  Every GPU counts its byte range into its own tables; then table
// entries are repartitioned by KEY OWNER so that each distinct word is reduced on exactly one GPU:
//   export : scan the word table, write 32-byte records {key0, key1, count, first} grouped by
//            owner = hash(key) * n_parts >> 32 into one send buffer (two passes: count, scatter);
//   (NCCL all-to-all of the send buffers -- done by the caller)
//   import : add received records into the owner's table (counts sum, first offsets take the min).
// Words of 17+ bytes live in the arena-backed long table; they are few and travel as
// {length, count, first, blob offset} + a byte blob straight to the gathering rank.
#pragma once
#include "wc_device.cuh"

namespace wc {

struct WordRecord { u64 k0, k1, count, first; };           // 32 bytes
struct LongRecord { u64 len, count, first, blob_off; };    // 32 bytes; bytes at blob + blob_off, padded to 8

__host__ __device__ __forceinline__ u32 key_owner(u64 k0, u64 k1, u32 n_parts) {
    const u32 h = hash_key128((u32)k0, (u32)(k0 >> 32), (u32)k1, (u32)(k1 >> 32));
    return (u32)(((u64)(h ^ (h >> 16) * 0x45D9F3Bu) * n_parts) >> 32);
}

__global__ void export_count_kernel(Tables T, u32 n_parts, u64 *counts) {
    const u64 n = 1ull << T.w_log2;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const ulonglong2 k = T.w_keys[i];
        if (k.x == 0 || T.w_counts[i] == 0) continue;
        atomicAdd(&counts[key_owner(k.x, k.y, n_parts)], 1ull);
    }
}

// offsets[p] = first record index of partition p in `out`; cursors[p] starts at 0
__global__ void export_scatter_kernel(Tables T, u32 n_parts, const u64 *offsets, u64 *cursors, WordRecord *out) {
    const u64 n = 1ull << T.w_log2;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const ulonglong2 k = T.w_keys[i];
        const u64 c = T.w_counts[i];
        if (k.x == 0 || c == 0) continue;
        const u32 p = key_owner(k.x, k.y, n_parts);
        const u64 at = offsets[p] + atomicAdd(&cursors[p], 1ull);
        out[at] = WordRecord{k.x, k.y, c, ~T.w_finv[i]};
    }
}

__global__ void import_words_kernel(Tables T, const WordRecord *in, u64 n) {
    u32 made = 0;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const WordRecord r = in[i];
        if (r.k0 == 0 || r.count == 0) continue;
        const u32 len = r.k1 ? 8 + key_len8(r.k1) : key_len8(r.k0);
        made += word_insert(T, r.k0, r.k1, len, r.count, r.first);
    }
    made = __reduce_add_sync(0xFFFFFFFFu, made);
    if ((threadIdx.x & 31) == 0 && made) atomicAdd(&T.st->w_entries, (u64)made);
}

// cursors[0] = entries written, cursors[1] = blob bytes used
__global__ void export_long_kernel(Tables T, u64 *cursors, LongRecord *meta, u64 *blob) {
    const u64 n = 1ull << T.x_log2;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 r = T.x_refs[i];
        const u64 c = T.x_counts[i];
        if (r == 0 || c == 0) continue;
        const u64 unit = r & ((1ull << 40) - 1);
        const u64 len = T.arena[unit], units = (len + 7) >> 3;
        const u64 at = atomicAdd(&cursors[0], 1ull);
        const u64 off = atomicAdd(&cursors[1], units * 8);
        meta[at] = LongRecord{len, c, ~T.x_finv[i], off};
        for (u64 j = 0; j < units; j++) blob[(off >> 3) + j] = T.arena[unit + 1 + j];
    }
}

__global__ void import_long_kernel(Tables T, const LongRecord *meta, u64 n, const uint8_t *blob) {
    u32 made = 0;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const LongRecord r = meta[i];
        if (r.len == 0 || r.count == 0) continue;
        made += long_insert(&T, blob, r.blob_off, r.len, r.count, r.first);
    }
    made = __reduce_add_sync(0xFFFFFFFFu, made);
    if ((threadIdx.x & 31) == 0 && made) atomicAdd(&T.st->x_entries, (u64)made);
}

// Table growth: every entry of the old word table into the new one (T already points at the new arrays).
__global__ void rehash_words_kernel(Tables T, const ulonglong2 *old_keys, const u64 *old_counts, const u64 *old_finv,
                                    u64 n_old) {
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n_old; i += (u64)gridDim.x * blockDim.x) {
        const ulonglong2 k = old_keys[i];
        if (k.x == 0) continue;
        const u32 len = k.y ? 8 + key_len8(k.y) : key_len8(k.x);
        word_insert(T, k.x, k.y, len, old_counts[i], ~old_finv[i]);
    }
}

// wc_finish: sort keys for the exported records, so that the host reduce starts from sorted lists.
// Record index space: word record i -> i, long record j -> n_w_max + j.  hi/lo = the word's first 16
// bytes as two big-endian integers (zero padded), fst = its first offset.  Slots past the number of
// records actually written get all-ones keys (the host skips them by index).
__device__ __forceinline__ u64 bswap64(u64 v) {
    const u32 a = (u32)v, b = (u32)(v >> 32);
    return ((u64)__byte_perm(a, 0, 0x0123) << 32) | (u64)__byte_perm(b, 0, 0x0123);
}

__global__ void order_keys_kernel(const WordRecord *w, u64 n_w_max, const u64 *n_w_ptr, const LongRecord *x, u64 n_x_max,
                                  const u64 *n_x_ptr, const u64 *blob, u64 *hi, u64 *lo, u64 *fst, u32 *idx) {
    const u64 n_w = *n_w_ptr, n_x = *n_x_ptr, n = n_w_max + n_x_max;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 h = ~0ull, l = ~0ull, f = ~0ull;
        if (i < n_w_max) {
            if (i < n_w) { h = bswap64(w[i].k0); l = bswap64(w[i].k1); f = w[i].first; }
        } else if (i - n_w_max < n_x) {
            const LongRecord r = x[i - n_w_max];
            h = bswap64(blob[r.blob_off >> 3]); l = bswap64(blob[(r.blob_off >> 3) + 1]); f = r.first;
        }
        hi[i] = h; lo[i] = l; fst[i] = f; idx[i] = (u32)i;
    }
}

__global__ void gather_u64_kernel(const u64 *src, const u32 *idx, u64 n, u64 *dst) {
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) dst[i] = src[idx[i]];
}

__global__ void count_long_kernel(Tables T, u64 *out) { // out[0] entries, out[1] blob bytes (padded to 8 each)
    const u64 n = 1ull << T.x_log2;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 r = T.x_refs[i];
        if (r == 0 || T.x_counts[i] == 0) continue;
        const u64 len = T.arena[r & ((1ull << 40) - 1)];
        atomicAdd(&out[0], 1ull);
        atomicAdd(&out[1], ((len + 7) >> 3) * 8);
    }
}

} // namespace wc
