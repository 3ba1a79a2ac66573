// wc_sort.cuh -- the radix-sort + segmented-reduce aggregation path (WC_PATH_SORT).
//
// No counterpart in the reference (its reduce is one thread's linear scan, main.cu:69-108).  The
// tokenizer writes one record per word (wc_device.cuh, EmitBuf); at the end of the job the records are
// sorted by a 64-bit hash of the word (cub::DeviceRadixSort), runs of EQUAL WORDS -- the 16-byte keys
// are compared, not the hashes, so two words that share a hash simply form separate runs -- are
// marked, numbered with a prefix sum, and reduced: count = run length, first = min offset.  Cost is
// independent of how many distinct words there are, which is what the hash path cannot offer when
// nearly every token is new.
#pragma once
#include "wc_device.cuh"

namespace wc {

// flags[i] = 1 iff sorted record i starts a run (its word differs from the previous record's)
// sort key = 64-bit hash of the 16-byte word; payload = the record's index
__global__ void sort_prepare_kernel(const ulonglong2 *key, u64 n, u64 *hash, u32 *idx) {
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const ulonglong2 k = key[i];
        hash[i] = mix64(k.x ^ mix64(k.y + 0x9E3779B97F4A7C15ull));
        idx[i] = (u32)i;
    }
}

// collisions: runs that start although the hash did not change (two different words, one hash) -- then
// the same word may form more than one run and the host merges them.
__global__ void sort_mark_heads_kernel(const u64 *hash, const u32 *idx, const ulonglong2 *key, u64 n, u32 *flags,
                                       u64 *collisions) {
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u32 head = 1;
        if (i > 0) {
            const ulonglong2 a = key[idx[i]], b = key[idx[i - 1]];
            head = (a.x != b.x) || (a.y != b.y);
            if (head && hash[i] == hash[i - 1]) atomicAdd(collisions, 1ull);
        }
        flags[i] = head;
    }
}

// run[i] = inclusive prefix sum of flags (1-based run number).  One (count, first) per run, with the
// partial sums of consecutive records of one run combined inside the warp before any atomic.
// Records flushed at or after `stop` (main.cu:185: the reference had stopped reading) count for nothing.
__global__ void sort_reduce_runs_kernel(const u32 *idx, const ulonglong2 *key, const u64 *end, const u32 *run, u64 n,
                                        u64 stop, ulonglong2 *out_key, u64 *out_count, u64 *out_first) {
    const u32 lane = threadIdx.x & 31;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    const u64 rounds = (n + stride - 1) / stride;
    for (u64 r = 0; r < rounds; r++) {
        const u64 i = r * stride + blockIdx.x * (u64)blockDim.x + threadIdx.x;
        const bool valid = i < n;
        const u32 my_run = valid ? run[i] : 0xFFFFFFFFu;
        const u32 rec = valid ? idx[i] : 0;
        u64 first = valid ? end[rec] : ~0ull;
        u32 count = valid ? 1u : 0u;
        if (first >= stop) { first = ~0ull; count = 0; }
        // segmented reduction over lanes holding the same run (records are sorted: such lanes are adjacent)
        const u32 same = __match_any_sync(0xFFFFFFFFu, my_run);
        const u32 leader = __ffs(same) - 1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u64 f2 = __shfl_down_sync(0xFFFFFFFFu, first, d);
            const u32 c2 = __shfl_down_sync(0xFFFFFFFFu, count, d);
            const u32 r2 = __shfl_down_sync(0xFFFFFFFFu, my_run, d);
            if (lane + d < 32 && r2 == my_run) { first = f2 < first ? f2 : first; count += c2; }
        }
        if (valid && lane == leader) {
            const u32 o = my_run - 1;
            atomicAdd(&out_count[o], (u64)count);
            atomicMin(&out_first[o], first);
            if (i == 0 || run[i - 1] != my_run) out_key[o] = key[rec]; // the run's first record names it
        }
    }
}

} // namespace wc
