// wc_jobs.cuh -- a second job through the same skeleton.
//
// The reference is nominally a MapReduce skeleton whose mapper() (/root/reference/main.cu:37-54) and
// reducer() (main.cu:69-108) a maintainer may swap for another job.  WC_JOB_LINE_LENGTHS is such a
// swap: map = every line of the input (what fgets delivers piece by piece, main.cu:179) to the key
// "its length in bytes, line feed excluded", reduce = count per key.  It runs on the tokenizer's front
// end -- the same byte-range split over warps, the same SWAR line-feed classification, the same
// look-back at the start of a warp's range and the same "previous one from the nearest lane below"
// carry as the word loop -- but its reducer is a histogram: a shared-memory array per CTA for the
// lengths below kLineBins, flushed with one atomic per used bin, and a list for the rare longer lines.
#pragma once
#include "wc_device.cuh"

namespace wc {

constexpr u32 kLineBins = 4096;       // lengths 0 .. kLineBins-1 are counted in bins
constexpr int kLineThreads = 512;

struct LineJob {
    u64 *hist;       // [kLineBins]
    u64 *ovf;        // lengths >= kLineBins, one entry per such line
    u64 ovf_cap;
    u64 *ovf_count;  // entries appended (may exceed ovf_cap: then err is set)
    u32 *err;
};

__device__ __forceinline__ u32 lf_flags4(u32 x) { // 0x80 in every byte that is 0x0A
    const u32 b = ((x & 0x7F7F7F7Fu) ^ 0x0A0A0A0Au) + 0x7F7F7F7Fu;
    return ~(b | x) & 0x80808080u;
}

__global__ void __launch_bounds__(kLineThreads, 2)
line_length_kernel(const uint8_t *buf, u64 n, u32 tiles_per_warp, int last_piece, LineJob J) {
    __shared__ u32 s_hist[kLineBins];
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (u32 i = tid; i < kLineBins; i += kLineThreads) s_hist[i] = 0;
    __syncthreads();
    const u32 tiles_total = (u32)((n + kTile - 1) / kTile);
    const u32 gwarp = blockIdx.x * (kLineThreads / 32) + warp;
    const u32 t0 = gwarp * tiles_per_warp < tiles_total ? gwarp * tiles_per_warp : tiles_total;
    const u32 t1 = t0 + tiles_per_warp < tiles_total ? t0 + tiles_per_warp : tiles_total;
    const u32 lt_mask = (1u << lane) - 1;
    auto record = [&](u64 len) {
        if (len < kLineBins) atomicAdd(&s_hist[(u32)len], 1u);
        else {
            const u64 at = atom_add_u64(J.ovf_count, 1ull);
            if (at < J.ovf_cap) J.ovf[at] = len; else atomicOr(J.err, kErrEmitFull);
        }
    };
    if (t0 < t1) {
        // buffer-relative position of the last line feed before the range (-1: the range starts the piece)
        long long prev_lf = -1;
        if (t0 != 0) prev_lf = warp_lookback(buf, (u64)t0 * kTile).line_start - 1;
        for (u32 t = t0; t < t1; ++t) {
            const u64 T0 = (u64)t * kTile, off = T0 + lane * 32;
            const uint4 va = load16(buf, n, off), vb = load16(buf, n, off + 16); // bytes at or past n read as 'x'
            const u32 lf = movemask16(lf_flags4(va.x), lf_flags4(va.y), lf_flags4(va.z), lf_flags4(va.w)) |
                           (movemask16(lf_flags4(vb.x), lf_flags4(vb.y), lf_flags4(vb.z), lf_flags4(vb.w)) << 16);
            // the line feed before this lane's first one: the last one of the nearest lane below that has any
            const int mylast = (int)(lane * 32 + 31) - (int)__clz(lf);
            const u32 have = __ballot_sync(0xFFFFFFFFu, lf != 0);
            const u32 lower = have & lt_mask;
            const int got = __shfl_sync(0xFFFFFFFFu, mylast, (31 - (int)__clz(lower)) & 31);
            long long q = lower ? (long long)T0 + got : prev_lf;
            for (u32 m = lf; m;) {
                const u32 p = __ffs(m) - 1; m &= m - 1;
                const long long pos = (long long)off + p;
                record((u64)(pos - q - 1));
                q = pos;
            }
            if (have) prev_lf = (long long)T0 + __shfl_sync(0xFFFFFFFFu, mylast, 31 - __clz(have));
        }
        // the end of the input without a final line feed: what is left is a line too (fgets returns it)
        if (last_piece && t1 == tiles_total && lane == 0 && (long long)n - 1 > prev_lf) record((u64)((long long)n - 1 - prev_lf));
    }
    __syncthreads();
    for (u32 i = tid; i < kLineBins; i += kLineThreads)
        if (s_hist[i]) red_add_u64(&J.hist[i], (u64)s_hist[i]);
}

} // namespace wc
