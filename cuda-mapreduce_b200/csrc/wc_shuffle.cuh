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
        made += long_insert<MapIdentity>(&T, blob, r.blob_off, r.len, r.count, r.first);
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

// ---- peer gather (one node, NVLink): every non-root GPU writes its table entries straight into a
// landing area in the ROOT GPU's memory, mapped into its address space through CUDA IPC.  No staging
// buffer, no size exchange: the stores travel over NVLink while the scan of the table is still going.
// The root's block: [PeerCtl] then, per sending rank, TWO areas [PeerHeader][word records][long records][blob]:
// job k lands in area k & 1, so that a sender that is one job ahead of the root never writes where the
// root is still reading; a sender that would be two ahead waits for PeerCtl::done_seq (the root stores
// the number of the last job it has finished importing there).
struct PeerCtl { u64 done_seq; u64 pad[31]; }; // 256 bytes
static_assert(sizeof(PeerCtl) == 256, "control block size");
struct PeerHeader {          // 256 bytes, written last (by the last block of the push kernel)
    u64 n_words, n_long, blob_bytes;
    u64 flags;               // bit 0: an area was too small; bits 8..: the sender's Status.err; bit 1: overflow log not empty
    u64 base, bytes, n_flush, n_empty, lost_empty, p0, stop, p0_found; // the sender's shard counters
    u64 seq;                 // the push number (the root checks it against its pull number)
    u64 pad[19];
};
static_assert(sizeof(PeerHeader) == 256, "header size");
struct PeerArea { PeerHeader *hdr; WordRecord *words; LongRecord *longs; u64 *blob; u64 cap_words, cap_long, cap_blob_units; };

// cur (local, zeroed): [0] word records, [1] long records, [2] blob bytes, [3] blocks done
__global__ void peer_push_kernel(Tables T, PeerArea A, const PeerCtl *ctl, u64 base, u64 bytes, u64 seq, u64 *cur,
                                 u32 *flag_out) {
    const u32 lane = threadIdx.x & 31;
    // flow control: area seq & 1 was last used by job seq - 2; the root must be done with that one
    if (seq > 2) {
        if (threadIdx.x == 0) {
            const volatile u64 *done = &ctl->done_seq; // in the root GPU's memory
            for (u32 spins = 0; *done + 2 < seq && spins < (1u << 24); ++spins) __nanosleep(200);
        }
        __syncthreads();
    }
    const u64 n_w = 1ull << T.w_log2, n_x = 1ull << T.x_log2;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i0 = (u64)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); i0 < n_w; i0 += stride) { // warp-uniform trip count
        const u64 i = i0 + lane;
        ulonglong2 k = make_ulonglong2(0, 0);
        u64 c = 0, finv = 0;
        if (i < n_w) { k = T.w_keys[i]; c = T.w_counts[i]; finv = T.w_finv[i]; }
        const bool used = k.x != 0 && c != 0;
        const u32 m = __ballot_sync(0xFFFFFFFFu, used);
        if (m == 0) continue;
        u64 at = 0;
        if (lane == 0) at = atomicAdd(&cur[0], (u64)__popc(m));
        at = __shfl_sync(0xFFFFFFFFu, at, 0) + __popc(m & ((1u << lane) - 1));
        if (used && at < A.cap_words) {
            ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(&A.words[at]); // two 16-byte stores over NVLink
            dst[0] = k;
            dst[1] = make_ulonglong2(c, ~finv);
        }
    }
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n_x; i += stride) {
        const u64 r = T.x_refs[i];
        const u64 c = T.x_counts[i];
        if (r == 0 || c == 0) continue;
        const u64 unit = r & ((1ull << 40) - 1);
        const u64 len = T.arena[unit], units = (len + 7) >> 3;
        const u64 at = atomicAdd(&cur[1], 1ull);
        const u64 off = atomicAdd(&cur[2], units * 8);
        if (at < A.cap_long && (off >> 3) + units <= A.cap_blob_units) {
            A.longs[at] = LongRecord{len, c, ~T.x_finv[i], off};
            for (u64 j = 0; j < units; j++) A.blob[(off >> 3) + j] = T.arena[unit + 1 + j];
        }
    }
    // the header goes last, after every block's records are on their way
    __threadfence_system();
    __shared__ u32 s_last;
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&cur[3], 1ull) == (u64)gridDim.x - 1;
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        __threadfence();
        const u64 nw = atomicAdd(&cur[0], 0ull), nx = atomicAdd(&cur[1], 0ull), nb = atomicAdd(&cur[2], 0ull);
        const Status *st = T.st;
        u64 flags = (u64)st->err << 8;
        if (nw > A.cap_words || nx > A.cap_long || (nb >> 3) > A.cap_blob_units) flags |= 1;
        if (st->ovf_count != 0) flags |= 2;
        PeerHeader h;
        h.n_words = nw; h.n_long = nx; h.blob_bytes = nb; h.flags = flags;
        h.base = base; h.bytes = bytes; h.n_flush = st->n_flush; h.n_empty = st->n_empty; h.lost_empty = st->lost_empty;
        h.p0 = st->p0; h.stop = st->stop; h.p0_found = st->p0_found; h.seq = seq;
        for (int q = 0; q < 19; q++) h.pad[q] = 0;
        *A.hdr = h;
        if (flag_out != nullptr) *flag_out = flags != 0;
        __threadfence_system();
    }
}

// The root's side: every live sender's records into the root's tables, in one launch.
struct ImportList {
    const WordRecord *w[64]; u64 n_w[64];
    const LongRecord *x[64]; u64 n_x[64];
    const uint8_t *blob[64];
    u32 n;
};
__global__ void import_areas_kernel(Tables T, const __grid_constant__ ImportList L) {
    u32 made_w = 0, made_x = 0;
    const u64 tid = blockIdx.x * (u64)blockDim.x + threadIdx.x, stride = (u64)gridDim.x * blockDim.x;
    for (u32 a = 0; a < L.n; a++) {
        for (u64 i = tid; i < L.n_w[a]; i += stride) {
            const WordRecord r = L.w[a][i];
            if (r.k0 == 0 || r.count == 0) continue;
            const u32 len = r.k1 ? 8 + key_len8(r.k1) : key_len8(r.k0);
            made_w += word_insert(T, r.k0, r.k1, len, r.count, r.first);
        }
        for (u64 i = tid; i < L.n_x[a]; i += stride) {
            const LongRecord r = L.x[a][i];
            if (r.len == 0 || r.count == 0) continue;
            made_x += long_insert<MapIdentity>(&T, L.blob[a], r.blob_off, r.len, r.count, r.first);
        }
    }
    made_w = __reduce_add_sync(0xFFFFFFFFu, made_w);
    made_x = __reduce_add_sync(0xFFFFFFFFu, made_x);
    if ((threadIdx.x & 31) == 0) {
        if (made_w) atomicAdd(&T.st->w_entries, (u64)made_w);
        if (made_x) atomicAdd(&T.st->x_entries, (u64)made_x);
    }
}

// ---- the root's side without a host round trip (wc_peer_pull_device): one small kernel waits until
// every sender's header of job `seq` has arrived in this GPU's memory, reduces the shard counters the way
// wc_shard_stats_reduce() does on the host (main.cu:185: a shard at or after the first STOP contributes
// nothing; main.cu:81,103: empty words count as lost until the job's first non-empty word), installs the
// job's counters and leaves, in device memory, the list of areas to import.
struct PeerJob {
    u32 n_ranks, root_rank, root_dead, failed;
    u64 n_w[64], n_x[64]; // per rank: records to import (0: nothing, dead shard, or the root itself)
};

__global__ void peer_wait_reduce_kernel(uint8_t *area0, u64 stride, u32 n_ranks, u32 root_rank, u64 seq, Status *st,
                                        u64 root_base, u64 root_bytes, PeerJob *job) {
    __shared__ u32 s_failed;
    const u32 r = threadIdx.x;
    if (r == 0) s_failed = 0;
    __syncthreads();
    if (r < n_ranks && r != root_rank) {
        const volatile PeerHeader *h = reinterpret_cast<const volatile PeerHeader *>(area0 + (2ull * r + (seq & 1)) * stride);
        u32 spins = 0;
        while (h->seq != seq && ++spins < (1u << 24)) __nanosleep(100);
        if (h->seq != seq || h->flags != 0) atomicOr(&s_failed, 1u); // never arrived, or the sender overflowed
    }
    __threadfence_system();
    __syncthreads();
    if (r != 0) return;
    // the shard counters in input order (the ranks' shards tile the input in rank order)
    u64 n_flush = 0, n_empty = 0, lost = 0, p0 = ~0ull, stop = ~0ull, bytes = 0, next_base = 0;
    u32 p0_found = 0;
    job->n_ranks = n_ranks; job->root_rank = root_rank; job->root_dead = 0; job->failed = s_failed;
    for (u32 i = 0; i < n_ranks; i++) {
        u64 s_base, s_bytes, s_flush, s_empty, s_lost, s_p0, s_stop, s_nw = 0, s_nx = 0; u32 s_p0f;
        if (i == root_rank) {
            s_base = root_base; s_bytes = root_bytes; s_flush = st->n_flush; s_empty = st->n_empty; s_lost = st->lost_empty;
            s_p0 = st->p0; s_stop = st->stop; s_p0f = st->p0_found;
        } else {
            const PeerHeader *h = reinterpret_cast<const PeerHeader *>(area0 + (2ull * i + (seq & 1)) * stride);
            s_base = h->base; s_bytes = h->bytes; s_flush = h->n_flush; s_empty = h->n_empty; s_lost = h->lost_empty;
            s_p0 = h->p0; s_stop = h->stop; s_p0f = (u32)h->p0_found; s_nw = h->n_words; s_nx = h->n_long;
        }
        if (s_base != next_base && i > 0) { atomicOr(&st->err, kErrPeer); } // the ranks' shards must tile the input
        next_base = s_base + s_bytes;
        bytes += s_bytes;
        const bool dead = stop != ~0ull; // a shard after the job's STOP (shards come in input order)
        if (dead && s_bytes > 0) { if (i == root_rank) job->root_dead = 1; s_nw = 0; s_nx = 0; }
        job->n_w[i] = s_failed ? 0 : s_nw; job->n_x[i] = s_failed ? 0 : s_nx;
        if (dead) continue;
        n_flush += s_flush; n_empty += s_empty;
        if (!p0_found) {
            if (s_p0f && s_p0 < s_stop) { p0_found = 1; p0 = s_p0; lost += s_lost; }
            else lost += s_empty;
        }
        if (s_stop != ~0ull) stop = s_stop;
    }
    st->own_words = st->n_flush - st->n_empty;
    st->n_flush = n_flush; st->n_empty = n_empty; st->lost_empty = lost; st->p0 = p0; st->stop = stop; st->p0_found = p0_found;
    st->job_bytes = bytes; st->job_set = 1;
    if (s_failed) atomicOr(&st->err, kErrPeer);
}

// The root's own shard lies after the job's STOP: what it counted does not belong to the job.
__global__ void peer_clear_root_kernel(const PeerJob *job, u64 *table_block, u64 n_words, Status *st) {
    if (!job->root_dead) return;
    for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < n_words; i += (u64)gridDim.x * blockDim.x) table_block[i] = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) { st->w_entries = 0; st->x_entries = 0; st->arena_used = 1; st->ovf_count = 0; }
}

__global__ void import_areas_dev_kernel(Tables T, const PeerJob *job, uint8_t *area0, u64 stride, u64 seq, u64 cap_words,
                                        u64 cap_long) {
    u32 made_w = 0, made_x = 0;
    const u64 tid = blockIdx.x * (u64)blockDim.x + threadIdx.x, nthr = (u64)gridDim.x * blockDim.x;
    for (u32 a = 0; a < job->n_ranks; a++) {
        const u64 n_w = job->n_w[a], n_x = job->n_x[a];
        if (n_w == 0 && n_x == 0) continue;
        const uint8_t *ar = area0 + (2ull * a + (seq & 1)) * stride;
        const WordRecord *w = reinterpret_cast<const WordRecord *>(ar + sizeof(PeerHeader));
        const LongRecord *x = reinterpret_cast<const LongRecord *>(ar + sizeof(PeerHeader) + cap_words * sizeof(WordRecord));
        const uint8_t *blob = ar + sizeof(PeerHeader) + cap_words * sizeof(WordRecord) + cap_long * sizeof(LongRecord);
        // four records per thread and step: their home slots are read before any of them is used (the import is
        // bound by the latency of those reads, not by arithmetic)
        for (u64 i0 = tid; i0 < n_w; i0 += 4 * nthr) {
            WordRecord r[4]; u32 slot[4]; Bucket b[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const u64 i = i0 + (u64)j * nthr;
                r[j] = i < n_w ? w[i] : WordRecord{0, 0, 0, 0};
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const u32 len = r[j].k1 ? 8 + key_len8(r[j].k1) : key_len8(r[j].k0 | 1);
                slot[j] = word_home(T, r[j].k0, len);
                b[j] = load_bucket(T, slot[j]);
            }
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (r[j].k0 != 0 && r[j].count != 0) made_w += word_insert_at(T, r[j].k0, r[j].k1, r[j].count, r[j].first, slot[j], b[j]);
        }
        for (u64 i = tid; i < n_x; i += nthr) {
            const LongRecord r = x[i];
            if (r.len == 0 || r.count == 0) continue;
            made_x += long_insert<MapIdentity>(&T, blob, r.blob_off, r.len, r.count, r.first);
        }
    }
    made_w = __reduce_add_sync(0xFFFFFFFFu, made_w);
    made_x = __reduce_add_sync(0xFFFFFFFFu, made_x);
    if ((threadIdx.x & 31) == 0) {
        if (made_w) atomicAdd(&T.st->w_entries, (u64)made_w);
        if (made_x) atomicAdd(&T.st->x_entries, (u64)made_x);
    }
}

// The root is done with job `seq`: its areas may be written again (read by the senders' push kernels).
__global__ void peer_done_kernel(PeerCtl *ctl, u64 seq) {
    ctl->done_seq = seq;
    __threadfence_system();
}

// Installs the job-wide counters without touching what the import kernels are updating.
__global__ void set_job_counters_kernel(Status *st, u64 n_flush, u64 n_empty, u64 lost_empty, u64 p0, u64 stop, u32 p0_found) {
    st->n_flush = n_flush; st->n_empty = n_empty; st->lost_empty = lost_empty; st->p0 = p0; st->stop = stop;
    st->p0_found = p0_found;
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
