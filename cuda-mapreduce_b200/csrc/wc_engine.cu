// wc_engine.cu -- host side of libwc.so: the C ABI of include/wc.h over the kernels of
// wc_device.cuh.  Replaces the reference's driver runMapReduce() (/root/reference/main.cu:133-162:
// cudaMalloc / blocking cudaMemcpy / two <<<100,100>>> launches / cudaFree, with no error checks)
// by an engine that keeps its device tables alive, streams host text through pinned, double-buffered
// cudaMemcpyAsync chunks, and checks every CUDA call.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/wc.h"
#include "wc_device.cuh"
#include "wc_count2.cuh"
#include "wc_jobs.cuh"
#include "wc_shuffle.cuh"
#include "wc_sort.cuh"
#include <cub/cub.cuh>
#include <chrono>
#include <mutex>
#include <unordered_map>
#include "wc_finalize.h"
#include "wc_reduce.cuh"
#include <nvtx3/nvToolsExt.h> // header-only (v3): a no-op unless a tool (nsys, ncu --nvtx) injects itself

namespace {
// One NVTX range per ABI call and per phase of wc_finish, so that a timeline shows the job's structure
// (tracing).  Costs two predictable branches when no tool is attached.
struct Range {
    explicit Range(const char *name) { nvtxRangePushA(name); }
    ~Range() { nvtxRangePop(); }
    Range(const Range &) = delete;
    Range &operator=(const Range &) = delete;
};
} // namespace

using wc::u32;
using wc::u64;

struct wc_engine {
    wc_config cfg{};
    int device = 0;
    int num_sms = 0;
    int grid = 0;
    cudaStream_t stream = nullptr;      // kernels
    cudaStream_t copy_stream = nullptr; // host -> device chunks
    std::string err;

    // device tables: one allocation, so that a reset is a single memset
    void *table_block = nullptr;
    size_t table_block_bytes = 0;
    wc::Tables T{};
    wc::Status *d_status = nullptr;
    wc::WarpSummary *d_ws = nullptr;
    u32 n_ws = 0;
    wc::Status *h_status = nullptr; // pinned
    wc::Status *h_prev = nullptr;   // pinned: the counters of a job that ended in wc_reset (poll_job_snapshot)
    cudaEvent_t ev_prev = nullptr;
    bool prev_pending = false;

    // streaming (allocated on first wc_feed)
    size_t chunk_cap = 0;
    uint8_t *d_chunk[2] = {nullptr, nullptr};
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_counted[2] = {nullptr, nullptr};
    int ring = 0;
    std::vector<uint8_t> carry; // bytes after the last '\n' seen so far (an unfinished line)
    uint8_t *h_stage[2] = {nullptr, nullptr}; // pinned staging for callers that feed pageable memory (allocated on demand)

    // sort path (WC_PATH_SORT): one record per word of 1..16 bytes, sorted and reduced in wc_finish
    bool sort_mode = false;
    // WC_PATH_AUTO: the path is chosen from the share of distinct words -- of a probe over the head of an
    // engine's first job, then of every finished job (the next job of the same engine follows it)
    bool auto_path = false, auto_decided = false;
    double auto_ratio = -1.0;   // distinct words / words last seen (probe or finished job); < 0: none yet
    u32 producers = wc::kProducers; // count_kernel's warp split for the next launch (note_consumer_load)
    u32 last_tokenizing_warps = 0;  // what the job's last forward launch ran with (wc_stats.tokenizing_warps)
    wc::EmitBuf em{};           // device buffers; em.cursor is always allocated in sort mode
    u64 em_bound = 0;           // upper bound of records written so far (pieces launched)
    ulonglong2 *so_key = nullptr; u64 *so_count = nullptr, *so_first = nullptr; // wc_aggregate()'s output
    u32 so_runs = 0; u64 so_coll = 0, so_words = 0; bool so_valid = false;
    u64 *d_ovf = nullptr;       // overflow log (wc::Tables::ovf_log)
    u64 ovf_cap = 0;
    u64 *d_cursors = nullptr;   // shuffle: per-partition counts / cursors (2 * kMaxParts + 2 words)
    void *d_fin = nullptr, *h_fin = nullptr; // wc_finish: the used table entries, compacted (device / pinned host)
    size_t d_fin_cap = 0, h_fin_cap = 0;
    u64 finish_d2h = 0;                      // bytes the last wc_finish brought to the host

    // peer gather (wc_peer_*): the root owns `area`; the others hold a mapping of it
    struct Peer {
        uint8_t *area = nullptr;   // root: cudaMalloc'd landing areas, one per rank
        uint8_t *mapped = nullptr; // sender: the root's area in this process (IPC mapping or the same pointer)
        bool ipc = false;
        u32 n_ranks = 0, root_rank = 0, my_rank = 0;
        u64 cap_words = 0, cap_long = 0, cap_blob = 0, stride = 0;
        u64 seq = 0;               // pushes made / pulls made
        u64 *d_cur = nullptr;      // sender: the push kernel's cursors
        wc::PeerHeader *h_hdr = nullptr; // root: pinned copy of the headers
        wc::PeerJob *d_job = nullptr;    // root: what wc_peer_pull_device's first kernel leaves for the import
    } peer;
    void *d_ord = nullptr;                   // wc_finish: sort keys and cub scratch
    size_t d_ord_cap = 0;
    void *d_fz = nullptr;                    // wc_finish: arrays of the device reduce (wc_reduce.cuh)
    size_t d_fz_cap = 0;
    void *d_out = nullptr, *h_out = nullptr; // wc_finish: the finished rows and their words (device / pinned host)
    size_t d_out_cap = 0, h_out_cap = 0;
    bool job_stats_set = false; // wc_shard_stats_set(): the counters describe the whole job, not this shard
    u64 job_bytes = 0;

    // WC_JOB_LINE_LENGTHS (wc_jobs.cuh): histogram bins, the list of lines longer than the bins, its cursor, error flag
    wc::LineJob line{};
    void *line_block = nullptr; size_t line_block_bytes = 0;

    u64 consumed = 0; // bytes handed to the device so far
    u64 fed = 0;      // bytes given to wc_feed / wc_count_device
    u64 launches = 0;
    u64 pieces = 0;      // pieces launched since the last reset
    bool settle_due = false; // the table check for the last launched piece is still to be done
    u64 grows = 0;

    // timing: per piece, events before the count kernel, after it, and after its roll-back companion
    struct EvTriple { cudaEvent_t a, mid, b; };
    std::vector<cudaEvent_t> ev_pool;
    std::vector<EvTriple> ev_used;
    double device_ms = 0;       // a -> b
    double count_kernel_ms = 0; // a -> mid: the tokenizer/count kernel alone
};

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t ck_e = (call);                                                                       \
        if (ck_e != cudaSuccess) {                                                                       \
            e->err = std::string(#call) + ": " + cudaGetErrorString(ck_e);                              \
            return ck_e == cudaErrorMemoryAllocation ? WC_ERR_NOMEM : WC_ERR_CUDA;                       \
        }                                                                                                \
    } while (0)

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
namespace {
constexpr u32 kMaxParts = 64;
constexpr u32 kMaxTokenWarps = wc::k2Warps > wc::kProducers ? wc::k2Warps : wc::kProducers; // per CTA, over both count kernels
}
static int ensure_cursors(wc_engine *e);

// ----------------------------------------------------------------- tables

static int alloc_tables(wc_engine *e, u32 s_log2, u32 l_log2, u64 arena_bytes, wc::Tables *out, void **block,
                        size_t *block_bytes) {
    // word table (1..16 bytes): 16-byte keys | counts | ~first;  long table (17+): refs | counts | ~first
    const size_t s_n = (size_t)1 << s_log2, l_n = (size_t)1 << l_log2;
    const size_t bytes = (4 * s_n + 3 * l_n) * sizeof(u64);
    void *p = nullptr;
    CK(cudaMalloc(&p, bytes));
    u64 *q = (u64 *)p;
    wc::Tables T = e->T;
    T.w_keys = (ulonglong2 *)q; T.w_counts = q + 2 * s_n; T.w_finv = q + 3 * s_n;
    q += 4 * s_n;
    T.x_refs = q; T.x_counts = q + l_n; T.x_finv = q + 2 * l_n;
    T.w_log2 = s_log2; T.x_log2 = l_log2;
    if (T.arena == nullptr) {
        const u64 units = arena_bytes / 8;
        CK(cudaMalloc((void **)&T.arena, units * 8));
        T.arena_units = units;
    }
    *out = T; *block = p; *block_bytes = bytes;
    return WC_OK;
}

static int reset_device_state(wc_engine *e) {
    CK(cudaMemsetAsync(e->table_block, 0, e->table_block_bytes, e->stream));
    wc::Status s;
    memset(&s, 0, sizeof(s));
    s.stop = wc::kNone; s.p0 = wc::kNone; s.arena_used = 1;
    *e->h_status = s;
    CK(cudaMemcpyAsync(e->d_status, e->h_status, sizeof(s), cudaMemcpyHostToDevice, e->stream));
    if (e->em.cursor) CK(cudaMemsetAsync(e->em.cursor, 0, sizeof(u64), e->stream));
    if (e->line_block) CK(cudaMemsetAsync(e->line_block, 0, (wc::kLineBins + 2) * sizeof(u64), e->stream));
    e->em_bound = 0;
    e->so_valid = false;
    return WC_OK;
}

static void poll_job_snapshot(wc_engine *e);

// ----------------------------------------------------------------- C ABI

extern "C" {

void wc_config_default(wc_config *cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->struct_size = (uint32_t)sizeof(*cfg);
    cfg->device = 0;
    cfg->mode = WC_MODE_REF;
    cfg->path = WC_PATH_AUTO;
    cfg->chunk_bytes = 64ull << 20;
    cfg->short_table_log2 = 21;
    cfg->long_table_log2 = 18;
    cfg->arena_bytes = 256ull << 20;
}

static thread_local std::string g_create_err; // why the last wc_create on this thread failed

const char *wc_last_error(const wc_engine *e) { return e ? e->err.c_str() : g_create_err.c_str(); }

void wc_destroy(wc_engine *e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->copy_stream) cudaStreamSynchronize(e->copy_stream);
    for (auto ev : e->ev_pool) cudaEventDestroy(ev);
    for (auto &p : e->ev_used) { cudaEventDestroy(p.a); cudaEventDestroy(p.mid); cudaEventDestroy(p.b); }
    for (int i = 0; i < 2; i++) {
        if (e->d_chunk[i]) cudaFree(e->d_chunk[i]);
        if (e->ev_copied[i]) cudaEventDestroy(e->ev_copied[i]);
        if (e->ev_counted[i]) cudaEventDestroy(e->ev_counted[i]);
    }
    if (e->line_block) cudaFree(e->line_block);
    if (e->table_block) cudaFree(e->table_block);
    if (e->T.arena) cudaFree(e->T.arena);
    if (e->d_status) cudaFree(e->d_status);
    if (e->d_ws) cudaFree(e->d_ws);
    if (e->d_cursors) cudaFree(e->d_cursors);
    for (int i = 0; i < 2; i++) if (e->h_stage[i]) cudaFreeHost(e->h_stage[i]);
    if (e->peer.area) cudaFree(e->peer.area);
    if (e->peer.mapped && e->peer.ipc) cudaIpcCloseMemHandle(e->peer.mapped);
    if (e->peer.d_cur) cudaFree(e->peer.d_cur);
    if (e->peer.h_hdr) cudaFreeHost(e->peer.h_hdr);
    if (e->peer.d_job) cudaFree(e->peer.d_job);
    if (e->d_fin) cudaFree(e->d_fin);
    if (e->d_ord) cudaFree(e->d_ord);
    if (e->d_fz) cudaFree(e->d_fz);
    if (e->d_out) cudaFree(e->d_out);
    if (e->h_out) cudaFreeHost(e->h_out);
    if (e->h_fin) cudaFreeHost(e->h_fin);
    if (e->so_key) { cudaFree(e->so_key); cudaFree(e->so_count); cudaFree(e->so_first); }
    if (e->em.cursor) cudaFree(e->em.cursor);
    if (e->em.hash) cudaFree(e->em.hash);
    if (e->em.idx) cudaFree(e->em.idx);
    if (e->em.key) cudaFree(e->em.key);
    if (e->em.end) cudaFree(e->em.end);
    if (e->d_ovf) cudaFree(e->d_ovf);
    if (e->h_status) cudaFreeHost(e->h_status);
    if (e->h_prev) cudaFreeHost(e->h_prev);
    if (e->ev_prev) cudaEventDestroy(e->ev_prev);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    delete e;
}

static int create_impl(wc_engine *e, const wc_config *cfg) {
    wc_config_default(&e->cfg);
    if (cfg) {
        if (cfg->struct_size != sizeof(wc_config)) { e->err = "wc_config.struct_size mismatch"; return WC_ERR_INVALID; }
        e->cfg = *cfg;
    }
    wc_config &c = e->cfg;
    if (c.chunk_bytes == 0) c.chunk_bytes = 64ull << 20;
    if (c.short_table_log2 == 0) c.short_table_log2 = 21;
    if (c.long_table_log2 == 0) c.long_table_log2 = 18;
    if (c.arena_bytes == 0) c.arena_bytes = 256ull << 20;
    if (c.short_table_log2 < 10 || c.short_table_log2 > 32 || c.long_table_log2 < 10 || c.long_table_log2 > 34 ||
        c.arena_bytes < 4096 || (c.mode != WC_MODE_REF && c.mode != WC_MODE_EXACT) || (c.flags & ~(WC_FLAG_STRICT_REF | WC_FLAG_LANE_SERIAL)) != 0 || c.mapper > WC_MAP_LOWER_ASCII || c.job > WC_JOB_LINE_LENGTHS) {
        e->err = "wc_config out of range";
        return WC_ERR_INVALID;
    }
    c.chunk_bytes = align_up(c.chunk_bytes, 4096);
    e->device = c.device;
    CK(cudaSetDevice(e->device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, e->device));
    if (prop.major < 10) { e->err = "libwc.so needs an sm_100a GPU (B200)"; return WC_ERR_CUDA; }
    e->num_sms = prop.multiProcessorCount;
    e->grid = e->num_sms; // persistent: one CTA per SM, the shared-memory table fills the SM
    CK(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
    CK(cudaFuncSetAttribute(wc::count_kernel<wc::MapIdentity>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wc::kSmemTotal));
    CK(cudaFuncSetAttribute(wc::count_kernel<wc::MapLowerAscii>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wc::kSmemTotal));
    CK(cudaFuncSetAttribute(wc::count_kernel2<wc::MapIdentity>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wc::k2SmemTotal));
    CK(cudaFuncSetAttribute(wc::count_kernel2<wc::MapLowerAscii>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wc::k2SmemTotal));
    CK(cudaMalloc((void **)&e->d_status, sizeof(wc::Status)));
    CK(cudaMallocHost((void **)&e->h_status, sizeof(wc::Status)));
    CK(cudaMallocHost((void **)&e->h_prev, sizeof(wc::Status)));
    CK(cudaEventCreateWithFlags(&e->ev_prev, cudaEventDisableTiming));
    e->T.st = e->d_status;
    e->n_ws = (u32)e->grid * kMaxTokenWarps;
    CK(cudaMalloc((void **)&e->d_ws, sizeof(wc::WarpSummary) * e->n_ws));
    e->ovf_cap = 8ull << 20; // 8 Mi records (256 MiB): more words than a 64 MiB chunk can hold
    CK(cudaMalloc((void **)&e->d_ovf, e->ovf_cap * 32));
    e->T.ovf_log = e->d_ovf; e->T.ovf_cap = e->ovf_cap;
    int rc = alloc_tables(e, c.short_table_log2, c.long_table_log2, c.arena_bytes, &e->T, &e->table_block,
                          &e->table_block_bytes);
    if (rc != WC_OK) return rc;
    if (c.job == WC_JOB_LINE_LENGTHS) { // [bins][cursor, error flag][list of longer lines]
        const u64 ovf_cap = 1ull << 20;
        e->line_block_bytes = (wc::kLineBins + 2 + ovf_cap) * sizeof(u64);
        CK(cudaMalloc(&e->line_block, e->line_block_bytes));
        u64 *q = (u64 *)e->line_block;
        e->line.hist = q; e->line.ovf_count = q + wc::kLineBins; e->line.err = (u32 *)(q + wc::kLineBins + 1);
        e->line.ovf = q + wc::kLineBins + 2; e->line.ovf_cap = ovf_cap;
    }
    e->sort_mode = c.path == WC_PATH_SORT && c.job == WC_JOB_WORDCOUNT;
    e->auto_path = c.path == WC_PATH_AUTO && c.job == WC_JOB_WORDCOUNT;
    if (e->sort_mode) CK(cudaMalloc((void **)&e->em.cursor, sizeof(u64)));
    rc = reset_device_state(e);
    if (rc != WC_OK) return rc;
    CK(cudaStreamSynchronize(e->stream));
    return WC_OK;
}

int wc_create(wc_engine **out, const wc_config *cfg) {
    if (!out) return WC_ERR_INVALID;
    *out = nullptr;
    wc_engine *e = new (std::nothrow) wc_engine();
    if (!e) return WC_ERR_NOMEM;
    int rc = create_impl(e, cfg);
    if (rc != WC_OK) {
        g_create_err = e->err;
        wc_destroy(e);
        return rc;
    }
    *out = e;
    return WC_OK;
}

int wc_reset(wc_engine *e) {
    if (!e) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->copy_stream));
    poll_job_snapshot(e);
    if (e->consumed > 0 && !e->prev_pending && !e->sort_mode && e->cfg.job == WC_JOB_WORDCOUNT) {
        CK(cudaMemcpyAsync(e->h_prev, e->d_status, sizeof(wc::Status), cudaMemcpyDeviceToHost, e->stream));
        CK(cudaEventRecord(e->ev_prev, e->stream));
        e->prev_pending = true;
    }
    int rc = reset_device_state(e);
    if (rc != WC_OK) return rc;
    e->carry.clear();
    e->auto_decided = false;
    e->job_stats_set = false; e->job_bytes = 0;
    e->consumed = 0; e->fed = 0; e->launches = 0; e->pieces = 0; e->settle_due = false; e->device_ms = 0; e->count_kernel_ms = 0;
    for (auto &p : e->ev_used) { e->ev_pool.push_back(p.a); e->ev_pool.push_back(p.mid); e->ev_pool.push_back(p.b); }
    e->ev_used.clear();
    return WC_OK;
}

} // extern "C"

// Re-hashes the word table into one with 2^new_log2 slots (stream ordered).
static int grow_word_table(wc_engine *e, u32 new_log2) {
    if (new_log2 > 32) { e->err = "word table cannot grow beyond 2^32 slots"; return WC_ERR_TABLE_FULL; }
    const wc::Tables old = e->T;
    void *old_block = e->table_block;
    wc::Tables nt; void *nb = nullptr; size_t nbytes = 0;
    int rc = alloc_tables(e, new_log2, old.x_log2, 0, &nt, &nb, &nbytes);
    if (rc != WC_OK) return rc;
    CK(cudaMemsetAsync(nb, 0, nbytes, e->stream));
    // the long table moves as it is (same capacity): a plain copy of its three arrays
    const size_t l_n = (size_t)1 << old.x_log2;
    CK(cudaMemcpyAsync(nt.x_refs, old.x_refs, 3 * l_n * sizeof(u64), cudaMemcpyDeviceToDevice, e->stream));
    wc::rehash_words_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(nt, old.w_keys, old.w_counts, old.w_finv,
                                                                  1ull << old.w_log2);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(e->stream));
    CK(cudaFree(old_block));
    e->T = nt; e->table_block = nb; e->table_block_bytes = nbytes;
    e->cfg.short_table_log2 = new_log2;
    e->grows++;
    return WC_OK;
}

// count_kernel's warp split follows the text: the share of words that the tokenizing warps hand to the consumer
// warps (misses of the CTA's table and words of 9+ bytes) is read from the counters whenever the host looks at
// them anyway -- after the first pieces of a fed job, and at wc_finish for the job after it.  Measured on one
// B200 (profiles/r2_kernel_experiments.txt, F): at 0.17 (config 2) 20 + 8 warps is the fastest split, 18 + 10
// costs 5 %; at 0.27 (config 3) 18 + 10 is 13 % faster; at 0.78 (config 5: almost every word is a new table
// entry, the consumers wait for DRAM whatever their number) more consumers do not help.
constexpr u64 kLoadMinWords = 100000;
static void note_consumer_load(wc_engine *e, const wc::Status &st) {
    static const int forced = [] { const char *v = getenv("WC_PRODUCERS"); return v ? atoi(v) : 0; }();
    if (forced) { e->producers = (u32)forced; return; } // (launch_count falls back to the default split if it is out of range)
    const u64 words = st.job_set ? st.own_words : st.n_flush - st.n_empty; // (after a peer pull the counters are the job's)
    if (words < kLoadMinWords) return;
    const double load = (double)(st.n_spill + st.n_long) / (double)words;
    e->producers = (load >= 0.22 && load < 0.60) ? (u32)wc::kProducersHeavy : (u32)wc::kProducers;
}

// A job that ends in wc_reset without a wc_finish (a caller that only wants device-side tables, a benchmark
// loop) is looked at too, without a host wait: wc_reset queues a copy of the counters in front of the reset,
// and whoever launches next reads it once the copy has completed.
static void poll_job_snapshot(wc_engine *e) {
    if (!e->prev_pending) return;
    const cudaError_t q = cudaEventQuery(e->ev_prev);
    if (q == cudaErrorNotReady) { (void)cudaGetLastError(); return; } // (not an error: keep it out of the next CK(cudaGetLastError()))
    if (q != cudaSuccess) return;                                     // (a real failure surfaces at the caller's next checked call)
    e->prev_pending = false;
    note_consumer_load(e, *e->h_prev);
}

// After a sync: if words overflowed the table (or it is more than half full), grow it and replay the
// overflow log.  Leaves e->h_status up to date.  No word is ever dropped: wc_device.cuh logs what the
// table could not take.
static int settle_tables(wc_engine *e) {
    Range nvtx_range("wc: settle tables");
    for (int round = 0; round < 8; round++) {
        CK(cudaMemcpyAsync(e->h_status, e->d_status, sizeof(wc::Status), cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        const wc::Status st = *e->h_status;
        const u64 cap = 1ull << e->T.w_log2;
        note_consumer_load(e, st);
        if ((st.err & wc::kErrWordFull) != 0) return WC_OK; // the log itself overflowed: reported by the caller
        if (st.ovf_count == 0 && st.w_entries * 2 <= cap) return WC_OK;
        u32 want = e->T.w_log2;
        while ((1ull << want) < 4 * (st.w_entries + st.ovf_count) && want < 32) want++;
        if (want == e->T.w_log2 && st.ovf_count == 0) return WC_OK;
        if (want > e->T.w_log2) { int rc = grow_word_table(e, want); if (rc != WC_OK) return rc; }
        if (st.ovf_count) {
            // The replay may overflow again (then: next round) and appends to the log from its start, so it
            // reads a copy: records not yet replayed must not be overwritten by the ones it produces.
            void *copy = nullptr;
            CK(cudaMallocAsync(&copy, st.ovf_count * sizeof(wc::WordRecord), e->stream));
            CK(cudaMemcpyAsync(copy, e->d_ovf, st.ovf_count * sizeof(wc::WordRecord), cudaMemcpyDeviceToDevice, e->stream));
            e->h_status->ovf_count = 0;
            CK(cudaMemcpyAsync(&e->d_status->ovf_count, &e->h_status->ovf_count, sizeof(u64), cudaMemcpyHostToDevice, e->stream));
            wc::import_words_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(e->T, (const wc::WordRecord *)copy, st.ovf_count);
            CK(cudaGetLastError());
            CK(cudaFreeAsync(copy, e->stream));
        }
    }
    return WC_OK;
}

// Sort path: room for the records of a piece of n more bytes (at most one word per two bytes).  The
// number written so far is read back, so the bound stays tight from piece to piece.
static int ensure_emit_capacity(wc_engine *e, size_t n) {
    u64 have = 0;
    if (e->em_bound > 0) {
        CK(cudaStreamSynchronize(e->stream));
        CK(cudaMemcpy(&have, e->em.cursor, sizeof(u64), cudaMemcpyDeviceToHost));
    }
    const u64 need = have + n / 2 + 64;
    if (need >= 0xFFFFFFF0ull) { e->err = "sort path: more than 2^32 words in one job"; return WC_ERR_TOO_LARGE; }
    if (need > e->em.cap) {
        u64 cap = e->em.cap ? e->em.cap : 1;
        while (cap < need) cap *= 2;
        if (cap > 0xFFFFFFF0ull) cap = 0xFFFFFFF0ull;
        wc::EmitBuf nb = e->em;
        CK(cudaMalloc((void **)&nb.hash, cap * sizeof(u64)));
        CK(cudaMalloc((void **)&nb.idx, cap * sizeof(u32)));
        CK(cudaMalloc((void **)&nb.key, cap * sizeof(ulonglong2)));
        CK(cudaMalloc((void **)&nb.end, cap * sizeof(u64)));
        if (have) {
            CK(cudaMemcpy(nb.key, e->em.key, have * sizeof(ulonglong2), cudaMemcpyDeviceToDevice));
            CK(cudaMemcpy(nb.end, e->em.end, have * sizeof(u64), cudaMemcpyDeviceToDevice));
        }
        if (e->em.hash) { cudaFree(e->em.hash); cudaFree(e->em.idx); cudaFree(e->em.key); cudaFree(e->em.end); }
        nb.cap = cap;
        e->em = nb;
    }
    e->em_bound = need;
    return WC_OK;
}

// Sort path, device half: sort the records by hash, mark and number the runs of equal words, reduce
// every run to (word, count, first offset).  The result stays in device memory (e->so_*).
static int sort_aggregate(wc_engine *e, u64 stop_abs) {
    Range nvtx_range("wc: sort + segmented reduce");
    if (e->so_valid) return WC_OK;
    if (e->so_key) { cudaFree(e->so_key); cudaFree(e->so_count); cudaFree(e->so_first); e->so_key = nullptr; }
    e->so_runs = 0; e->so_coll = 0;
    u64 n = 0;
    CK(cudaStreamSynchronize(e->stream));
    CK(cudaMemcpy(&n, e->em.cursor, sizeof(u64), cudaMemcpyDeviceToHost));
    if (n > e->em.cap) n = e->em.cap;
    e->so_words = n;
    e->so_valid = true;
    if (n == 0) return WC_OK;
    u64 *hash_alt = nullptr; u32 *idx_alt = nullptr, *flags = nullptr; void *temp = nullptr; u64 *d_coll = nullptr;
    CK(cudaMalloc((void **)&hash_alt, n * sizeof(u64)));
    CK(cudaMalloc((void **)&idx_alt, n * sizeof(u32)));
    CK(cudaMalloc((void **)&flags, n * sizeof(u32)));
    CK(cudaMalloc((void **)&d_coll, sizeof(u64)));
    CK(cudaMemsetAsync(d_coll, 0, sizeof(u64), e->stream));
    wc::sort_prepare_kernel<<<e->num_sms * 8, 256, 0, e->stream>>>(e->em.key, n, e->em.hash, e->em.idx);
    cub::DoubleBuffer<u64> dk(e->em.hash, hash_alt);
    cub::DoubleBuffer<u32> dv(e->em.idx, idx_alt);
    size_t tb_sort = 0, tb_scan = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tb_sort, dk, dv, n, 0, 64, e->stream));
    CK(cub::DeviceScan::InclusiveSum(nullptr, tb_scan, flags, flags, n, e->stream));
    const size_t tb = tb_sort > tb_scan ? tb_sort : tb_scan;
    CK(cudaMalloc(&temp, tb ? tb : 16));
    size_t tbx = tb;
    CK(cub::DeviceRadixSort::SortPairs(temp, tbx, dk, dv, n, 0, 64, e->stream));
    wc::sort_mark_heads_kernel<<<e->num_sms * 8, 256, 0, e->stream>>>(dk.Current(), dv.Current(), e->em.key, n, flags, d_coll);
    tbx = tb;
    CK(cub::DeviceScan::InclusiveSum(temp, tbx, flags, flags, n, e->stream)); // flags -> 1-based run numbers
    u32 n_runs = 0; u64 coll = 0;
    CK(cudaMemcpyAsync(&n_runs, flags + (n - 1), sizeof(u32), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(&coll, d_coll, sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    CK(cudaMalloc((void **)&e->so_key, (size_t)n_runs * sizeof(ulonglong2)));
    CK(cudaMalloc((void **)&e->so_count, (size_t)n_runs * sizeof(u64)));
    CK(cudaMalloc((void **)&e->so_first, (size_t)n_runs * sizeof(u64)));
    CK(cudaMemsetAsync(e->so_count, 0, (size_t)n_runs * sizeof(u64), e->stream));
    CK(cudaMemsetAsync(e->so_first, 0xFF, (size_t)n_runs * sizeof(u64), e->stream));
    wc::sort_reduce_runs_kernel<<<e->num_sms * 8, 256, 0, e->stream>>>(dv.Current(), e->em.key, e->em.end, flags, n, stop_abs,
                                                                      e->so_key, e->so_count, e->so_first);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(e->stream));
    e->launches += 3;
    e->so_runs = n_runs; e->so_coll = coll;
    cudaFree(hash_alt); cudaFree(idx_alt); cudaFree(flags); cudaFree(temp); cudaFree(d_coll);
    return WC_OK;
}

// Sort path, host half: bring the runs over and append them to `entries` / `blob`.
static int finish_sorted(wc_engine *e, u64 stop_abs, std::string &blob, std::vector<wc::HostEntry> &entries, u64 *n_words_out) {
    int rc = sort_aggregate(e, stop_abs);
    if (rc != WC_OK) return rc;
    *n_words_out = e->so_words;
    const u32 n_runs = e->so_runs;
    const u64 coll = e->so_coll;
    if (n_runs == 0) return WC_OK;
    std::vector<u64> hk((size_t)n_runs * 2), hc(n_runs), hf(n_runs);
    CK(cudaMemcpyAsync(hk.data(), e->so_key, hk.size() * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(hc.data(), e->so_count, hc.size() * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(hf.data(), e->so_first, hf.size() * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));

    struct KeyHash { size_t operator()(const std::pair<u64, u64> &k) const { return (size_t)(k.first * 0x9E3779B97F4A7C15ull ^ k.second); } };
    std::unordered_map<std::pair<u64, u64>, size_t, KeyHash> seen; // only if two words shared a hash
    entries.reserve(entries.size() + n_runs);
    for (u32 r = 0; r < n_runs; r++) {
        const u64 k[2] = {hk[2 * (size_t)r], hk[2 * (size_t)r + 1]};
        if (hc[r] == 0) continue;
        if (coll) {
            auto it = seen.find({k[0], k[1]});
            if (it != seen.end()) {
                wc::HostEntry &he = entries[it->second];
                he.count += (int64_t)hc[r];
                if (hf[r] < he.first) he.first = hf[r];
                continue;
            }
            seen.emplace(std::make_pair(k[0], k[1]), entries.size());
        }
        const unsigned char *kb = (const unsigned char *)k;
        u32 len = 16;
        while (len > 1 && kb[len - 1] == 0) len--;
        entries.push_back(wc::HostEntry{blob.size(), len, (int64_t)hc[r], hf[r]});
        blob.append((const char *)kb, len);
    }
    return WC_OK;
}

// Which kernel counts a forward launch: count_kernel (wc_device.cuh), or -- with WC_FLAG_LANE_SERIAL, or
// WC_KERNEL=k2 in the environment (A/B timing) -- count_kernel2 (wc_count2.cuh) where it applies: the hash
// path outside strict-ref mode.  Roll-back launches always run count_kernel.
static bool forward_uses_kernel2(uint32_t flags, bool sort_mode, bool strict) {
    static const bool env_k2 = [] { const char *k = getenv("WC_KERNEL"); return k && strcmp(k, "k2") == 0; }();
    return (env_k2 || (flags & WC_FLAG_LANE_SERIAL) != 0) && !sort_mode && !strict;
}

// One count launch (forward or roll-back) over `n` bytes.  Both kernels split the piece evenly over
// their tokenizing warps; L.tiles_per_warp is set here for the kernel that runs.  Returns false if the
// piece is too large for one launch (a CTA's range must fit 32-bit offsets).
static bool launch_count(cudaStream_t stream, int grid, uint32_t mapper, wc::Launch L, const wc::Tables &T, bool kernel2) {
    const u64 tiles = (L.n + wc::kTile - 1) / wc::kTile;
    if (L.producers < (u32)wc::kMinProducers || L.producers > (u32)wc::kMaxProducers) L.producers = wc::kProducers;
    L.smem_entries = wc::smem_entries_for((int)L.producers);
    const u64 per_cta = kernel2 ? wc::k2Warps : L.producers;
    const u64 warps = (u64)grid * per_cta;
    const u64 tpw = (tiles + warps - 1) / warps;
    if (tpw * per_cta * wc::kTile >= 0xFFFFFFF0ull) return false;
    L.tiles_per_warp = (u32)tpw;
    if (kernel2) {
        if (mapper == WC_MAP_LOWER_ASCII)
            wc::count_kernel2<wc::MapLowerAscii><<<grid, wc::k2Threads, wc::k2SmemTotal, stream>>>(L, T);
        else
            wc::count_kernel2<wc::MapIdentity><<<grid, wc::k2Threads, wc::k2SmemTotal, stream>>>(L, T);
    } else {
        if (mapper == WC_MAP_LOWER_ASCII)
            wc::count_kernel<wc::MapLowerAscii><<<grid, wc::kThreads, wc::kSmemTotal, stream>>>(L, T);
        else
            wc::count_kernel<wc::MapIdentity><<<grid, wc::kThreads, wc::kSmemTotal, stream>>>(L, T);
    }
    return true;
}

// ---- WC_PATH_AUTO.  The hash path costs one table update per word but each NEW word costs a 128-bit CAS
// and, once the vocabulary has outgrown the L2, random DRAM sectors; the sort path costs the same for every
// word whatever the vocabulary (a 32-byte record written, sorted and read back).  The first job of an engine
// is probed: the head of its first piece is counted on the hash path, the counters are read, the tables are
// reset; later jobs follow the ratio of the job before them.  Measured on one B200
// (profiles/r2_path_crossover.txt) the hash path wins at every share of distinct words that was tried --
// 0.0003 (config 2: 1.6 ms against 30 ms per GiB) up to 0.80 (config 5 at full size: 55 ms against 125 ms per
// step) -- so the default threshold is "never" (2.0) and the sort path is taken when the caller asks for it
// (WC_PATH_SORT: no table to size) or lowers the threshold (WC_AUTO_SORT_RATIO).
static double auto_sort_ratio() {
    const char *v = getenv("WC_AUTO_SORT_RATIO");
    return v ? atof(v) : 2.0;
}
constexpr size_t kAutoProbeBytes = (size_t)8 << 20, kAutoProbeMinWords = 100000;

static int set_sort_mode(wc_engine *e, bool on) {
    if (on && !e->em.cursor) {
        CK(cudaMalloc((void **)&e->em.cursor, sizeof(u64)));
        CK(cudaMemsetAsync(e->em.cursor, 0, sizeof(u64), e->stream));
    }
    e->sort_mode = on;
    return WC_OK;
}

static bool launch_count(cudaStream_t stream, int grid, uint32_t mapper, wc::Launch L, const wc::Tables &T, bool kernel2);
static int auto_choose_path(wc_engine *e, const uint8_t *dev, size_t n) {
    if (e->auto_ratio < 0) { // the engine's first job: probe the head of the piece on the hash path
        e->auto_decided = true;
        if (n < ((size_t)1 << 20)) return set_sort_mode(e, false);
        wc::Launch L{};
        L.buf = dev; L.n = n < kAutoProbeBytes ? n : kAutoProbeBytes; L.base = e->cfg.base_offset; L.ws = e->d_ws;
        L.n_ws = (u32)e->grid * wc::kProducers;
        if (!launch_count(e->stream, e->grid, e->cfg.mapper, L, e->T, false)) return WC_ERR_TOO_LARGE;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(e->h_status, e->d_status, sizeof(wc::Status), cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        const wc::Status st = *e->h_status;
        const u64 words = st.n_flush - st.n_empty, distinct = st.w_entries + st.x_entries;
        int rc = reset_device_state(e); // the probe leaves no trace: the job starts from empty tables
        if (rc != WC_OK) return rc;
        e->launches += 1;
        // (a probe that overflowed a small table says nothing: the overflow log counts occurrences, not words)
        if (words >= kAutoProbeMinWords && st.ovf_count == 0 && st.err == 0) e->auto_ratio = (double)distinct / (double)words;
    }
    e->auto_decided = true;
    return set_sort_mode(e, e->auto_ratio >= auto_sort_ratio());
}

// Launches the forward count of one piece and its roll-back/summary companion on e->stream.
static int launch_piece(wc_engine *e, const uint8_t *dev, size_t n, int last_piece) {
    Range nvtx_range("wc: count piece");
    if (n == 0) return WC_OK;
    if (e->cfg.job == WC_JOB_LINE_LENGTHS) { // the other job: one kernel, its own reducer (wc_jobs.cuh)
        const int grid = e->num_sms * 2;
        const u64 tiles = (n + wc::kTile - 1) / wc::kTile, warps = (u64)grid * (wc::kLineThreads / 32);
        const u64 tpw = (tiles + warps - 1) / warps;
        if (tpw >= 0xFFFFFFF0ull) { e->err = "piece too large for one launch"; return WC_ERR_TOO_LARGE; }
        wc::line_length_kernel<<<grid, wc::kLineThreads, 0, e->stream>>>(dev, n, (u32)tpw, last_piece, e->line);
        CK(cudaGetLastError());
        e->consumed += n;
        e->launches += 1;
        return WC_OK;
    }
    if (e->auto_path && !e->auto_decided) { // the first piece of a job
        int rc = auto_choose_path(e, dev, n);
        if (rc != WC_OK) return rc;
    }
    const bool strict = (e->cfg.flags & WC_FLAG_STRICT_REF) != 0;
    const bool k2 = forward_uses_kernel2(e->cfg.flags, e->sort_mode, strict);
    poll_job_snapshot(e);
    wc::Launch L{};
    L.buf = dev; L.n = n; L.base = e->cfg.base_offset + e->consumed;
    L.last_piece = last_piece; L.negate = 0;
    L.strict = strict;
    L.producers = (k2 || strict || e->sort_mode) ? (u32)wc::kProducers : e->producers;
    L.ws = e->d_ws; L.n_ws = (u32)e->grid * (k2 ? wc::k2Warps : L.producers); // one summary per tokenizing warp
    e->last_tokenizing_warps = k2 ? wc::k2Warps : L.producers;
    if (e->sort_mode) {
        int rc = ensure_emit_capacity(e, n);
        if (rc != WC_OK) return rc;
        L.em = e->em;
        e->so_valid = false;
    }

    cudaEvent_t a, mid, b;
    for (cudaEvent_t *ev : {&a, &mid, &b}) {
        if (!e->ev_pool.empty()) { *ev = e->ev_pool.back(); e->ev_pool.pop_back(); }
        else CK(cudaEventCreate(ev));
    }
    CK(cudaEventRecord(a, e->stream));
    if (!launch_count(e->stream, e->grid, e->cfg.mapper, L, e->T, k2)) { e->err = "piece too large for one launch"; return WC_ERR_TOO_LARGE; }
    CK(cudaEventRecord(mid, e->stream));
    L.negate = 1; // (reads the forward launch's summaries: same L.ws / L.n_ws)
    if (!launch_count(e->stream, e->grid, e->cfg.mapper, L, e->T, false)) { e->err = "piece too large for one launch"; return WC_ERR_TOO_LARGE; }
    CK(cudaGetLastError());
    CK(cudaEventRecord(b, e->stream));
    e->ev_used.push_back(wc_engine::EvTriple{a, mid, b});
    e->consumed += n;
    e->launches += 2;
    return WC_OK;
}

static int ensure_chunks(wc_engine *e) {
    if (e->d_chunk[0]) return WC_OK;
    e->chunk_cap = e->cfg.chunk_bytes;
    for (int i = 0; i < 2; i++) {
        CK(cudaMalloc((void **)&e->d_chunk[i], e->chunk_cap + 64));
        CK(cudaEventCreateWithFlags(&e->ev_copied[i], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&e->ev_counted[i], cudaEventDisableTiming));
        CK(cudaEventRecord(e->ev_counted[i], e->stream));
        CK(cudaEventRecord(e->ev_copied[i], e->copy_stream));
    }
    return WC_OK;
}

// One piece (whole lines, or the end of the input) from host memory through the chunk ring.
static int feed_piece(wc_engine *e, const uint8_t *src, size_t n, int last_piece) {
    if (n == 0) return WC_OK;
    int rc = ensure_chunks(e);
    if (rc != WC_OK) return rc;
    if (n <= e->chunk_cap) {
        const int s = e->ring;
        // Pageable source (malloc'd buffers, std::vector, the carry): cudaMemcpyAsync would stage it through
        // the driver's small bounce buffer on this thread.  Copy it into a pinned slot with several threads
        // instead and let the DMA engine take it from there, overlapped with the next slot's fill.
        if (n >= ((size_t)1 << 20)) {
            cudaPointerAttributes at;
            const bool pageable = cudaPointerGetAttributes(&at, src) != cudaSuccess || at.type == cudaMemoryTypeUnregistered;
            (void)cudaGetLastError();
            if (pageable) {
                if (!e->h_stage[s]) CK(cudaMallocHost((void **)&e->h_stage[s], e->chunk_cap + 64));
                CK(cudaEventSynchronize(e->ev_copied[s])); // the copy that last read this slot is done
                wc::parallel_copy(e->h_stage[s], src, n);
                src = e->h_stage[s];
            }
        }
        CK(cudaStreamWaitEvent(e->copy_stream, e->ev_counted[s], 0)); // the kernel that read this buffer is done
        CK(cudaMemcpyAsync(e->d_chunk[s], src, n, cudaMemcpyHostToDevice, e->copy_stream));
        CK(cudaEventRecord(e->ev_copied[s], e->copy_stream));
        // Keep the word table ahead of the vocabulary: after each of the first pieces, and then every
        // eighth, wait for the piece and grow the table / replay the overflow log if needed.  One chunk
        // can never overflow the log (it holds more records than a chunk has words).  The wait for the
        // PREVIOUS piece happens here, while this piece's copy is already running.
        if (e->settle_due) {
            e->settle_due = false;
            CK(cudaStreamSynchronize(e->stream));
            rc = settle_tables(e);
            if (rc != WC_OK) return rc;
        }
        CK(cudaStreamWaitEvent(e->stream, e->ev_copied[s], 0));
        rc = launch_piece(e, e->d_chunk[s], n, last_piece);
        if (rc != WC_OK) return rc;
        CK(cudaEventRecord(e->ev_counted[s], e->stream));
        e->ring ^= 1;
        e->pieces++;
        if (e->pieces <= 4 || (e->pieces & 7) == 0) e->settle_due = true;
        return WC_OK;
    }
    // a single line longer than the chunk: give it a buffer of its own
    uint8_t *tmp = nullptr;
    CK(cudaMallocAsync((void **)&tmp, n + 64, e->stream));
    CK(cudaMemcpyAsync(tmp, src, n, cudaMemcpyHostToDevice, e->stream));
    rc = launch_piece(e, tmp, n, last_piece);
    if (rc != WC_OK) return rc;
    CK(cudaFreeAsync(tmp, e->stream));
    CK(cudaStreamSynchronize(e->stream)); // src must stay valid until the copy has run
    return WC_OK;
}

extern "C" {

int wc_feed(wc_engine *e, const void *bytes, size_t n) {
    Range nvtx_range("wc_feed");
    if (!e || (!bytes && n)) return WC_ERR_INVALID;
    if (n == 0) return WC_OK;
    CK(cudaSetDevice(e->device));
    const uint8_t *p = (const uint8_t *)bytes;
    size_t left = n;
    e->fed += n;
    int rc = WC_OK;

    // finish the line left over from the previous call
    if (!e->carry.empty()) {
        const uint8_t *nl = (const uint8_t *)memchr(p, '\n', left);
        const size_t take = nl ? (size_t)(nl - p) + 1 : left;
        e->carry.insert(e->carry.end(), p, p + take);
        p += take; left -= take;
        if (!nl) return WC_OK;
        rc = feed_piece(e, e->carry.data(), e->carry.size(), 0);
        if (rc != WC_OK) return rc;
        CK(cudaStreamSynchronize(e->copy_stream)); // carry is about to be reused
        e->carry.clear();
    }
    // whole-line pieces straight from the caller's buffer
    while (left > 0) {
        const size_t take = left < e->cfg.chunk_bytes ? left : (size_t)e->cfg.chunk_bytes;
        const uint8_t *nl = (const uint8_t *)memrchr(p, '\n', take);
        size_t piece;
        if (nl) piece = (size_t)(nl - p) + 1;
        else {
            const uint8_t *fw = take < left ? (const uint8_t *)memchr(p + take, '\n', left - take) : nullptr;
            if (!fw) break; // no line end in what is left: keep it for the next call
            piece = (size_t)(fw - p) + 1;
        }
        rc = feed_piece(e, p, piece, 0);
        if (rc != WC_OK) return rc;
        p += piece; left -= piece;
    }
    if (left > 0) e->carry.insert(e->carry.end(), p, p + left);
    CK(cudaStreamSynchronize(e->copy_stream)); // the caller may reuse `bytes` now
    return WC_OK;
}

int wc_count_device(wc_engine *e, const void *dev_bytes, size_t n, int last_piece) {
    Range nvtx_range("wc_count_device");
    if (!e || (!dev_bytes && n)) return WC_ERR_INVALID;
    if (!e->carry.empty()) { e->err = "wc_count_device after an unfinished wc_feed line"; return WC_ERR_INVALID; }
    if (((uintptr_t)dev_bytes & 15) != 0) { e->err = "device buffer must be 16-byte aligned"; return WC_ERR_UNALIGNED; }
    CK(cudaSetDevice(e->device));
    e->fed += n;
    return launch_piece(e, (const uint8_t *)dev_bytes, n, last_piece);
}

int wc_sync(wc_engine *e) {
    if (!e) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->copy_stream));
    CK(cudaStreamSynchronize(e->stream));
    for (auto &p : e->ev_used) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, p.a, p.b));
        e->device_ms += ms;
        CK(cudaEventElapsedTime(&ms, p.a, p.mid));
        e->count_kernel_ms += ms;
        e->ev_pool.push_back(p.a); e->ev_pool.push_back(p.mid); e->ev_pool.push_back(p.b);
    }
    e->ev_used.clear();
    return WC_OK;
}

int wc_aggregate(wc_engine *e) {
    if (!e) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    if (!e->sort_mode) return WC_OK; // the hash tables are the aggregate already
    int rc = wc_end_input(e);
    if (rc != WC_OK) return rc;
    CK(cudaMemcpyAsync(e->h_status, e->d_status, sizeof(wc::Status), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return sort_aggregate(e, e->h_status->stop);
}

int wc_device_ms(wc_engine *e, double *ms, double *count_kernel_ms, uint64_t *kernel_launches) {
    if (!e) return WC_ERR_INVALID;
    if (ms) *ms = e->device_ms;
    if (count_kernel_ms) *count_kernel_ms = e->count_kernel_ms;
    if (kernel_launches) *kernel_launches = e->launches;
    return WC_OK;
}

uint64_t wc_finish_d2h_bytes(const wc_engine *e) { return e ? e->finish_d2h : 0; }

void *wc_stream(wc_engine *e) { return e ? (void *)e->stream : nullptr; }

int wc_host_alloc(void **p, size_t n) {
    if (!p) return WC_ERR_INVALID;
    return cudaMallocHost(p, n ? n : 1) == cudaSuccess ? WC_OK : WC_ERR_NOMEM;
}

int wc_host_free(void *p) { return cudaFreeHost(p) == cudaSuccess ? WC_OK : WC_ERR_CUDA; }

int wc_end_input(wc_engine *e) {
    if (!e) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    if (!e->carry.empty()) { // the unterminated end of the input (main.cu:198-203 drops its last word)
        int rc = feed_piece(e, e->carry.data(), e->carry.size(), 1);
        if (rc != WC_OK) return rc;
        CK(cudaStreamSynchronize(e->copy_stream));
        CK(cudaStreamSynchronize(e->stream));
        e->carry.clear();
    }
    return WC_OK;
}

// WC_JOB_LINE_LENGTHS: the histogram comes to the host, one row per length that occurs (key in decimal).
static int finish_line_lengths(wc_engine *e, wc_result **out) {
    std::vector<u64> bins(wc::kLineBins + 2);
    CK(cudaMemcpyAsync(bins.data(), e->line_block, bins.size() * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    const u64 n_ovf = bins[wc::kLineBins];
    if (*(const u32 *)&bins[wc::kLineBins + 1] != 0 || n_ovf > e->line.ovf_cap) {
        e->err = "line-length job: more lines of 4096+ bytes than its list holds";
        return WC_ERR_TABLE_FULL;
    }
    std::vector<u64> longs((size_t)n_ovf);
    if (n_ovf) CK(cudaMemcpy(longs.data(), e->line.ovf, n_ovf * sizeof(u64), cudaMemcpyDeviceToHost));
    std::sort(longs.begin(), longs.end());
    wc_result *r = new (std::nothrow) wc_result();
    if (!r) return WC_ERR_NOMEM;
    auto add = [r](u64 len, u64 count) {
        char key[24];
        const int k = snprintf(key, sizeof(key), "%llu", (unsigned long long)len);
        r->rows.push_back(wc::Row{r->blob.size(), (uint32_t)k, (int64_t)count, len});
        r->blob.append(key, (size_t)k);
        r->total += (int64_t)count;
    };
    for (u32 len = 0; len < wc::kLineBins; len++)
        if (bins[len]) add(len, bins[len]);
    for (size_t i = 0; i < longs.size();) {
        size_t j = i;
        while (j < longs.size() && longs[j] == longs[i]) j++;
        add(longs[i], j - i);
        i = j;
    }
    r->stats.bytes = e->consumed;
    r->stats.n_flush = (u64)r->total; r->stats.stop = e->consumed; r->stats.n_distinct = r->rows.size();
    e->finish_d2h = (bins.size() + n_ovf) * sizeof(u64);
    *out = r;
    return WC_OK;
}

// The path this job took, and what it says about the next one (WC_PATH_AUTO).
static void note_job_path(wc_engine *e, wc_stats *s) {
    s->path_used = e->sort_mode ? WC_PATH_SORT : WC_PATH_HASH;
    s->tokenizing_warps = e->last_tokenizing_warps;
    const u64 words = s->n_flush - s->n_empty;
    if (words >= kAutoProbeMinWords) e->auto_ratio = (double)s->n_distinct / (double)words;
    s->distinct_ratio = words ? (double)s->n_distinct / (double)words : 0.0;
}

int wc_finish(wc_engine *e, wc_result **out) {
    Range nvtx_range("wc_finish");
    if (!e || !out) return WC_ERR_INVALID;
    *out = nullptr;
    CK(cudaSetDevice(e->device));
    static const bool timing = getenv("WC_TIMING") != nullptr; // stderr: where wc_finish spends its time
    const auto t0 = std::chrono::steady_clock::now();
    auto since = [&t0] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    int rc = wc_end_input(e);
    if (rc != WC_OK) return rc;
    rc = wc_sync(e);
    if (rc != WC_OK) return rc;
    const double t_sync = since();
    if (e->cfg.job == WC_JOB_LINE_LENGTHS) return finish_line_lengths(e, out);
    rc = settle_tables(e);
    if (rc != WC_OK) return rc;
    const double t_settle = since();
    e->carry.clear();
    const wc::Status st = *e->h_status;
    note_consumer_load(e, st);
    if (st.job_set) { e->job_stats_set = true; e->job_bytes = st.job_bytes; } // (wc_peer_pull_device set the job's counters on the device)
    if (st.err & wc::kErrPeer) {
        e->err = "peer gather: a sender's push did not arrive, or its tables / landing area overflowed";
        return WC_ERR_TOO_LARGE;
    }
    if (st.err) {
        char msg[160];
        snprintf(msg, sizeof(msg), "device tables full (flags 0x%x: words=%llu long=%llu arena=%llu units)", st.err,
                 (unsigned long long)st.w_entries, (unsigned long long)st.x_entries, (unsigned long long)st.arena_used);
        e->err = msg;
        return WC_ERR_TABLE_FULL;
    }

    // ---- bring the USED table entries to the host: compacted on the device into 32-byte records
    // (the same kernels the multi-GPU exchange exports with), then one copy into pinned memory.
    // The tables themselves can be tens of MB; the entries are a few.
    const size_t n_w_max = (size_t)st.w_entries, n_x_max = (size_t)st.x_entries;
    const size_t blob_max = (size_t)st.arena_used * 8;
    const size_t n_ord = e->sort_mode ? 0 : n_w_max + n_x_max; // (sort mode adds its entries on the host: no device order)
    const size_t off_w = 64, off_x = off_w + n_w_max * sizeof(wc::WordRecord),
                 off_b = off_x + n_x_max * sizeof(wc::LongRecord), off_ri = off_b + blob_max,
                 off_rc = off_ri + align_up(n_ord * sizeof(u32), 8), off_pk = off_rc + n_ord * sizeof(u64),
                 off_pf = off_pk + align_up(n_ord * sizeof(u32), 8), need = off_pf + align_up(n_ord * sizeof(u32), 8);
    // [cursors][word records][long records][blob][row record indices][row counts] go to the host;
    // [by_key][by_first] only when the host does the reduce (WC_HOST_REDUCE=1, a cross-check)
    static const bool host_reduce = getenv("WC_HOST_REDUCE") != nullptr;
    const size_t copy_bytes = n_ord > 1 ? need : off_pk; // (host reduce with device orders: the two orders come along)
    rc = ensure_cursors(e);
    if (rc != WC_OK) return rc;
    if (need > e->d_fin_cap) {
        if (e->d_fin) cudaFree(e->d_fin);
        e->d_fin = nullptr; e->d_fin_cap = 0;
        CK(cudaMalloc(&e->d_fin, need + need / 4));
        e->d_fin_cap = need + need / 4;
    }
    if (need > e->h_fin_cap && (host_reduce || n_ord <= 1)) { // (the device reduce has its own, smaller, host block)
        if (e->h_fin) cudaFreeHost(e->h_fin);
        e->h_fin = nullptr; e->h_fin_cap = 0;
        CK(cudaMallocHost(&e->h_fin, need + need / 4));
        e->h_fin_cap = need + need / 4;
    }
    uint8_t *dfin = (uint8_t *)e->d_fin, *hfin = (uint8_t *)e->h_fin;
    CK(cudaMemsetAsync(e->d_cursors, 0, (2 * kMaxParts + 2) * sizeof(u64), e->stream)); // offsets[0] = 0, cursors = 0
    if (n_w_max)
        wc::export_scatter_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(e->T, 1, e->d_cursors, e->d_cursors + kMaxParts,
                                                                        (wc::WordRecord *)(dfin + off_w));
    if (n_x_max)
        wc::export_long_kernel<<<e->num_sms, 256, 0, e->stream>>>(e->T, e->d_cursors + 2 * kMaxParts,
                                                                 (wc::LongRecord *)(dfin + off_x), (u64 *)(dfin + off_b));
    CK(cudaGetLastError());
    e->launches += (n_w_max != 0) + (n_x_max != 0);
    if (n_ord > 1) {
        // Two orders of the records for the host reduce, computed here where sorting is cheap: by the
        // word's first 16 bytes (two stable radix passes: low half, then high half) and by first offset.
        if (n_ord >= 0x7FFFFFF0ull) { e->err = "too many distinct words for wc_finish"; return WC_ERR_TOO_LARGE; }
        const int n = (int)n_ord;
        size_t tb = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tb, (const u64 *)nullptr, (u64 *)nullptr, (const u32 *)nullptr,
                                           (u32 *)nullptr, n, 0, 64, e->stream));
        const size_t a8 = align_up(n_ord * sizeof(u64), 256), a4 = align_up(n_ord * sizeof(u32), 256);
        const size_t need_ord = 5 * a8 + 2 * a4 + align_up(tb, 256);
        if (need_ord > e->d_ord_cap) {
            if (e->d_ord) cudaFree(e->d_ord);
            e->d_ord = nullptr; e->d_ord_cap = 0;
            CK(cudaMalloc(&e->d_ord, need_ord + need_ord / 4));
            e->d_ord_cap = need_ord + need_ord / 4;
        }
        uint8_t *q = (uint8_t *)e->d_ord;
        u64 *hi = (u64 *)q, *lo = (u64 *)(q + a8), *fst = (u64 *)(q + 2 * a8), *k_out = (u64 *)(q + 3 * a8),
            *hi2 = (u64 *)(q + 4 * a8);
        u32 *iota = (u32 *)(q + 5 * a8), *by_lo = (u32 *)(q + 5 * a8 + a4);
        void *temp = q + 5 * a8 + 2 * a4;
        wc::order_keys_kernel<<<e->num_sms * 2, 256, 0, e->stream>>>(
            (const wc::WordRecord *)(dfin + off_w), n_w_max, e->d_cursors + kMaxParts, (const wc::LongRecord *)(dfin + off_x),
            n_x_max, e->d_cursors + 2 * kMaxParts, (const u64 *)(dfin + off_b), hi, lo, fst, iota);
        size_t tbx = tb;
        CK(cub::DeviceRadixSort::SortPairs(temp, tbx, (const u64 *)fst, k_out, (const u32 *)iota, (u32 *)(dfin + off_pf), n, 0,
                                           64, e->stream));
        tbx = tb;
        CK(cub::DeviceRadixSort::SortPairs(temp, tbx, (const u64 *)lo, k_out, (const u32 *)iota, by_lo, n, 0, 64, e->stream));
        wc::gather_u64_kernel<<<e->num_sms * 2, 256, 0, e->stream>>>(hi, by_lo, n_ord, hi2);
        tbx = tb;
        CK(cub::DeviceRadixSort::SortPairs(temp, tbx, (const u64 *)hi2, k_out, (const u32 *)by_lo, (u32 *)(dfin + off_pk), n, 0,
                                           64, e->stream)); // k_out = the sorted high halves
        CK(cudaGetLastError());
        e->launches += 5;

        if (!host_reduce) {
            // ---- the reduce itself (wc_reduce.cuh): owners by range-minimum queries in key order, counts
            // to the owners, rows compacted in order of first occurrence
            const size_t n1 = (n_ord + wc::kRmqBlock - 1) / wc::kRmqBlock, n2 = (n1 + wc::kRmqBlock - 1) / wc::kRmqBlock;
            size_t tb_sel = 0, tb_scan = 0;
            CK(cub::DeviceSelect::Flagged(nullptr, tb_sel, (const u32 *)nullptr, (const unsigned char *)nullptr, (u32 *)nullptr,
                                          (u64 *)nullptr, n, e->stream));
            CK(cub::DeviceScan::ExclusiveSum(nullptr, tb_scan, (const u64 *)nullptr, (u64 *)nullptr, n, e->stream));
            if (tb_scan > tb_sel) tb_sel = tb_scan;
            const size_t b1 = align_up(n1 * sizeof(u64), 256), b2 = align_up(n2 * sizeof(u64), 256);
            const size_t need_fz = 3 * a8 + 3 * a4 + align_up(n_ord, 256) + 2 * b1 + 2 * b2 + align_up(tb_sel, 256) + 256;
            if (need_fz > e->d_fz_cap) {
                if (e->d_fz) cudaFree(e->d_fz);
                e->d_fz = nullptr; e->d_fz_cap = 0;
                CK(cudaMalloc(&e->d_fz, need_fz + need_fz / 4));
                e->d_fz_cap = need_fz + need_fz / 4;
            }
            uint8_t *z = (uint8_t *)e->d_fz;
            wc::ReduceArrays A;
            A.w = (const wc::WordRecord *)(dfin + off_w); A.n_w_max = n_w_max; A.n_w_ptr = e->d_cursors + kMaxParts;
            A.x = (const wc::LongRecord *)(dfin + off_x); A.n_x_max = n_x_max; A.n_x_ptr = e->d_cursors + 2 * kMaxParts;
            A.blob = (const u64 *)(dfin + off_b);
            A.by_key = (const u32 *)(dfin + off_pk); A.by_first = (const u32 *)(dfin + off_pf);
            A.n = n_ord;
            A.s_hi = k_out;
            A.s_lo = hi2;                       // (free again after the last sort)
            A.s_first = (u64 *)z; A.s_count = (u64 *)(z + a8); A.acc = (u64 *)(z + 2 * a8);
            A.s_len = (u32 *)(z + 3 * a8); A.pos_of_idx = (u32 *)(z + 3 * a8 + a4); A.owner = (u32 *)(z + 3 * a8 + 2 * a4);
            unsigned char *flags = z + 3 * a8 + 3 * a4;
            uint8_t *zz = flags + align_up(n_ord, 256);
            A.bmin1 = (u64 *)zz; A.bpos1 = (u32 *)(zz + b1); A.bmin2 = (u64 *)(zz + 2 * b1); A.bpos2 = (u32 *)(zz + 2 * b1 + b2);
            void *sel_temp = zz + 2 * b1 + 2 * b2;
            A.counters = (u64 *)(zz + 2 * b1 + 2 * b2 + align_up(tb_sel, 256));
            CK(cudaMemsetAsync(A.counters, 0, 4 * sizeof(u64), e->stream));
            // the output block: [64-byte head][rows, room for n_ord][the rows' words, packed]
            const size_t out_rows = 64, out_blob = out_rows + n_ord * sizeof(wc::RowOut), out_need = out_blob + n_w_max * 16 + blob_max;
            if (out_need > e->d_out_cap) {
                if (e->d_out) cudaFree(e->d_out);
                e->d_out = nullptr; e->d_out_cap = 0;
                CK(cudaMalloc(&e->d_out, out_need + out_need / 4));
                e->d_out_cap = out_need + out_need / 4;
            }
            uint8_t *dout = (uint8_t *)e->d_out;
            const int gb = e->num_sms * 8;
            wc::reduce_gather_kernel<<<gb, 256, 0, e->stream>>>(A, lo);
            wc::reduce_blockmin_kernel<<<(unsigned)(n1 < 65535 ? n1 : 65535), wc::kRmqBlock, 0, e->stream>>>(A.s_first, nullptr, n_ord, A.bmin1, A.bpos1);
            wc::reduce_blockmin_kernel<<<(unsigned)(n2 < 65535 ? n2 : 65535), wc::kRmqBlock, 0, e->stream>>>(A.bmin1, A.bpos1, n1, A.bmin2, A.bpos2);
            wc::reduce_owner_kernel<<<gb, 256, 0, e->stream>>>(A, e->cfg.mode == WC_MODE_EXACT);
            wc::reduce_rowflags_kernel<<<gb, 256, 0, e->stream>>>(A, flags);
            size_t tbs = tb_sel;
            CK(cub::DeviceSelect::Flagged(sel_temp, tbs, A.by_first, (const unsigned char *)flags, (u32 *)(dfin + off_ri),
                                          A.counters + 1, n, e->stream));
            // the rows as the host wants them: word lengths -> prefix sum -> rows + packed words
            u64 *lens = A.s_count, *offs = (u64 *)(dfin + off_rc); // (s_count is not needed any more; off_rc: n_ord u64)
            wc::reduce_rowlens_kernel<<<gb, 256, 0, e->stream>>>(A, (const u32 *)(dfin + off_ri), A.counters + 1, lens);
            tbs = tb_sel;
            CK(cub::DeviceScan::ExclusiveSum(sel_temp, tbs, (const u64 *)lens, offs, n, e->stream));
            wc::reduce_emit_rows_kernel<<<gb, 256, 0, e->stream>>>(A, (const u32 *)(dfin + off_ri), A.counters + 1, offs,
                                                                   (wc::RowOut *)(dout + out_rows), dout + out_blob);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(dout, A.counters, 4 * sizeof(u64), cudaMemcpyDeviceToDevice, e->stream));
            e->launches += 9;
        }
    }
    // cursors into the head of the block: [0] word records, [1] long records, [2] blob bytes
    CK(cudaMemcpyAsync(dfin, e->d_cursors + kMaxParts, sizeof(u64), cudaMemcpyDeviceToDevice, e->stream));
    CK(cudaMemcpyAsync(dfin + 8, e->d_cursors + 2 * kMaxParts, 2 * sizeof(u64), cudaMemcpyDeviceToDevice, e->stream));
    const bool device_reduce = n_ord > 1 && !host_reduce;
    if (device_reduce) {
        static_assert(sizeof(wc::RowOut) == sizeof(wc::Row) && offsetof(wc::Row, len) == 8 && offsetof(wc::Row, count) == 16 &&
                          offsetof(wc::Row, first) == 24, "RowOut must mirror wc::Row");
        const size_t out_rows = 64, out_blob = out_rows + n_ord * sizeof(wc::RowOut), out_need = out_blob + n_w_max * 16 + blob_max;
        if (out_need > e->h_out_cap) {
            if (e->h_out) cudaFreeHost(e->h_out);
            e->h_out = nullptr; e->h_out_cap = 0;
            CK(cudaMallocHost(&e->h_out, out_need + out_need / 4));
            e->h_out_cap = out_need + out_need / 4;
        }
        uint8_t *dout = (uint8_t *)e->d_out, *hout = (uint8_t *)e->h_out;
        // the head first (how many rows, how many bytes of words), then exactly those
        CK(cudaMemcpyAsync(hout, dout, 64, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        const u64 n_valid = ((const u64 *)hout)[0], n_rows = ((const u64 *)hout)[1], sum_counts = ((const u64 *)hout)[2],
                  words_bytes = n_rows ? ((const u64 *)hout)[3] : 0;
        if (n_rows > n_ord || words_bytes > n_w_max * 16 + blob_max) { e->err = "row counters are inconsistent"; return WC_ERR_CUDA; }
        if (n_rows) {
            CK(cudaMemcpyAsync(hout + out_rows, dout + out_rows, n_rows * sizeof(wc::RowOut), cudaMemcpyDeviceToHost, e->stream));
            CK(cudaMemcpyAsync(hout + out_blob, dout + out_blob, words_bytes, cudaMemcpyDeviceToHost, e->stream));
            CK(cudaStreamSynchronize(e->stream));
        }
        e->finish_d2h = 64 + n_rows * sizeof(wc::RowOut) + words_bytes + sizeof(wc::Status);
        const double t_d2h = since();
        wc_result *r = new (std::nothrow) wc_result();
        if (!r) return WC_ERR_NOMEM;
        const wc::Row *rp = (const wc::Row *)(hout + out_rows);
        wc::Row *row0 = nullptr;
        const char *big_env = getenv("WC_BIG_RESULT_BYTES"); // (tests lower the threshold to reach this path)
        const size_t big_at = big_env ? (size_t)strtoull(big_env, nullptr, 10) : ((size_t)256 << 20);
        if (n_rows * sizeof(wc::Row) + words_bytes >= big_at) { // large: uninitialised arrays, filled by several threads
            r->big_rows = (wc::Row *)malloc(n_rows * sizeof(wc::Row));
            r->big_blob = (char *)malloc(words_bytes ? words_bytes : 1);
            if (!r->big_rows || !r->big_blob) { delete r; return WC_ERR_NOMEM; }
            r->n_big_rows = n_rows;
            wc::parallel_copy(r->big_rows, rp, n_rows * sizeof(wc::Row));
            wc::parallel_copy(r->big_blob, hout + out_blob, words_bytes);
            row0 = r->big_rows;
        } else {
            r->rows.assign(rp, rp + n_rows);
            r->blob.assign((const char *)(hout + out_blob), (size_t)words_bytes);
            row0 = r->rows.data();
        }
        const double t_decode = since();
        const u64 total_bytes = e->job_stats_set ? e->job_bytes : e->consumed;
        const u64 stop_abs = st.stop == wc::kNone ? e->cfg.base_offset + total_bytes : st.stop;
        u64 lost = st.lost_empty;
        if (!st.p0_found || st.p0 >= stop_abs) lost = st.n_empty; // no non-empty word at all: every empty one is lost
        r->total = (int64_t)sum_counts; // every word is credited to exactly one row
        if (e->cfg.mode != WC_MODE_EXACT && n_rows) { // empty words after the first real one go to row 0 (main.cu:81-86)
            row0[0].count += (int64_t)(st.n_empty - lost);
            r->total += (int64_t)(st.n_empty - lost);
        }
        r->stats.bytes = total_bytes;
        r->stats.n_flush = st.n_flush; r->stats.n_empty = st.n_empty; r->stats.n_lost_empty = lost;
        r->stats.stop = stop_abs - e->cfg.base_offset;
        r->stats.n_distinct = n_valid;
        r->stats.n_long_words = st.n_long; r->stats.n_spilled = st.n_spill; r->stats.n_slow_tiles = st.n_slow_tiles;
        r->stats.table_grows = e->grows; r->stats.device_ms = e->device_ms; r->stats.count_kernel_ms = e->count_kernel_ms;
        note_job_path(e, &r->stats);
        if (timing)
            fprintf(stderr, "[wc] wc_finish: drain %.2f settle %.2f export+order+reduce+copy %.2f rows %.2f ms (%llu rows of %llu words)\n",
                    t_sync, t_settle - t_sync, t_d2h - t_settle, t_decode - t_d2h, (unsigned long long)n_rows,
                    (unsigned long long)n_valid);
        *out = r;
        return WC_OK;
    }
    CK(cudaMemcpyAsync(hfin, dfin, copy_bytes, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->finish_d2h = copy_bytes + sizeof(wc::Status);
    const double t_d2h = since();
    const u64 n_w = ((const u64 *)hfin)[0], n_x = ((const u64 *)hfin)[1];
    if (n_w > n_w_max || n_x > n_x_max) { e->err = "table entry counters are inconsistent"; return WC_ERR_CUDA; }

    // Host reduce (sort path, a single entry, or WC_HOST_REDUCE=1).  Entries keep the record index space
    // of the device order (word i -> i, long j -> n_w_max + j); slots without a record, and words whose
    // count went back to zero, stay as count-0 entries.
    const bool have_order = n_ord > 1;
    std::string blob;
    std::vector<wc::HostEntry> entries;
    blob.reserve((size_t)n_w * 8 + blob_max);
    const wc::WordRecord *wr = (const wc::WordRecord *)(hfin + off_w);
    const wc::LongRecord *xr = (const wc::LongRecord *)(hfin + off_x);
    if (have_order) entries.assign(n_ord, wc::HostEntry{0, 0, 0, ~0ull});
    else entries.reserve((size_t)(n_w + n_x));
    for (size_t i = 0; i < n_w; i++) {
        const u64 k[2] = {wr[i].k0, wr[i].k1};
        if ((int64_t)wr[i].count == 0) continue;
        const u32 len = k[1] ? 16 - (u32)(__builtin_clzll(k[1]) >> 3) : 8 - (u32)(__builtin_clzll(k[0]) >> 3);
        const wc::HostEntry he{blob.size(), len, (int64_t)wr[i].count, wr[i].first};
        if (have_order) entries[i] = he; else entries.push_back(he);
        blob.append((const char *)k, len); // the word itself (little-endian words: byte order is text order)
    }
    if (e->sort_mode) {
        u64 n_sorted = 0;
        rc = finish_sorted(e, st.stop, blob, entries, &n_sorted);
        if (rc != WC_OK) return rc;
    }
    for (size_t i = 0; i < n_x; i++) {
        if ((int64_t)xr[i].count == 0) continue;
        const wc::HostEntry he{blob.size(), (u32)xr[i].len, (int64_t)xr[i].count, xr[i].first};
        if (have_order) entries[n_w_max + i] = he; else entries.push_back(he);
        blob.append((const char *)(hfin + off_b + xr[i].blob_off), (size_t)xr[i].len);
    }
    size_t n_distinct = 0;
    for (const wc::HostEntry &he : entries) n_distinct += he.count != 0;

    const double t_decode = since();
    wc_result *r = new (std::nothrow) wc_result();
    if (!r) return WC_ERR_NOMEM;
    const u64 total_bytes = e->job_stats_set ? e->job_bytes : e->consumed;
    const u64 stop_abs = st.stop == wc::kNone ? e->cfg.base_offset + total_bytes : st.stop;
    u64 lost = st.lost_empty;
    if (!st.p0_found || st.p0 >= stop_abs) lost = st.n_empty; // no non-empty word at all: every empty one is lost
    r->stats.bytes = total_bytes;
    r->stats.n_flush = st.n_flush;
    r->stats.n_empty = st.n_empty;
    r->stats.n_lost_empty = lost;
    r->stats.stop = stop_abs - e->cfg.base_offset;
    r->stats.n_distinct = n_distinct;
    r->stats.n_long_words = st.n_long;
    r->stats.n_spilled = st.n_spill;
    r->stats.n_slow_tiles = st.n_slow_tiles;
    r->stats.table_grows = e->grows;
    r->stats.device_ms = e->device_ms;
    r->stats.count_kernel_ms = e->count_kernel_ms;
    note_job_path(e, &r->stats);
    wc::finalize_rows(std::move(blob), entries, e->cfg.mode, st.n_empty, lost, r,
                      have_order ? (const uint32_t *)(hfin + off_pk) : nullptr,
                      have_order ? (const uint32_t *)(hfin + off_pf) : nullptr);
    if (timing)
        fprintf(stderr, "[wc] wc_finish: drain %.2f settle %.2f export+order+copy %.2f decode %.2f reduce %.2f ms\n", t_sync,
                t_settle - t_sync, t_d2h - t_settle, t_decode - t_d2h, since() - t_decode);
    *out = r;
    return WC_OK;
}

} // extern "C"

// Test hook (no GPU needed): the SWAR byte classification and flag->mask compaction of the
// tokenizer kernel, run on the host from the same source (classify4 / movemask16 in wc_device.cuh).
extern "C" int wc_debug_classify16(const uint8_t *bytes16, uint32_t *delim_mask, uint32_t *lf_mask,
                                   uint32_t *kill_mask) {
    if (!bytes16 || !delim_mask || !lf_mask || !kill_mask) return WC_ERR_INVALID;
    u32 w[4];
    memcpy(w, bytes16, 16);
    wc::ByteFlags f0 = wc::classify4(w[0]), f1 = wc::classify4(w[1]), f2 = wc::classify4(w[2]), f3 = wc::classify4(w[3]);
    *delim_mask = wc::movemask16(f0.delim, f1.delim, f2.delim, f3.delim);
    *lf_mask = wc::movemask16(f0.lf, f1.lf, f2.lf, f3.lf);
    *kill_mask = wc::movemask16(f0.kill, f1.kill, f2.kill, f3.kill);
    return WC_OK;
}

extern "C" uint32_t wc_debug_smem_entries(void) { return wc::kSmemEntries; }

// WC_PROFILE builds: the cycle counters of the last job (call after wc_sync / wc_finish)
extern "C" int wc_debug_counters(wc_engine *e, uint64_t *out8) {
    if (!e || !out8) return WC_ERR_INVALID;
    wc::Status st;
    if (cudaMemcpy(&st, e->d_status, sizeof(st), cudaMemcpyDeviceToHost) != cudaSuccess) return WC_ERR_CUDA;
    for (int i = 0; i < 8; i++) out8[i] = st.dbg[i];
    return WC_OK;
}

extern "C" uint32_t wc_debug_smem_slot(uint64_t key) {
    return (uint32_t)(((uint64_t)wc::hash32((u32)key, (u32)(key >> 32)) * wc::kSmemEntries) >> 32);
}


// ------------------------------------------------------------------------------------------------
// Multi-GPU: shard counters and the key-owner exchange.  One engine per GPU counts
// its byte range (wc_config.base_offset = where the range starts); the caller moves the exported
// buffers between GPUs (NCCL all-to-all / gather through torch.distributed or any other transport).

static int ensure_cursors(wc_engine *e) {
    if (!e->d_cursors) CK(cudaMalloc((void **)&e->d_cursors, (2 * kMaxParts + 2) * sizeof(u64)));
    return WC_OK;
}

extern "C" {

int wc_set_base_offset(wc_engine *e, uint64_t base_offset) {
    if (!e) return WC_ERR_INVALID;
    if (e->fed != 0 || e->consumed != 0) { e->err = "wc_set_base_offset after input was fed"; return WC_ERR_INVALID; }
    e->cfg.base_offset = base_offset;
    return WC_OK;
}

int wc_shard_stats_get(wc_engine *e, wc_shard_stats *out) {
    if (!e || !out) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    { int rc0 = wc_end_input(e); if (rc0 != WC_OK) return rc0; } // a shard's counters describe all of its bytes
    int rc = wc_sync(e);
    if (rc != WC_OK) return rc;
    rc = settle_tables(e);
    if (rc != WC_OK) return rc;
    const wc::Status st = *e->h_status;
    if (st.err) { e->err = "device tables full"; return WC_ERR_TABLE_FULL; }
    memset(out, 0, sizeof(*out));
    out->base = e->cfg.base_offset; out->bytes = e->consumed;
    out->n_flush = st.n_flush; out->n_empty = st.n_empty; out->lost_empty = st.lost_empty;
    out->p0 = st.p0; out->p0_found = st.p0_found;
    out->stop = st.stop; // absolute, ~0 if this shard saw no reason to stop
    return WC_OK;
}

// Pure host logic: what the counters of the whole job are, given every shard's (in input order).
// A shard at or after the first STOP contributes nothing (main.cu:185: the reference stops reading);
// the shard holding STOP has already rolled its own tail back.  Empty words count as lost until the
// first non-empty word of the whole job (main.cu:81,103).
int wc_shard_stats_reduce(const wc_shard_stats *s, size_t n, wc_shard_stats *out) {
    if (!s || !out || n == 0) return WC_ERR_INVALID;
    for (size_t i = 1; i < n; i++)
        if (s[i].base != s[i - 1].base + s[i - 1].bytes) return WC_ERR_INVALID; // shards must tile the input
    wc_shard_stats j;
    memset(&j, 0, sizeof(j));
    j.base = s[0].base; j.stop = ~0ull; j.p0 = ~0ull;
    for (size_t i = 0; i < n; i++) {
        j.bytes += s[i].bytes;
        if (j.stop != ~0ull) continue; // after STOP: ignored
        j.n_flush += s[i].n_flush; j.n_empty += s[i].n_empty;
        if (!j.p0_found) {
            if (s[i].p0_found && s[i].p0 < s[i].stop) { j.p0_found = 1; j.p0 = s[i].p0; j.lost_empty += s[i].lost_empty; }
            else j.lost_empty += s[i].n_empty;
        }
        if (s[i].stop != ~0ull) j.stop = s[i].stop;
    }
    *out = j;
    return WC_OK;
}

int wc_shard_stats_set(wc_engine *e, const wc_shard_stats *job) {
    if (!e || !job) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaMemcpyAsync(e->h_status, e->d_status, sizeof(wc::Status), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    wc::Status st = *e->h_status;
    st.n_flush = job->n_flush; st.n_empty = job->n_empty; st.lost_empty = job->lost_empty;
    st.p0 = job->p0; st.p0_found = job->p0_found; st.stop = job->stop;
    *e->h_status = st;
    CK(cudaMemcpyAsync(e->d_status, e->h_status, sizeof(st), cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->job_stats_set = true; e->job_bytes = job->bytes;
    return WC_OK;
}

int wc_clear_tables(wc_engine *e) {
    if (!e) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaMemsetAsync(e->table_block, 0, e->table_block_bytes, e->stream));
    CK(cudaMemcpyAsync(e->h_status, e->d_status, sizeof(wc::Status), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->h_status->w_entries = 0; e->h_status->x_entries = 0; e->h_status->arena_used = 1;
    CK(cudaMemcpyAsync(e->d_status, e->h_status, sizeof(wc::Status), cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return WC_OK;
}

int wc_export_counts(wc_engine *e, uint32_t n_parts, uint64_t *word_counts, uint64_t *long_entries,
                     uint64_t *long_blob_bytes) {
    if (!e || !word_counts || n_parts == 0 || n_parts > kMaxParts) return WC_ERR_INVALID;
    if (e->sort_mode) { e->err = "the key-owner exchange needs the hash path (wc_config.path = WC_PATH_HASH)"; return WC_ERR_INVALID; }
    CK(cudaSetDevice(e->device));
    int rc = ensure_cursors(e);
    if (rc != WC_OK) return rc;
    CK(cudaMemsetAsync(e->d_cursors, 0, (2 * kMaxParts + 2) * sizeof(u64), e->stream));
    wc::export_count_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(e->T, n_parts, e->d_cursors);
    wc::count_long_kernel<<<e->num_sms, 256, 0, e->stream>>>(e->T, e->d_cursors + 2 * kMaxParts);
    CK(cudaGetLastError());
    u64 host[2 * kMaxParts + 2];
    CK(cudaMemcpyAsync(host, e->d_cursors, sizeof(host), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    for (u32 p = 0; p < n_parts; p++) word_counts[p] = host[p];
    if (long_entries) *long_entries = host[2 * kMaxParts];
    if (long_blob_bytes) *long_blob_bytes = host[2 * kMaxParts + 1];
    e->launches += 2;
    return WC_OK;
}

int wc_export_words(wc_engine *e, uint32_t n_parts, const uint64_t *part_offsets, void *dev_records) {
    if (!e || !part_offsets || !dev_records || n_parts == 0 || n_parts > kMaxParts) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    int rc = ensure_cursors(e);
    if (rc != WC_OK) return rc;
    u64 host[2 * kMaxParts];
    memset(host, 0, sizeof(host));
    for (u32 p = 0; p < n_parts; p++) host[p] = part_offsets[p]; // [0, kMaxParts): offsets, [kMaxParts, ..): cursors
    CK(cudaMemcpyAsync(e->d_cursors, host, sizeof(host), cudaMemcpyHostToDevice, e->stream));
    wc::export_scatter_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(e->T, n_parts, e->d_cursors, e->d_cursors + kMaxParts,
                                                                    (wc::WordRecord *)dev_records);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(e->stream)); // `host` is on this stack frame
    e->launches += 1;
    return WC_OK;
}

int wc_export_long(wc_engine *e, void *dev_meta, void *dev_blob) {
    if (!e || !dev_meta || !dev_blob) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    int rc = ensure_cursors(e);
    if (rc != WC_OK) return rc;
    CK(cudaMemsetAsync(e->d_cursors + 2 * kMaxParts, 0, 2 * sizeof(u64), e->stream));
    wc::export_long_kernel<<<e->num_sms, 256, 0, e->stream>>>(e->T, e->d_cursors + 2 * kMaxParts, (wc::LongRecord *)dev_meta,
                                                             (u64 *)dev_blob);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(e->stream));
    e->launches += 1;
    return WC_OK;
}

int wc_import_words(wc_engine *e, const void *dev_records, uint64_t n_records) {
    if (!e || (!dev_records && n_records)) return WC_ERR_INVALID;
    if (n_records == 0) return WC_OK;
    CK(cudaSetDevice(e->device));
    wc::import_words_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(e->T, (const wc::WordRecord *)dev_records, n_records);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(e->stream));
    e->launches += 1;
    return WC_OK;
}

int wc_import_long(wc_engine *e, const void *dev_meta, uint64_t n_entries, const void *dev_blob) {
    if (!e || ((!dev_meta || !dev_blob) && n_entries)) return WC_ERR_INVALID;
    if (n_entries == 0) return WC_OK;
    CK(cudaSetDevice(e->device));
    wc::import_long_kernel<<<e->num_sms, 256, 0, e->stream>>>(e->T, (const wc::LongRecord *)dev_meta, n_entries,
                                                             (const uint8_t *)dev_blob);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(e->stream));
    e->launches += 1;
    return WC_OK;
}

// ---- peer gather: entries written straight into the root GPU's memory over NVLink (wc_shuffle.cuh)

static u64 peer_stride(u64 cap_words, u64 cap_long, u64 cap_blob) {
    return align_up(sizeof(wc::PeerHeader) + cap_words * sizeof(wc::WordRecord) + cap_long * sizeof(wc::LongRecord) +
                        align_up(cap_blob, 8), 256);
}

// The root's block is [PeerCtl][rank 0: area 0, area 1][rank 1: area 0, area 1]...; job `seq` uses area seq & 1.
static wc::PeerArea peer_area(uint8_t *block, const wc_engine::Peer &p, u32 rank, u64 seq) {
    uint8_t *a = block + sizeof(wc::PeerCtl) + (2ull * rank + (seq & 1)) * p.stride;
    wc::PeerArea A;
    A.hdr = (wc::PeerHeader *)a;
    A.words = (wc::WordRecord *)(a + sizeof(wc::PeerHeader));
    A.longs = (wc::LongRecord *)(a + sizeof(wc::PeerHeader) + p.cap_words * sizeof(wc::WordRecord));
    A.blob = (u64 *)(a + sizeof(wc::PeerHeader) + p.cap_words * sizeof(wc::WordRecord) + p.cap_long * sizeof(wc::LongRecord));
    A.cap_words = p.cap_words; A.cap_long = p.cap_long; A.cap_blob_units = p.cap_blob / 8;
    return A;
}

int wc_peer_root_create(wc_engine *e, uint32_t n_ranks, uint32_t root_rank, uint64_t words_per_rank, uint64_t long_per_rank,
                        uint64_t blob_bytes_per_rank, void *handle_out) {
    if (!e || n_ranks == 0 || n_ranks > kMaxParts || root_rank >= n_ranks || words_per_rank == 0) return WC_ERR_INVALID;
    if (e->peer.area || e->peer.mapped) { e->err = "peer gather is already set up on this engine"; return WC_ERR_INVALID; }
    if (e->sort_mode) { e->err = "peer gather needs the hash path"; return WC_ERR_INVALID; }
    e->auto_path = false; // (the gather moves table entries: the engines of a multi-GPU job stay on the hash path)
    CK(cudaSetDevice(e->device));
    wc_engine::Peer p;
    p.n_ranks = n_ranks; p.root_rank = root_rank; p.my_rank = root_rank;
    p.cap_words = words_per_rank; p.cap_long = long_per_rank; p.cap_blob = align_up(blob_bytes_per_rank, 8);
    p.stride = peer_stride(p.cap_words, p.cap_long, p.cap_blob);
    CK(cudaMalloc((void **)&p.area, sizeof(wc::PeerCtl) + 2 * p.stride * n_ranks));
    CK(cudaMemset(p.area, 0, sizeof(wc::PeerCtl) + 2 * p.stride * n_ranks));
    CK(cudaMallocHost((void **)&p.h_hdr, sizeof(wc::PeerHeader) * n_ranks));
    CK(cudaMalloc((void **)&p.d_job, sizeof(wc::PeerJob)));
    if (handle_out) {
        static_assert(sizeof(cudaIpcMemHandle_t) == WC_PEER_HANDLE_BYTES, "handle size");
        cudaIpcMemHandle_t h;
        CK(cudaIpcGetMemHandle(&h, p.area));
        memcpy(handle_out, &h, sizeof(h));
    }
    e->peer = p;
    return WC_OK;
}

static int peer_sender_init(wc_engine *e, uint32_t my_rank, uint32_t n_ranks, uint32_t root_rank, uint64_t words_per_rank,
                            uint64_t long_per_rank, uint64_t blob_bytes_per_rank) {
    if (n_ranks == 0 || n_ranks > kMaxParts || my_rank >= n_ranks || root_rank >= n_ranks || my_rank == root_rank)
        return WC_ERR_INVALID;
    if (e->peer.area || e->peer.mapped) { e->err = "peer gather is already set up on this engine"; return WC_ERR_INVALID; }
    if (e->sort_mode) { e->err = "peer gather needs the hash path"; return WC_ERR_INVALID; }
    e->auto_path = false;
    wc_engine::Peer &p = e->peer;
    p.n_ranks = n_ranks; p.root_rank = root_rank; p.my_rank = my_rank;
    p.cap_words = words_per_rank; p.cap_long = long_per_rank; p.cap_blob = align_up(blob_bytes_per_rank, 8);
    p.stride = peer_stride(p.cap_words, p.cap_long, p.cap_blob);
    CK(cudaMalloc((void **)&p.d_cur, 4 * sizeof(u64)));
    return WC_OK;
}

int wc_peer_open(wc_engine *e, uint32_t my_rank, uint32_t n_ranks, uint32_t root_rank, uint64_t words_per_rank,
                 uint64_t long_per_rank, uint64_t blob_bytes_per_rank, const void *handle) {
    if (!e || !handle) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    int rc = peer_sender_init(e, my_rank, n_ranks, root_rank, words_per_rank, long_per_rank, blob_bytes_per_rank);
    if (rc != WC_OK) return rc;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *ptr = nullptr;
    CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    e->peer.mapped = (uint8_t *)ptr; e->peer.ipc = true;
    return WC_OK;
}

int wc_peer_attach(wc_engine *e, uint32_t my_rank, wc_engine *root) {
    if (!e || !root || e == root || !root->peer.area) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    const wc_engine::Peer &r = root->peer;
    int rc = peer_sender_init(e, my_rank, r.n_ranks, r.root_rank, r.cap_words, r.cap_long, r.cap_blob);
    if (rc != WC_OK) return rc;
    if (root->device != e->device) {
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, e->device, root->device));
        if (!can) { e->err = "no peer access to the root's GPU"; return WC_ERR_CUDA; }
        cudaError_t pe = cudaDeviceEnablePeerAccess(root->device, 0);
        if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) CK(pe);
        (void)cudaGetLastError();
    }
    e->peer.mapped = r.area; e->peer.ipc = false;
    return WC_OK;
}

int wc_peer_push(wc_engine *e, void *dev_flag) {
    Range nvtx_range("wc_peer_push");
    if (!e || !e->peer.mapped) return WC_ERR_INVALID;
    if (e->sort_mode) { e->err = "peer gather needs the hash path"; return WC_ERR_INVALID; }
    CK(cudaSetDevice(e->device));
    int rc = wc_end_input(e);
    if (rc != WC_OK) return rc;
    const wc_engine::Peer &p = e->peer;
    CK(cudaMemsetAsync(p.d_cur, 0, 4 * sizeof(u64), e->stream));
    e->peer.seq++;
    wc::peer_push_kernel<<<e->num_sms * 2, 256, 0, e->stream>>>(e->T, peer_area(p.mapped, p, p.my_rank, p.seq),
                                                               (const wc::PeerCtl *)p.mapped, e->cfg.base_offset, e->consumed,
                                                               p.seq, p.d_cur, (u32 *)dev_flag);
    CK(cudaGetLastError());
    e->launches += 1;
    return WC_OK;
}

int wc_peer_pull(wc_engine *e, wc_shard_stats *job_out) {
    Range nvtx_range("wc_peer_pull");
    if (!e || !e->peer.area) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    wc_engine::Peer &p = e->peer;
    p.seq++;
    int rc = wc_end_input(e);
    if (rc != WC_OK) return rc;
    // one wait for everything the host needs: this engine's counters and every sender's header
    CK(cudaMemcpyAsync(e->h_status, e->d_status, sizeof(wc::Status), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpy2DAsync(p.h_hdr, sizeof(wc::PeerHeader), p.area + sizeof(wc::PeerCtl) + (p.seq & 1) * p.stride, 2 * p.stride,
                         sizeof(wc::PeerHeader), p.n_ranks, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    wc_shard_stats mine;
    if (e->h_status->ovf_count != 0 || e->h_status->err != 0) { // rare: grow the table / replay its overflow log first
        rc = wc_shard_stats_get(e, &mine);
        if (rc != WC_OK) return rc;
    } else {
        const wc::Status &st = *e->h_status;
        memset(&mine, 0, sizeof(mine));
        mine.base = e->cfg.base_offset; mine.bytes = e->consumed;
        mine.n_flush = st.n_flush; mine.n_empty = st.n_empty; mine.lost_empty = st.lost_empty;
        mine.p0 = st.p0; mine.p0_found = st.p0_found; mine.stop = st.stop;
    }
    std::vector<wc_shard_stats> shards(p.n_ranks);
    for (u32 r = 0; r < p.n_ranks; r++) {
        if (r == p.root_rank) { shards[r] = mine; continue; }
        const wc::PeerHeader &h = p.h_hdr[r];
        if (h.seq != p.seq) { e->err = "peer gather: a rank's push has not arrived (call after the ranks' barrier)"; return WC_ERR_INVALID; }
        if (h.flags) { e->err = "peer gather: a sender's tables or its landing area overflowed"; return WC_ERR_TOO_LARGE; }
        wc_shard_stats s;
        memset(&s, 0, sizeof(s));
        s.base = h.base; s.bytes = h.bytes; s.n_flush = h.n_flush; s.n_empty = h.n_empty; s.lost_empty = h.lost_empty;
        s.p0 = h.p0; s.stop = h.stop; s.p0_found = (u32)h.p0_found;
        shards[r] = s;
    }
    wc_shard_stats job;
    rc = wc_shard_stats_reduce(shards.data(), shards.size(), &job);
    if (rc != WC_OK) { e->err = "peer gather: the ranks' shards do not tile the input"; return rc; }
    auto dead = [&job](const wc_shard_stats &s) { return job.stop != ~0ull && s.base >= job.stop && s.bytes > 0; };
    if (dead(mine)) { rc = wc_clear_tables(e); if (rc != WC_OK) return rc; }
    wc::ImportList L;
    memset(&L, 0, sizeof(L));
    for (u32 r = 0; r < p.n_ranks; r++) {
        if (r == p.root_rank || dead(shards[r])) continue;
        const wc::PeerHeader &h = p.h_hdr[r];
        if (h.n_words == 0 && h.n_long == 0) continue;
        const wc::PeerArea A = peer_area(p.area, p, r, p.seq);
        L.w[L.n] = A.words; L.n_w[L.n] = h.n_words;
        L.x[L.n] = A.longs; L.n_x[L.n] = h.n_long;
        L.blob[L.n] = (const uint8_t *)A.blob;
        L.n++;
    }
    if (L.n) {
        wc::import_areas_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(e->T, L);
        e->launches += 1;
    }
    // the root's counters become the job's (stream ordered; wc_finish reads them after its own sync)
    wc::set_job_counters_kernel<<<1, 1, 0, e->stream>>>(e->d_status, job.n_flush, job.n_empty, job.lost_empty, job.p0, job.stop,
                                                       job.p0_found);
    wc::peer_done_kernel<<<1, 1, 0, e->stream>>>((wc::PeerCtl *)p.area, p.seq); // the senders may reuse this job's areas
    CK(cudaGetLastError());
    e->launches += 2;
    e->job_stats_set = true; e->job_bytes = job.bytes;
    if (job_out) *job_out = job;
    return WC_OK;
}

int wc_peer_pull_device(wc_engine *e) {
    Range nvtx_range("wc_peer_pull_device");
    if (!e || !e->peer.area) return WC_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    wc_engine::Peer &p = e->peer;
    p.seq++;
    int rc = wc_end_input(e);
    if (rc != WC_OK) return rc;
    uint8_t *area0 = p.area + sizeof(wc::PeerCtl);
    wc::peer_wait_reduce_kernel<<<1, 64, 0, e->stream>>>(area0, p.stride, p.n_ranks, p.root_rank, p.seq, e->d_status,
                                                        e->cfg.base_offset, e->consumed, p.d_job);
    wc::peer_clear_root_kernel<<<e->num_sms, 256, 0, e->stream>>>(p.d_job, (u64 *)e->table_block, e->table_block_bytes / 8, e->d_status);
    wc::import_areas_dev_kernel<<<e->num_sms * 4, 256, 0, e->stream>>>(e->T, p.d_job, area0, p.stride, p.seq, p.cap_words, p.cap_long);
    wc::peer_done_kernel<<<1, 1, 0, e->stream>>>((wc::PeerCtl *)p.area, p.seq);
    CK(cudaGetLastError());
    e->launches += 4;
    return WC_OK;
}

uint32_t wc_key_owner(uint64_t k0, uint64_t k1, uint32_t n_parts) { return wc::key_owner(k0, k1, n_parts); }

} // extern "C"

// ------------------------------------------------------------------------------------------------
// wc_count_device_static: the count kernel for callers that must not allocate.
//
// The regular engine keeps the kernel's working state -- the Status block and one WarpSummary per
// warp -- in memory it cudaMallocs in wc_create().  Some users of this library cannot do that: a
// host application that has handed the whole device heap to its own pool allocator (a cudaMalloc
// there fails or steals from the pool), or code on a latency-critical path, where cudaMalloc and
// cudaFree stall the whole device while they run.  For them the same kernel is offered with its
// working state in a fixed array that is part of the library image; the caller supplies the text
// and every buffer the results are written to.
//
// g_static_scratch lives in every GPU's memory from the moment the library is loaded there until
// the process exits (kStaticScratchBytes, about 106 KiB).  In host code the name `g_static_scratch`
// is only a handle for the symbol, not a device address, so the address the kernel needs is looked
// up once per device with cudaGetSymbolAddress() while that device is current.
namespace {

constexpr u32 kStaticMaxSMs = 160;    // more than any sm_100 part has (B200: 148)
constexpr u32 kStaticMaxWarps = kStaticMaxSMs * kMaxTokenWarps; // one WarpSummary per tokenizing warp
constexpr size_t kStaticStatusBytes = 512;
constexpr size_t kStaticScratchBytes = kStaticStatusBytes + sizeof(wc::WarpSummary) * kStaticMaxWarps;
static_assert(sizeof(wc::Status) <= kStaticStatusBytes, "Status must fit its slot");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kStaticMaxDevices = 64;
struct StaticDevice {
    bool ready;
    unsigned char *scratch; // device address of g_static_scratch on this device
    int num_sms;
};
StaticDevice g_static_dev[kStaticMaxDevices];
std::mutex g_static_mu; // guards the first-call set-up of g_static_dev[] (host threads may race to it)
thread_local char g_static_err[192];

// Start-of-call state: everything zero except the "none yet" markers; arena unit 0 is never used.
__global__ void static_begin_kernel(wc::Status *st) {
    if (threadIdx.x == 0) {
        wc::Status s;
        memset(&s, 0, sizeof(s));
        s.stop = wc::kNone; s.p0 = wc::kNone; s.arena_used = 1;
        *st = s;
    }
}

__global__ void static_end_kernel(const wc::Status *st, u64 *out) {
    if (threadIdx.x == 0) {
        out[0] = st->n_flush; out[1] = st->n_empty; out[2] = st->stop; out[3] = st->p0;
        out[4] = st->lost_empty; out[5] = st->p0_found; out[6] = st->err; out[7] = st->w_entries;
        out[8] = st->x_entries; out[9] = st->arena_used; out[10] = st->n_long; out[11] = st->n_spill;
    }
}

int static_fail(int code, const char *what, cudaError_t ce) {
    if (ce != cudaSuccess) snprintf(g_static_err, sizeof(g_static_err), "%s: %s", what, cudaGetErrorString(ce));
    else snprintf(g_static_err, sizeof(g_static_err), "%s", what);
    return code;
}

} // namespace

extern "C" const char *wc_static_last_error(void) { return g_static_err; }

extern "C" int wc_count_device_static(int device, const void *dev_text, size_t n, uint64_t *dev_short_table,
                                      uint32_t short_log2, uint64_t *dev_long_table, uint32_t long_log2,
                                      uint64_t *dev_arena, uint64_t arena_bytes, uint64_t *dev_counters,
                                      void *stream_handle) {
    g_static_err[0] = 0;
    if (device < 0 || device >= kStaticMaxDevices || (!dev_text && n) || !dev_short_table || !dev_long_table ||
        !dev_arena || !dev_counters || short_log2 < 10 || short_log2 > 32 || long_log2 < 10 || long_log2 > 34 ||
        arena_bytes < 4096)
        return static_fail(WC_ERR_INVALID, "wc_count_device_static: bad argument", cudaSuccess);
    if (((uintptr_t)dev_text & 15) != 0 || ((uintptr_t)dev_short_table & 15) != 0)
        return static_fail(WC_ERR_UNALIGNED, "wc_count_device_static: text and short table must be 16-byte aligned", cudaSuccess);
    if (n > (size_t)WC_STATIC_MAX_BYTES)
        return static_fail(WC_ERR_TOO_LARGE, "wc_count_device_static: text larger than WC_STATIC_MAX_BYTES", cudaSuccess);

    int prev_device = 0;
    cudaError_t ce = cudaGetDevice(&prev_device);
    if (ce != cudaSuccess) return static_fail(WC_ERR_CUDA, "cudaGetDevice", ce);
    if ((ce = cudaSetDevice(device)) != cudaSuccess) return static_fail(WC_ERR_CUDA, "cudaSetDevice", ce);
    int rc = WC_OK;
    do {
        std::lock_guard<std::mutex> guard(g_static_mu);
        StaticDevice &sd = g_static_dev[device];
        if (!sd.ready) { // first call on this device: find the array's address here, raise the kernel's smem limit
            void *addr = nullptr;
            if ((ce = cudaGetSymbolAddress(&addr, g_static_scratch)) != cudaSuccess) {
                rc = static_fail(WC_ERR_CUDA, "cudaGetSymbolAddress(g_static_scratch)", ce); break;
            }
            int sms = 0;
            if ((ce = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) {
                rc = static_fail(WC_ERR_CUDA, "cudaDeviceGetAttribute", ce); break;
            }
            if ((ce = cudaFuncSetAttribute(wc::count_kernel<wc::MapIdentity>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)wc::kSmemTotal)) != cudaSuccess ||
                (ce = cudaFuncSetAttribute(wc::count_kernel2<wc::MapIdentity>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)wc::k2SmemTotal)) != cudaSuccess) {
                rc = static_fail(WC_ERR_CUDA, "cudaFuncSetAttribute", ce); break;
            }
            sd.scratch = (unsigned char *)addr; sd.num_sms = sms; sd.ready = true;
        }
        const bool k2 = forward_uses_kernel2(0, false, false);
        const u32 n_ws = (u32)sd.num_sms * (k2 ? wc::k2Warps : wc::kProducers);
        if ((u32)sd.num_sms > kStaticMaxSMs) {
            rc = static_fail(WC_ERR_TOO_LARGE, "wc_count_device_static: this GPU has more SMs than the static scratch covers", cudaSuccess);
            break;
        }
        cudaStream_t stream = (cudaStream_t)stream_handle;
        wc::Status *st = (wc::Status *)sd.scratch;
        wc::WarpSummary *ws = (wc::WarpSummary *)(sd.scratch + kStaticStatusBytes);
        const size_t s_n = (size_t)1 << short_log2, l_n = (size_t)1 << long_log2;
        wc::Tables T{}; // (no overflow log here: ovf_log = nullptr, ovf_cap = 0 -> a full table raises err, counters[6])
        T.w_keys = (ulonglong2 *)dev_short_table; T.w_counts = (u64 *)dev_short_table + 2 * s_n;
        T.w_finv = (u64 *)dev_short_table + 3 * s_n; T.w_log2 = short_log2;
        T.x_refs = (u64 *)dev_long_table; T.x_counts = T.x_refs + l_n; T.x_finv = T.x_refs + 2 * l_n; T.x_log2 = long_log2;
        T.arena = (u64 *)dev_arena; T.arena_units = arena_bytes / 8; T.st = st;

        // results start from zero; the shared working state is re-initialised on every call
        if ((ce = cudaMemsetAsync(dev_short_table, 0, 4 * s_n * sizeof(u64), stream)) != cudaSuccess ||
            (ce = cudaMemsetAsync(dev_long_table, 0, 3 * l_n * sizeof(u64), stream)) != cudaSuccess) {
            rc = static_fail(WC_ERR_CUDA, "cudaMemsetAsync", ce); break;
        }
        static_begin_kernel<<<1, 32, 0, stream>>>(st);
        if (n > 0) {
            wc::Launch L{}; // (em.cursor = nullptr: the counting path, not the sort path's record emitter)
            L.buf = (const uint8_t *)dev_text; L.n = n; L.base = 0;
            L.last_piece = 1; L.negate = 0; L.strict = 0; L.ws = ws; L.n_ws = n_ws;
            bool ok = launch_count(stream, sd.num_sms, WC_MAP_IDENTITY, L, T, k2);
            L.negate = 1;
            ok = ok && launch_count(stream, sd.num_sms, WC_MAP_IDENTITY, L, T, false);
            if (!ok) { // a CTA's range must fit 32-bit offsets
                rc = static_fail(WC_ERR_TOO_LARGE, "wc_count_device_static: text too large for this GPU's SM count", cudaSuccess);
                break;
            }
        }
        static_end_kernel<<<1, 32, 0, stream>>>(st, (u64 *)dev_counters);
        if ((ce = cudaGetLastError()) != cudaSuccess) { rc = static_fail(WC_ERR_CUDA, "kernel launch", ce); break; }
    } while (0);
    cudaSetDevice(prev_device);
    return rc;
}
