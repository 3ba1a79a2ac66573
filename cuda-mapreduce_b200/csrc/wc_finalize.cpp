// wc_finalize.cpp -- see wc_finalize.h.
#include "wc_finalize.h"

#include <algorithm>
#include <atomic>
#include <memory>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <thread>
#include <chrono>
#include <cstdlib>

wc_result::~wc_result() { free(big_rows); free(big_blob); }

namespace wc {

void parallel_copy(void *dst, const void *src, size_t bytes) {
    const size_t kMin = (size_t)4 << 20;
    unsigned hw = std::thread::hardware_concurrency();
    size_t nt = bytes / kMin;
    if (nt > 8) nt = 8;
    if (hw && nt > hw) nt = hw;
    if (nt < 2) { memcpy(dst, src, bytes); return; }
    std::vector<std::thread> th;
    const size_t step = ((bytes / nt) + 4095) & ~(size_t)4095;
    for (size_t t = 0; t < nt; t++) {
        const size_t lo = t * step, hi = (t + 1 == nt || (t + 1) * step > bytes) ? bytes : (t + 1) * step;
        if (lo >= hi) break;
        th.emplace_back([=] { memcpy((char *)dst + lo, (const char *)src + lo, hi - lo); });
    }
    for (auto &x : th) x.join();
}

// ---- the large host reduce (the sort path's wc_finish at 10^8 distinct words): every pass on all cores ----------

// How many threads a host pass over n items gets.
static unsigned host_threads(size_t n) {
    if (n < ((size_t)1 << 18)) return 1;
    unsigned hw = std::thread::hardware_concurrency();
    if (hw == 0) hw = 1;
    return hw > 32 ? 32 : hw;
}

// f(t) for t in [0, nt) on nt threads (the caller's included).
template <class F>
static void run_threads(unsigned nt, F f) {
    if (nt <= 1) { f(0u); return; }
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; t++) th.emplace_back([&f, t] { f(t); });
    f(0u);
    for (auto &x : th) x.join();
}

// Sorts a[0, n) by `less`, where bucket(x) in [0, nb) is monotone in that order: per-thread histograms over
// contiguous slices, a stable scatter into a second array, then the buckets sorted on all threads (a
// most-significant-digit radix pass followed by comparison sorts).  Arrays are left uninitialised on purpose:
// zero-filling 10^8 elements on one core costs more than the pass that overwrites them.
template <class T, class Bucket, class Less>
static void bucket_sort(std::unique_ptr<T[]> &a, size_t n, uint32_t nb, Bucket bucket, Less less) {
    const unsigned nt = host_threads(n);
    if (nt <= 1) { std::sort(a.get(), a.get() + n, less); return; }
    std::vector<std::vector<size_t>> cnt(nt, std::vector<size_t>(nb, 0));
    const size_t step = (n + nt - 1) / nt;
    T *src = a.get();
    run_threads(nt, [&](unsigned t) {
        const size_t lo = t * step, hi = lo + step < n ? lo + step : n;
        for (size_t i = lo; i < hi; i++) cnt[t][bucket(src[i])]++;
    });
    std::vector<size_t> start(nb + 1, 0);
    for (uint32_t b = 0; b < nb; b++) {
        size_t at = start[b];
        for (unsigned t = 0; t < nt; t++) { const size_t c = cnt[t][b]; cnt[t][b] = at; at += c; }
        start[b + 1] = at;
    }
    std::unique_ptr<T[]> out(new T[n]);
    T *dst = out.get();
    run_threads(nt, [&](unsigned t) {
        const size_t lo = t * step, hi = lo + step < n ? lo + step : n;
        for (size_t i = lo; i < hi; i++) dst[cnt[t][bucket(src[i])]++] = src[i];
    });
    std::atomic<uint32_t> next{0};
    run_threads(nt, [&](unsigned) {
        for (;;) {
            const uint32_t b = next.fetch_add(1);
            if (b >= nb) return;
            if (start[b + 1] - start[b] > 1) std::sort(dst + start[b], dst + start[b + 1], less);
        }
    });
    a.swap(out);
}

// Rows in order of first appearance.
static void sort_rows_by_first(std::vector<Row> &rows) {
    const size_t n = rows.size();
    if (host_threads(n) <= 1) { // LSD radix sort on the 64-bit offset (11-bit digits, only the digits some offset uses)
        std::vector<Row> tmp(n);
        uint64_t all_or = 0;
        for (const Row &r : rows) all_or |= r.first;
        for (int shift = 0; shift < 64 && (all_or >> shift) != 0; shift += 11) {
            size_t cnt[2049] = {0};
            for (const Row &r : rows) cnt[((r.first >> shift) & 2047) + 1]++;
            for (int d = 0; d < 2048; d++) cnt[d + 1] += cnt[d];
            for (const Row &r : rows) tmp[cnt[(r.first >> shift) & 2047]++] = r;
            rows.swap(tmp);
        }
        return;
    }
    uint64_t mx = 0;
    for (const Row &r : rows) mx = r.first > mx ? r.first : mx;
    int shift = 0;
    while ((mx >> shift) >= 65536) shift++;
    std::unique_ptr<Row[]> a(new Row[n]);
    parallel_copy(a.get(), rows.data(), n * sizeof(Row));
    bucket_sort(a, n, 65536u, [shift](const Row &r) { return (uint32_t)(r.first >> shift); },
                [](const Row &x, const Row &y) { return x.first < y.first; });
    parallel_copy(rows.data(), a.get(), n * sizeof(Row));
}

// One distinct word in bytewise order: its first 16 bytes as two big-endian integers (zero padded: a word
// holds no NUL byte, so padding sorts a word before its extensions) and what the sweep needs of its entry, so
// that neither the sort nor the sweep goes back to the entry array or -- for words of up to 16 bytes -- the text.
struct Ordered { uint64_t hi, lo, first, blob_off; int64_t count; uint32_t len, pad; };

// main.cu's reduce (see finalize_rows) over entries in no particular order: rows, unsorted, appended to `rows`.
static void reduce_unordered(const char *blob, const std::vector<HostEntry> &v, std::vector<Row> &rows, double *t_keys_ms, double *t_sort_ms) {
    const auto t0 = std::chrono::steady_clock::now();
    const size_t nd = v.size();
    const unsigned nt = host_threads(nd);
    const size_t step = (nd + nt - 1) / nt;
    std::vector<size_t> live(nt + 1, 0);
    run_threads(nt, [&](unsigned t) {
        const size_t lo = t * step, hi = lo + step < nd ? lo + step : nd;
        size_t c = 0;
        for (size_t i = lo; i < hi; i++) c += v[i].count != 0;
        live[t + 1] = c;
    });
    for (unsigned t = 0; t < nt; t++) live[t + 1] += live[t];
    const size_t no = live[nt];
    std::unique_ptr<Ordered[]> a(new Ordered[no ? no : 1]);
    run_threads(nt, [&](unsigned t) {
        const size_t lo = t * step, hi = lo + step < nd ? lo + step : nd;
        size_t at = live[t];
        for (size_t i = lo; i < hi; i++) {
            if (v[i].count == 0) continue;
            unsigned char pad[16] = {0};
            memcpy(pad, blob + v[i].blob_off, v[i].len < 16 ? v[i].len : 16);
            uint64_t x, y;
            memcpy(&x, pad, 8); memcpy(&y, pad + 8, 8);
            a[at++] = Ordered{__builtin_bswap64(x), __builtin_bswap64(y), v[i].first, v[i].blob_off, v[i].count, v[i].len, 0};
        }
    });
    const auto t1 = std::chrono::steady_clock::now();
    bucket_sort(a, no, 65536u, [](const Ordered &k) { return (uint32_t)(k.hi >> 48); },
                [blob](const Ordered &x, const Ordered &y) {
                    if (x.hi != y.hi) return x.hi < y.hi;
                    if (x.lo != y.lo) return x.lo < y.lo;
                    if (x.len <= 16 || y.len <= 16) return x.len < y.len; // (equal padded heads: the shorter is the prefix)
                    const uint32_t l = (x.len < y.len ? x.len : y.len) - 16;
                    const int c = memcmp(blob + x.blob_off + 16, blob + y.blob_off + 16, l);
                    if (c) return c < 0;
                    return x.len < y.len;
                });
    const auto t2 = std::chrono::steady_clock::now();
    *t_keys_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    *t_sort_ms = std::chrono::duration<double, std::milli>(t2 - t1).count();
    // is word p a prefix of word q?  (p.len <= q.len checked by the caller)
    auto is_prefix = [blob, &a](uint32_t p, uint32_t q) {
        const Ordered &t = a[p], &w = a[q];
        if (t.len <= 8) return ((t.hi ^ w.hi) >> (64 - 8 * t.len)) == 0;
        if (t.hi != w.hi) return false;
        if (t.len < 16) return ((t.lo ^ w.lo) >> (128 - 8 * t.len)) == 0;
        if (t.lo != w.lo) return false;
        return t.len == 16 || memcmp(blob + t.blob_off + 16, blob + w.blob_off + 16, t.len - 16) == 0;
    };
    std::unique_ptr<uint32_t[]> owner(new uint32_t[no ? no : 1]);
    std::vector<uint32_t> stk, best; // positions in a[]
    stk.reserve(64); best.reserve(64);
    for (size_t i = 0; i <= no; i++) {
        while (!stk.empty()) {
            const bool open = i < no && a[stk.back()].len <= a[i].len && is_prefix(stk.back(), (uint32_t)i);
            if (open) break;
            owner[stk.back()] = best.back();
            const uint32_t b = best.back();
            stk.pop_back(); best.pop_back();
            if (!best.empty() && a[b].first < a[best.back()].first) best.back() = b;
        }
        if (i < no) { stk.push_back((uint32_t)i); best.push_back((uint32_t)i); }
    }
    // counts to the owners, then the rows (the words that own themselves)
    std::unique_ptr<int64_t[]> acc(new int64_t[no ? no : 1]);
    const size_t ostep = (no + nt - 1) / nt;
    run_threads(nt, [&](unsigned t) {
        const size_t lo = t * ostep, hi = lo + ostep < no ? lo + ostep : no;
        if (lo < hi) memset(acc.get() + lo, 0, (hi - lo) * sizeof(int64_t));
    });
    size_t n_rows = 0;
    for (size_t i = 0; i < no; i++) { acc[owner[i]] += a[i].count; n_rows += owner[i] == i; }
    rows.reserve(rows.size() + n_rows);
    for (size_t i = 0; i < no; i++)
        if (owner[i] == i) rows.push_back(Row{a[i].blob_off, a[i].len, acc[i], a[i].first});
}

// The reference's reducer (main.cu:69-108) walks the words in input order and credits each one to
// the FIRST output row of which it is a prefix (compare(), main.cu:57-67), appending a row when
// there is none.  Order-free restatement: with exact[w] and first[w] per distinct
// word, owner(w) is the distinct word with the smallest `first` among those having prefix w; rows
// are the words that own themselves, in order of `first`; a row's count is the sum of exact[] over
// what it owns.  Empty words after the first non-empty one go to row 0 (main.cu:81-86 with the
// all-zero slot of main.cu:154).
void finalize_rows(std::string &&blob_in, std::vector<HostEntry> &v, int mode, uint64_t n_empty,
                   uint64_t n_lost_empty, wc_result *out, const uint32_t *by_key, const uint32_t *by_first) {
    static const bool timing = getenv("WC_TIMING") != nullptr; // stderr: where the host reduce spends its time
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
    };
    const auto t_start = now();
    auto t_keys = t_start, t_sort = t_start, t_sweep = t_start;
    out->blob = std::move(blob_in);
    out->rows.clear();
    out->total = 0;
    const char *blob = out->blob.data();
    const size_t nd = v.size();
    if (nd == 0) return;

    const bool pre_ordered = by_key != nullptr && by_first != nullptr;
    std::vector<uint8_t> is_row;        // pre-ordered input: which entries become rows ...
    std::vector<int64_t> row_count;     // ... and with what count
    if (mode == WC_MODE_EXACT) {
        if (pre_ordered) {
            is_row.assign(nd, 0); row_count.assign(nd, 0);
            for (size_t i = 0; i < nd; i++) if (v[i].count != 0) { is_row[i] = 1; row_count[i] = v[i].count; }
        } else {
            out->rows.reserve(nd);
            for (const HostEntry &e : v) out->rows.push_back(Row{e.blob_off, e.len, e.count, e.first});
        }
    } else if (!pre_ordered) {
        double k_ms = 0, s_ms = 0;
        reduce_unordered(blob, v, out->rows, &k_ms, &s_ms);
        t_keys = t_start + std::chrono::duration_cast<std::chrono::steady_clock::duration>(std::chrono::duration<double, std::milli>(k_ms));
        t_sort = t_keys + std::chrono::duration_cast<std::chrono::steady_clock::duration>(std::chrono::duration<double, std::milli>(s_ms));
    } else {
        // The device's two orders came along (wc_finish, WC_HOST_REDUCE=1).  `order`: entry indices in bytewise
        // order of the words, so that every word is immediately followed by the words it is a prefix of.
        std::vector<uint32_t> order;
        order.reserve(nd);
        auto tail_less = [blob, &v](uint32_t x, uint32_t y) { // words that agree on their first 16 bytes
            const HostEntry &a = v[x], &b = v[y];
            if (a.len <= 16 || b.len <= 16) return a.len < b.len;
            const uint32_t l = (a.len < b.len ? a.len : b.len) - 16;
            const int c = memcmp(blob + a.blob_off + 16, blob + b.blob_off + 16, l);
            if (c) return c < 0;
            return a.len < b.len;
        };
        for (size_t i = 0; i < nd; i++) if (by_key[i] < nd && v[by_key[i]].count != 0) order.push_back(by_key[i]);
        // the device order looked at 16 bytes: put runs of longer words with equal heads right
        for (size_t i = 0; i < order.size();) {
            size_t j = i + 1;
            if (v[order[i]].len >= 16)
                while (j < order.size() && v[order[j]].len >= 16 &&
                       memcmp(blob + v[order[i]].blob_off, blob + v[order[j]].blob_off, 16) == 0) j++;
            if (j - i > 1) std::sort(order.begin() + i, order.begin() + j, tail_less);
            i = j;
        }
        t_sort = now();
        const size_t no = order.size();
        // is word t a prefix of word w?  (t.len <= w.len checked by the caller)
        auto is_prefix = [blob](const HostEntry &t, const HostEntry &w) {
            return memcmp(blob + t.blob_off, blob + w.blob_off, t.len) == 0;
        };
        std::vector<uint32_t> owner(no), stk, best; // all three hold positions in `order`
        stk.reserve(64); best.reserve(64);
        for (size_t i = 0; i <= no; i++) {
            while (!stk.empty()) {
                const HostEntry &t = v[order[stk.back()]];
                const bool open = i < no && t.len <= v[order[i]].len && is_prefix(t, v[order[i]]);
                if (open) break;
                owner[stk.back()] = best.back();
                const uint32_t b = best.back();
                stk.pop_back(); best.pop_back();
                if (!best.empty() && v[order[b]].first < v[order[best.back()]].first) best.back() = b;
            }
            if (i < no) { stk.push_back((uint32_t)i); best.push_back((uint32_t)i); }
        }
        std::vector<int64_t> acc(no, 0);
        for (size_t i = 0; i < no; i++) acc[owner[i]] += v[order[i]].count;
        is_row.assign(nd, 0); row_count.assign(nd, 0);
        for (size_t i = 0; i < no; i++) if (owner[i] == i) { is_row[order[i]] = 1; row_count[order[i]] = acc[i]; }
    }
    t_sweep = now();
    if (pre_ordered) { // rows in order of first appearance: walk the device's order
        size_t n_rows = 0;
        for (uint8_t b : is_row) n_rows += b;
        out->rows.reserve(n_rows);
        for (size_t i = 0; i < nd; i++) {
            const uint32_t x = by_first[i];
            if (x < nd && is_row[x]) out->rows.push_back(Row{v[x].blob_off, v[x].len, row_count[x], v[x].first});
        }
    } else {
        sort_rows_by_first(out->rows);
    }
    if (mode != WC_MODE_EXACT && !out->rows.empty()) out->rows[0].count += (int64_t)(n_empty - n_lost_empty);
    for (const Row &r : out->rows) out->total += r.count;
    if (timing)
        fprintf(stderr, "[wc] finalize_rows: %zu words -> %zu rows; keys %.2f sort %.2f sweep %.2f order %.2f ms\n", nd,
                out->rows.size(), ms(t_start, t_keys), ms(t_keys, t_sort), ms(t_sort, t_sweep), ms(t_sweep, now()));
}

} // namespace wc

// ------------------------------------------------------------------ accessors (C ABI)

extern "C" {

size_t wc_result_rows(const wc_result *r) { return r ? r->row_count() : 0; }

int wc_result_row(const wc_result *r, size_t i, const char **word, size_t *len, int64_t *count) {
    if (!r || i >= r->row_count()) return WC_ERR_INVALID;
    const wc::Row &row = r->row_data()[i];
    if (word) *word = r->blob_data() + row.blob_off;
    if (len) *len = row.len;
    if (count) *count = row.count;
    return WC_OK;
}

uint64_t wc_result_row_first(const wc_result *r, size_t i) {
    return (!r || i >= r->row_count()) ? ~0ull : r->row_data()[i].first;
}

int64_t wc_result_total(const wc_result *r) { return r ? r->total : 0; }

int wc_result_stats(const wc_result *r, wc_stats *out) {
    if (!r || !out) return WC_ERR_INVALID;
    *out = r->stats;
    return WC_OK;
}

static inline uint64_t mix_u64(uint64_t x) {
    x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32; x *= 0xD6E8FEB86659FD93ull; x ^= x >> 32;
    return x;
}

int wc_result_checksum(const wc_result *r, uint64_t out[2]) {
    if (!r || !out) return WC_ERR_INVALID;
    const size_t n = r->row_count();
    const wc::Row *rows = r->row_data();
    const char *blob = r->blob_data();
    const unsigned nt = n > (1u << 20) ? std::max(1u, std::min(16u, std::thread::hardware_concurrency())) : 1u;
    std::vector<uint64_t> a(nt, 0), b(nt, 0);
    auto work = [&](unsigned t) {
        uint64_t sa = 0, sb = 0;
        for (size_t i = n * t / nt; i < n * (t + 1) / nt; i++) {
            uint64_t h = 0xCBF29CE484222325ull;
            const unsigned char *w = (const unsigned char *)blob + rows[i].blob_off;
            for (uint32_t k = 0; k < rows[i].len; k++) { h ^= w[k]; h *= 0x100000001B3ull; }
            sa += mix_u64(h) * (uint64_t)rows[i].count;
            sb += mix_u64(h + (uint64_t)i) * (rows[i].first + 1);
        }
        a[t] = sa; b[t] = sb;
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    out[0] = out[1] = 0;
    for (unsigned t = 0; t < nt; t++) { out[0] += a[t]; out[1] += b[t]; }
    return WC_OK;
}

void wc_result_free(wc_result *r) { delete r; }

// main.cu:210-218
size_t wc_result_render(const wc_result *r, char *buf, size_t cap) {
    static const char sep[] = "--------------------------\n";
    size_t len = 0;
    auto put = [&](const char *p, size_t n) {
        if (buf && len + n <= cap) memcpy(buf + len, p, n);
        len += n;
    };
    char num[48];
    put(sep, sizeof(sep) - 1);
    if (r)
        for (size_t i = 0; i < r->row_count(); i++) {
            const wc::Row &row = r->row_data()[i];
            put(r->blob_data() + row.blob_off, row.len);
            int k = snprintf(num, sizeof(num), "\t%lld\n", (long long)row.count);
            put(num, (size_t)k);
        }
    put(sep, sizeof(sep) - 1);
    int k = snprintf(num, sizeof(num), "Total Count:%lld\n", (long long)(r ? r->total : 0));
    put(num, (size_t)k);
    return len;
}

// Test hook: run the host reduce on a caller-supplied table of exact counts.  `words` is the
// concatenation of the distinct words, lens/counts/firsts are per word.
int wc_debug_finalize(const char *words, const uint32_t *lens, const int64_t *counts, const uint64_t *firsts,
                      size_t n, int mode, uint64_t n_empty, uint64_t n_lost_empty, wc_result **out) {
    if (!out) return WC_ERR_INVALID;
    std::vector<wc::HostEntry> v(n);
    uint64_t off = 0;
    for (size_t i = 0; i < n; i++) {
        v[i] = wc::HostEntry{off, lens[i], counts[i], firsts[i]};
        off += lens[i];
    }
    wc_result *r = new wc_result();
    wc::finalize_rows(std::string(words, off), v, mode, n_empty, n_lost_empty, r);
    *out = r;
    return WC_OK;
}

const char *wc_strerror(int status) {
    switch (status) {
    case WC_OK: return "ok";
    case WC_ERR_INVALID: return "invalid argument or call order";
    case WC_ERR_CUDA: return "CUDA runtime error";
    case WC_ERR_NOMEM: return "out of memory";
    case WC_ERR_TABLE_FULL: return "device table or long-word arena full";
    case WC_ERR_UNALIGNED: return "device buffer is not 16-byte aligned";
    case WC_ERR_NCCL: return "NCCL error";
    case WC_ERR_TOO_LARGE: return "size beyond what this entry point supports";
    default: return "unknown status";
    }
}

} // extern "C"
