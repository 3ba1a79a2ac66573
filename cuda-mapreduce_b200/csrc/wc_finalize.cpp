// wc_finalize.cpp -- see wc_finalize.h.
#include "wc_finalize.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <numeric>

namespace wc {

// The reference's reducer (main.cu:69-108) walks the words in input order and credits each one to
// the FIRST output row of which it is a prefix (compare(), main.cu:57-67), appending a row when
// there is none.  Order-free restatement: with exact[w] and first[w] per distinct
// word, owner(w) is the distinct word with the smallest `first` among those having prefix w; rows
// are the words that own themselves, in order of `first`; a row's count is the sum of exact[] over
// what it owns.  Empty words after the first non-empty one go to row 0 (main.cu:81-86 with the
// all-zero slot of main.cu:154).
void finalize_rows(std::string &&blob_in, std::vector<HostEntry> &v, int mode, uint64_t n_empty,
                   uint64_t n_lost_empty, wc_result *out) {
    out->blob = std::move(blob_in);
    out->rows.clear();
    out->total = 0;
    const char *blob = out->blob.data();
    const size_t nd = v.size();
    if (nd == 0) return;

    if (mode == WC_MODE_EXACT) {
        out->rows.reserve(nd);
        for (const HostEntry &e : v) out->rows.push_back(Row{e.blob_off, e.len, e.count, e.first});
    } else {
        // bytewise order: every word is immediately followed by the words it is a prefix of
        std::sort(v.begin(), v.end(), [blob](const HostEntry &a, const HostEntry &b) {
            const uint32_t l = a.len < b.len ? a.len : b.len;
            const int c = memcmp(blob + a.blob_off, blob + b.blob_off, l);
            if (c) return c < 0;
            return a.len < b.len;
        });
        std::vector<uint32_t> owner(nd), stk, best;
        stk.reserve(64); best.reserve(64);
        for (size_t i = 0; i <= nd; i++) {
            while (!stk.empty()) {
                const HostEntry &t = v[stk.back()];
                const bool open = i < nd && t.len <= v[i].len &&
                                  memcmp(blob + t.blob_off, blob + v[i].blob_off, t.len) == 0;
                if (open) break;
                owner[stk.back()] = best.back();
                const uint32_t b = best.back();
                stk.pop_back(); best.pop_back();
                if (!best.empty() && v[b].first < v[best.back()].first) best.back() = b;
            }
            if (i < nd) { stk.push_back((uint32_t)i); best.push_back((uint32_t)i); }
        }
        std::vector<int64_t> acc(nd, 0);
        for (size_t i = 0; i < nd; i++) acc[owner[i]] += v[i].count;
        for (size_t i = 0; i < nd; i++)
            if (owner[i] == i) out->rows.push_back(Row{v[i].blob_off, v[i].len, acc[i], v[i].first});
    }
    std::sort(out->rows.begin(), out->rows.end(), [](const Row &a, const Row &b) { return a.first < b.first; });
    if (mode != WC_MODE_EXACT && !out->rows.empty()) out->rows[0].count += (int64_t)(n_empty - n_lost_empty);
    for (const Row &r : out->rows) out->total += r.count;
}

} // namespace wc

// ------------------------------------------------------------------ accessors (C ABI)

extern "C" {

size_t wc_result_rows(const wc_result *r) { return r ? r->rows.size() : 0; }

int wc_result_row(const wc_result *r, size_t i, const char **word, size_t *len, int64_t *count) {
    if (!r || i >= r->rows.size()) return WC_ERR_INVALID;
    if (word) *word = r->blob.data() + r->rows[i].blob_off;
    if (len) *len = r->rows[i].len;
    if (count) *count = r->rows[i].count;
    return WC_OK;
}

uint64_t wc_result_row_first(const wc_result *r, size_t i) {
    return (!r || i >= r->rows.size()) ? ~0ull : r->rows[i].first;
}

int64_t wc_result_total(const wc_result *r) { return r ? r->total : 0; }

int wc_result_stats(const wc_result *r, wc_stats *out) {
    if (!r || !out) return WC_ERR_INVALID;
    *out = r->stats;
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
        for (const wc::Row &row : r->rows) {
            put(r->blob.data() + row.blob_off, row.len);
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
