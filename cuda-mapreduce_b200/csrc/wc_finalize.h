// wc_finalize.h -- host side of the reduce: turns the device tables' exact counts into the rows
// the reference prints.  Pure C++ (no CUDA), so the CPU tests can drive it.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/wc.h"

namespace wc {

struct HostEntry {      // one distinct non-empty word with its exact count
    uint64_t blob_off;  // where its bytes start in the blob
    uint32_t len;
    int64_t count;
    uint64_t first;     // offset of the delimiter that flushed its first occurrence
};

struct Row {
    uint64_t blob_off;
    uint32_t len;
    int64_t count;
    uint64_t first;
};

} // namespace wc

struct wc_result {
    std::string blob; // bytes of every row's word
    std::vector<wc::Row> rows;
    // Large results (wc_finish with 10^8 rows) are copied into plain malloc'd arrays by several
    // threads instead: a vector would first zero-fill every page on one core.
    wc::Row *big_rows = nullptr; size_t n_big_rows = 0; char *big_blob = nullptr;
    int64_t total = 0;
    wc_stats stats{};
    const wc::Row *row_data() const { return big_rows ? big_rows : rows.data(); }
    size_t row_count() const { return big_rows ? n_big_rows : rows.size(); }
    const char *blob_data() const { return big_blob ? big_blob : blob.data(); }
    ~wc_result();
    wc_result() = default;
    wc_result(const wc_result &) = delete;
    wc_result &operator=(const wc_result &) = delete;
};

namespace wc {

// Applies the reduce semantics of /root/reference/main.cu:57-108 in its order-free form
// (SURVEY.md A.3).  `entries` index into `blob`; the result takes ownership of `blob`.
// Optional, both or neither (wc_finish computes them on the GPU): `by_key` lists entry indices
// ordered by the words' first 16 bytes (bytewise, shorter first; ties between longer words in any
// order), `by_first` ordered by `first`; entries with count 0 are placeholders and are skipped.
// dst <- src with several threads when it is worth it (large, freshly allocated destinations)
void parallel_copy(void *dst, const void *src, size_t bytes);

void finalize_rows(std::string &&blob, std::vector<HostEntry> &entries, int mode, uint64_t n_empty,
                   uint64_t n_lost_empty, wc_result *out, const uint32_t *by_key = nullptr,
                   const uint32_t *by_first = nullptr);

} // namespace wc
