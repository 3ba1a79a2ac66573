// wc_compat.cpp -- runMapReduce() with the reference's exact signature and struct layouts
// (/root/reference/main.cu:133, structs main.cu:16-33), implemented on the engine.
//
// The reference's main() pre-splits each line into words on the host (main.cu:187-201) and hands
// runMapReduce() the words.  Here the words are serialised back into a byte stream -- every word
// followed by one 0x20, no line ends -- so that the engine's splitter emits exactly the same word
// sequence (empty words included: an empty word becomes two adjacent delimiters), then counted on
// the GPU with the reference's reduce semantics.
#include <cstddef>
#include <cstring>
#include <string>

#include "../../include/wc.h"
#include "../../include/wc_compat.h"

static_assert(sizeof(wc_ref_word) == 30, "Word is 30 bytes in main.cu:16-18");
static_assert(sizeof(wc_ref_keypair) == 36 && offsetof(wc_ref_keypair, count) == 32, "KeyPair layout, main.cu:19-22");
static_assert(sizeof(wc_ref_input) == 604 && offsetof(wc_ref_input, n_words) == 600, "InputData layout, main.cu:24-27");

extern "C" void runMapReduce(wc_ref_input *input, int nInputCount, wc_ref_output *output, int *nOutputCount) {
    if (nOutputCount) *nOutputCount = 0;
    if (!input || !output || !nOutputCount || nInputCount < 0) return;
    std::string text;
    for (int i = 0; i < nInputCount; i++) {
        const int nw = input[i].n_words; // main.cu:39-40
        for (int w = 0; w < nw && w < WC_REF_WORDS_PER_LINE; w++) {
            const char *s = input[i].words[w].text;
            text.append(s, strnlen(s, WC_REF_WORD_BYTES)); // bytes up to the terminator, main.cu:46-51
            text.push_back(' ');
        }
    }
    // a 1-byte stream would read as "line shorter than 2" (main.cu:185); it holds one empty word,
    // which the reducer drops anyway (main.cu:81,103) -> no rows either way
    wc_config cfg;
    wc_config_default(&cfg);
    cfg.short_table_log2 = 12; cfg.long_table_log2 = 12; cfg.arena_bytes = 1 << 16; cfg.chunk_bytes = 1 << 16;
    wc_engine *e = nullptr;
    if (wc_create(&e, &cfg) != WC_OK) return;
    wc_result *r = nullptr;
    if (wc_feed(e, text.data(), text.size()) == WC_OK && wc_finish(e, &r) == WC_OK) {
        size_t n = wc_result_rows(r);
        if (n > WC_REF_MAX_OUTPUT) n = WC_REF_MAX_OUTPUT; // the caller's buffer holds 10 rows (main.cu:171)
        for (size_t i = 0; i < n; i++) {
            const char *w; size_t len; int64_t c;
            wc_result_row(r, i, &w, &len, &c);
            memset(output[i].text, 0, WC_REF_WORD_BYTES); // main.cu:154 leaves the tail zero
            memcpy(output[i].text, w, len < WC_REF_WORD_BYTES - 1 ? len : WC_REF_WORD_BYTES - 1);
            output[i].count = (int)c;
        }
        *nOutputCount = (int)n;
        wc_result_free(r);
    }
    wc_destroy(e);
}
