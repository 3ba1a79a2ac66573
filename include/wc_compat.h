/*
 * wc_compat.h -- the reference's own operator entry point, served by libwc.so.
 *
 *     void runMapReduce(InputData *input, int nInputCount, OutputData *output, int *nOutputCount);
 *
 * is the one function main() calls in /root/reference/main.cu (declared main.cu:133, called
 * main.cu:209).  libwc.so exports it with the same name, argument meaning and struct layouts
 * (main.cu:16-33: Word 30 B, KeyPair 36 B with nCount at 32, InputData 604 B with nWordCount at
 * 600), so the reference's main() links against libwc.so unchanged.  Like the
 * reference it reports no errors; on failure it leaves *nOutputCount = 0.
 *
 * The types below restate that layout under their own names; they are checked by static_asserts in
 * wc_compat.cpp.
 */
#ifndef WC_COMPAT_H
#define WC_COMPAT_H

#ifdef __cplusplus
extern "C" {
#endif

#define WC_REF_WORD_BYTES 30     /* main.cu:17,20 */
#define WC_REF_WORDS_PER_LINE 20 /* main.cu:25 */
#define WC_REF_MAX_OUTPUT 10     /* main.cu:13 */

typedef struct wc_ref_word { char text[WC_REF_WORD_BYTES]; } wc_ref_word;             /* main.cu:16-18 */
typedef struct wc_ref_keypair { char text[WC_REF_WORD_BYTES]; int count; } wc_ref_keypair; /* main.cu:19-22 */
typedef struct wc_ref_input {                                                          /* main.cu:24-27 */
    wc_ref_word words[WC_REF_WORDS_PER_LINE];
    int n_words;
} wc_ref_input;
typedef wc_ref_keypair wc_ref_output; /* main.cu:28 */

#if defined(__GNUC__)
__attribute__((visibility("default")))
#endif
void runMapReduce(wc_ref_input *input, int nInputCount, wc_ref_output *output, int *nOutputCount);

#ifdef __cplusplus
}
#endif
#endif
