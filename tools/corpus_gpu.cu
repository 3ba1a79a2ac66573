// tools/corpus_gpu.cu -- the Zipf block writer of tools/corpus.c on the GPU (test / bench tooling, not
// part of the product).  SURVEY.md 8(d) G6: blocks are generated independently from a counter-based
// RNG, so the same bytes come out on the host and on the device; here one thread writes one 1 MiB
// block.  Used to put the 16 GB / 100 GB corpora of configs 3 and 4 straight into HBM without
// generating and copying them on the host.  tests/test_gpu_parity.py checks it byte for byte
// against corpus_fill().
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define ZIPF_BLOCK (1u << 20)
#define WORD_STRIDE 20

struct Tables {
    const uint8_t *words, *lens;
    const uint32_t *alias, *thresh;
    uint32_t by_len[20];
    uint64_t fill_max, text_seed;
    uint32_t vocab;
};

__host__ __device__ static inline uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

__device__ static uint8_t *fill_exact(const Tables &T, uint8_t *p, size_t rem) {
    while (rem > 0) {
        size_t line = rem < 99 ? rem : 99;
        if (rem - line == 1) line -= 1;
        size_t content = line - 1;
        while (content > 0) {
            size_t w = content < T.fill_max ? content : (size_t)T.fill_max;
            if (content - w == 1) w -= 1;
            const uint8_t *src = T.words + (size_t)T.by_len[w] * WORD_STRIDE;
            for (size_t i = 0; i < w; i++) p[i] = src[i];
            p += w; content -= w;
            if (content > 0) { *p++ = ' '; content--; }
        }
        *p++ = '\n';
        rem -= line;
    }
    return p;
}

__global__ void zipf_blocks_kernel(Tables T, uint8_t *dst, uint64_t off, uint64_t b0, uint64_t b1, uint64_t total) {
    for (uint64_t b = b0 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < b1; b += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t lo = b * ZIPF_BLOCK, hi = lo + ZIPF_BLOCK > total ? total : lo + ZIPF_BLOCK;
        uint8_t *p = dst + (lo - off), *end = p + (hi - lo);
        const uint64_t key = mix64(T.text_seed) ^ mix64(b * 0x632BE59BD9B4E019ull + 1);
        uint64_t ctr = 0;
        size_t col = 0, width = 0;
        while ((size_t)(end - p) > 64) {
            const uint64_t x = mix64(key + (ctr++) * 0xD1342543DE82EF95ull);
            if (col == 0) width = 40 + (size_t)((x >> 52) % 59);
            const uint32_t c0 = (uint32_t)(((x >> 32) * (uint64_t)T.vocab) >> 32);
            const uint32_t r = ((uint32_t)x <= T.thresh[c0]) ? c0 : T.alias[c0];
            const size_t len = T.lens[r];
            if (col && col + 1 + len > width) { *p++ = '\n'; col = 0; width = 40 + (size_t)((x >> 52) % 59); }
            if (col) { *p++ = ' '; col++; }
            const uint8_t *src = T.words + (size_t)r * WORD_STRIDE;
            for (size_t i = 0; i < len; i++) p[i] = src[i];
            p += len; col += len;
        }
        if (col) *p++ = '\n';
        fill_exact(T, p, (size_t)(end - p));
    }
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "corpus_gpu: %s: %s\n", #x, cudaGetErrorString(e_)); rc = -10; goto done; } } while (0)

// Writes bytes [off, off + n) of the `total`-byte Zipf corpus into device memory at dev_dst (same
// alignment rules as corpus_fill).  Tables are host pointers from corpus_tables().
extern "C" int corpus_gpu_fill(const uint8_t *words, const uint8_t *lens, const uint32_t *by_len, uint64_t fill_max,
                               const uint32_t *alias, const uint32_t *thresh, uint32_t vocab, uint64_t text_seed,
                               void *dev_dst, uint64_t off, uint64_t n, uint64_t total) {
    if (off % ZIPF_BLOCK) return -1;
    if (off + n != total && (off + n) % ZIPF_BLOCK) return -2;
    if (off + n > total) return -3;
    if (total % ZIPF_BLOCK && total % ZIPF_BLOCK < 2) return -4;
    if (n == 0) return 0;
    int rc = 0;
    Tables T;
    uint8_t *d_words = nullptr, *d_lens = nullptr;
    uint32_t *d_alias = nullptr, *d_thresh = nullptr;
    {
        CK(cudaMalloc((void **)&d_words, (size_t)vocab * WORD_STRIDE));
        CK(cudaMalloc((void **)&d_lens, vocab));
        CK(cudaMalloc((void **)&d_alias, (size_t)vocab * 4));
        CK(cudaMalloc((void **)&d_thresh, (size_t)vocab * 4));
        CK(cudaMemcpy(d_words, words, (size_t)vocab * WORD_STRIDE, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_lens, lens, vocab, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_alias, alias, (size_t)vocab * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_thresh, thresh, (size_t)vocab * 4, cudaMemcpyHostToDevice));
        T.words = d_words; T.lens = d_lens; T.alias = d_alias; T.thresh = d_thresh;
        for (int i = 0; i < 20; i++) T.by_len[i] = by_len[i];
        T.fill_max = fill_max; T.text_seed = text_seed; T.vocab = vocab;
        const uint64_t b0 = off / ZIPF_BLOCK, b1 = (off + n + ZIPF_BLOCK - 1) / ZIPF_BLOCK;
        const uint64_t nb = b1 - b0;
        const unsigned grid = (unsigned)((nb + 63) / 64 < 65535 ? (nb + 63) / 64 : 65535);
        zipf_blocks_kernel<<<grid, 64>>>(T, (uint8_t *)dev_dst, off, b0, b1, total);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
    }
done:
    cudaFree(d_words); cudaFree(d_lens); cudaFree(d_alias); cudaFree(d_thresh);
    return rc;
}
