/*
 * TEST INFRASTRUCTURE ONLY (oracle/): a stand-in <cuda_runtime.h> that lets the
 * reference's unmodified main.cu be compiled by g++ and executed on host cores.
 *
 * It is written from scratch; nothing here comes from the reference or from the
 * CUDA toolkit.  It provides exactly the API surface main.cu touches
 * (/root/reference/main.cu:37-162): the __device__/__global__ qualifiers, the
 * threadIdx/blockIdx/blockDim/gridDim built-ins, cudaMalloc/cudaFree/
 * cudaMemcpy/cudaMemset, and a WC_SHIM_LAUNCH macro that replaces the two
 * `kernel << <G, B >> >(args)` launches (main.cu:125,130) after a one-line sed
 * in oracle/Makefile.  Blocks and threads are run one after the other, which is
 * a legal schedule of the reference's kernels (they do not synchronise).
 *
 * cudaMalloc returns zeroed memory: the reference relies on fresh device memory
 * being zero (main.cu:46-52 writes no terminator, main.cu:144 never memsets
 * dev_pairs) - SURVEY.md A.4 R1.
 */
#ifndef WC_ORACLE_SHIM_CUDA_RUNTIME_H
#define WC_ORACLE_SHIM_CUDA_RUNTIME_H

#include <cstdlib>
#include <cstring>
#include <cstddef>

#define __device__
#define __global__
#define __host__

struct wc_shim_dim3 { unsigned x, y, z; };
static wc_shim_dim3 threadIdx = {0, 0, 0}, blockIdx = {0, 0, 0};
static wc_shim_dim3 blockDim = {1, 1, 1}, gridDim = {1, 1, 1};

enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1,
                      cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };
typedef int cudaError_t;

template <class T> static inline cudaError_t cudaMalloc(T **p, size_t n) {
    *p = static_cast<T *>(calloc(1, n ? n : 1));
    return *p ? 0 : 2;
}
static inline cudaError_t cudaFree(void *p) { free(p); return 0; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) {
    memcpy(d, s, n);
    return 0;
}
static inline cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return 0; }

#define WC_SHIM_LAUNCH(kernel, G, B, ...)                                       \
    do {                                                                        \
        gridDim.x = (G); blockDim.x = (B);                                      \
        for (unsigned wc_b = 0; wc_b < (unsigned)(G); ++wc_b)                   \
            for (unsigned wc_t = 0; wc_t < (unsigned)(B); ++wc_t) {             \
                blockIdx.x = wc_b; threadIdx.x = wc_t;                          \
                kernel(__VA_ARGS__);                                            \
            }                                                                   \
    } while (0)

#endif
