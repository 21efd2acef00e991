// Shared helpers for the nanotensor_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/nanotensor_b200.h"

namespace ntb {

typedef __half op16_t;   // storage type of the split tensor-core operands (fp16 hi / lo planes)

extern thread_local char g_err[512];

inline int fail(int code, const char *fmt, const char *a = "", long long b = 0, long long c = 0) {
    snprintf(g_err, sizeof(g_err), fmt, a, b, c);
    return code;
}

#define NTB_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            snprintf(ntb::g_err, sizeof(ntb::g_err), "%s:%d: %s -> %s", __FILE__, __LINE__,   \
                     #expr, cudaGetErrorString(_e));                                         \
            return NTB_ECUDA;                                                                \
        }                                                                                    \
    } while (0)

#define NTB_LAUNCH_CHECK()                                                                   \
    do {                                                                                     \
        cudaError_t _e = cudaGetLastError();                                                 \
        if (_e != cudaSuccess) {                                                             \
            snprintf(ntb::g_err, sizeof(ntb::g_err), "%s:%d: kernel launch -> %s", __FILE__,  \
                     __LINE__, cudaGetErrorString(_e));                                      \
            return NTB_ECUDA;                                                                \
        }                                                                                    \
    } while (0)

#define NTB_TRY(expr)                \
    do {                             \
        int _rc = (expr);            \
        if (_rc != NTB_OK) return _rc; \
    } while (0)

// Ablation / A-B switches (NTB_LSTM_FLAGS, NTB_XPROJ_ABL, NTB_CF_ABL, ... -- several of them change RESULTS: they skip
// barrier waits, stores or whole stages to time what is left).  The production build never reads them: ntb_env()
// returns NULL unless the library was built with -DNTB_ABLATION (csrc/build.sh: NTB_ABLATION=1), so a stray
// environment variable cannot produce silently wrong labels with rc 0.
static inline const char *ntb_env(const char *name) {
#ifdef NTB_ABLATION
    return getenv(name);
#else
    (void)name;
    return nullptr;
#endif
}

static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }
static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// Bump allocator over a caller-provided device buffer (the library never cudaMallocs).
struct Arena {
    char *base;
    int64_t cap;
    int64_t off;
    Arena(void *p, int64_t bytes) : base((char *)p), cap(bytes), off(0) {}
    template <typename T>
    T *take(int64_t n) {
        int64_t bytes = align_up(n * (int64_t)sizeof(T), 256);
        T *p = (T *)(base ? base + off : nullptr);
        off += bytes;
        return p;
    }
    bool ok() const { return off <= cap; }
};

__device__ __forceinline__ float sigmoid_acc(float x) {
    // 1/(1+exp(-x)); ex2.approx + rcp.approx: abs error ~2e-7, far inside the 1e-3 logits budget
    return __fdividef(1.0f, 1.0f + __expf(-x));
}
__device__ __forceinline__ float tanh_acc(float x) {
    // tanh(x) = 2*sigmoid(2x) - 1 ; tanh.approx (2^-11) is too coarse for a 100-step recurrence
    return 2.0f * __fdividef(1.0f, 1.0f + __expf(-2.0f * x)) - 1.0f;
}

}  // namespace ntb
