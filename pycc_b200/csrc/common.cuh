// Internal definitions shared by the translation units of libpyccb200.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/pyccb200.h"

// an FP32 tensor [base, base + bytes) whose TF32 residual plane lo = x - trunc_tf32(x) lives at `lo`
struct pccb_lo_plane {
    const char* base;
    const char* lo;
    size_t bytes;
};

struct pccb_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    void* ws = nullptr;           // caller-owned split-K workspace
    size_t ws_bytes = 0;
    int impl = PCCB_IMPL_TENSOR;
    long long launches = 0;
    double* partial_dev = nullptr;   // small library-owned scratch for fixed-order reductions
    double* result_host = nullptr;   // pinned
    size_t partial_cap = 0;          // doubles
    std::vector<pccb_lo_plane> lo_planes;   // registered by pccb_register_lo_plane (tcgen05 3xTF32 path)
    std::string err;
};

extern thread_local std::string g_create_error;

#define PCCB_FAIL(ctx, code, ...)                                   \
    do {                                                            \
        char buf_[512];                                             \
        snprintf(buf_, sizeof(buf_), __VA_ARGS__);                  \
        (ctx)->err = buf_;                                          \
        return (code);                                              \
    } while (0)

#define PCCB_CUDA(ctx, call)                                                                   \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            PCCB_FAIL(ctx, PCCB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                      __FILE__, __LINE__);                                                     \
    } while (0)

#define PCCB_CHECK_LAUNCH(ctx, what)                                                            \
    do {                                                                                        \
        cudaError_t e_ = cudaGetLastError();                                                    \
        if (e_ != cudaSuccess)                                                                  \
            PCCB_FAIL(ctx, PCCB_ERR_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(e_)); \
        (ctx)->launches++;                                                                      \
    } while (0)

#define PCCB_REQUIRE(ctx, cond, ...)                          \
    do {                                                      \
        if (!(cond)) PCCB_FAIL(ctx, PCCB_ERR_ARG, __VA_ARGS__); \
    } while (0)

// device-side compound index (see pccb_cidx): offset(i) = (i / div) * s_hi + (i % div) * s_lo.
// The division is a multiply-high + shift with a host-computed magic number (exact for
// i < 2^31, 2 <= div < 2^31); a flat index has magic = 0 (quotient 0).
struct DIdx {
    long long s_hi, s_lo;
    unsigned div, magic, shift, flat;
};

__host__ __device__ __forceinline__ long long didx_off(const DIdx& c, unsigned i) {
#ifdef __CUDA_ARCH__
    const unsigned q = __umulhi(i, c.magic) >> c.shift;
#else
    const unsigned q = (unsigned)(((unsigned long long)i * c.magic) >> 32) >> c.shift;
#endif
    const unsigned r = i - q * c.div;
    return (long long)q * c.s_hi + (long long)r * c.s_lo;
}

static inline DIdx to_didx(const pccb_cidx& c, long long extent) {
    DIdx d;
    long long div = c.div;
    d.s_lo = c.s_lo;
    if (div <= 1 || div >= extent) {   // flat
        d.div = 0;
        d.magic = 0;
        d.shift = 0;
        d.s_hi = 0;
        d.flat = 1;
        if (div == 1 && extent > 1) {  // degenerate "inner extent 1": index is the outer one
            d.s_lo = c.s_hi;
        }
        return d;
    }
    unsigned s = 0;
    while ((1ll << s) < div) ++s;                         // s = ceil(log2 div) >= 1
    const unsigned long long m = ((1ull << (31 + s)) + (unsigned long long)div - 1) / (unsigned long long)div;
    d.div = (unsigned)div;
    d.magic = (unsigned)m;                               // < 2^32, see contract.cu header
    d.shift = s - 1;
    d.s_hi = c.s_hi;
    d.flat = 0;
    return d;
}

// fixed-order two-stage sum reduction helpers (reduce.cu)
int pccb_ensure_partials(pccb_ctx* ctx, size_t ndoubles);

// FP32-operand contraction (contract_f32.cu); arguments already validated by pccb_contract
int pccb_contract_f32(pccb_ctx* ctx, const pccb_contraction* desc);

// tcgen05 / TMEM 3xTF32 kernel for plain-matrix FP32 contractions whose operands have registered lo planes
// (contract_tc05.cu): 1 = launched, 0 = not eligible, < 0 = error
int pccb_contract_f32_tc05(pccb_ctx* ctx, const pccb_contraction* desc);
