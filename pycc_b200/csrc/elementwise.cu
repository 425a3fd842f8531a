// Memory-bound pieces of the CCSD iteration (HBM roofline): tau / spin-adapted amplitude
// combinations, pair symmetrisation, the denominator update + increment norm, the energy,
// and the DIIS vector algebra.  All kernels are grid-stride, coalesced along the last virtual
// index, 16-byte vectorised where the access pattern allows, and every reduction is two-stage
// with a fixed summation order (warp shuffle -> block -> one final block), so results are
// bitwise reproducible.  Reference formulas: pycc/ccwfn.py:287-294, 408-410, 621, 712-719;
// pycc/utils.py:26-46, 108-133.
#include "common.cuh"

namespace {

constexpr int RED_THREADS = 256;
constexpr int MAX_RED_BLOCKS = 148 * 8;

template <typename T>
struct Acc { using type = double; };

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// block-wide fixed-order sum; result valid in thread 0
__device__ __forceinline__ double block_sum(double v) {
    __shared__ double sh[RED_THREADS / 32];
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    double r = 0.0;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < RED_THREADS / 32; ++i) r += sh[i];
    }
    return r;
}

// second stage: one block sums `n` partials (stride `stride` apart, `nout` independent results)
__global__ void final_sum_kernel(const double* __restrict__ partials, int n, int nout, double* __restrict__ out) {
    for (int o = 0; o < nout; ++o) {
        double v = 0.0;
        for (int i = threadIdx.x; i < n; i += RED_THREADS) v += partials[(size_t)o * n + i];
        v = block_sum(v);
        if (threadIdx.x == 0) out[o] = v;
        __syncthreads();
    }
}

// ---- pair_combine: out[ijab] = c0 X[ijab] + c1 X[ijba] + c2 u[ia] w[jb] -----------------------
// One CTA per 32x32 (a,b) tile of one (i,j) pair; the transposed read goes through shared memory
// so both global reads are coalesced.
template <typename T>
__global__ void pair_combine_kernel(T* __restrict__ out, const T* __restrict__ X, double c0, double c1,
                                    const T* __restrict__ u, const T* __restrict__ w, double c2, int no, int nv,
                                    int tiles_per_dim) {
    __shared__ T tile[32][33];
    const long long ij = blockIdx.y;
    const int ta = blockIdx.x / tiles_per_dim, tb = blockIdx.x % tiles_per_dim;
    const int i = (int)(ij / no), j = (int)(ij % no);
    const T* Xij = X + ij * (long long)nv * nv;
    T* Oij = out + ij * (long long)nv * nv;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
    if (c1 != 0.0) {
        // tile[r][c] = X[b0 + r][a0 + c]  (rows are b, columns a)
        for (int r = ty; r < 32; r += 8) {
            const int b = tb * 32 + r, a = ta * 32 + tx;
            tile[r][tx] = (a < nv && b < nv) ? Xij[(long long)b * nv + a] : T(0);
        }
        __syncthreads();
    }
    for (int r = ty; r < 32; r += 8) {
        const int a = ta * 32 + r, b = tb * 32 + tx;
        if (a < nv && b < nv) {
            double v = 0.0;
            if (c0 != 0.0) v += c0 * (double)Xij[(long long)a * nv + b];
            if (c1 != 0.0) v += c1 * (double)tile[tx][r];
            if (u != nullptr) v += c2 * (double)u[(long long)i * nv + a] * (double)w[(long long)j * nv + b];
            Oij[(long long)a * nv + b] = (T)v;
        }
    }
}

// ---- pair_symmetrize: r2[ijab] += r2[jiba] in place --------------------------------------------
// CTA handles the tile pair {(i,j; A,B), (j,i; B,A)}; only canonical pairs do work.
template <typename T>
__global__ void pair_symmetrize_kernel(T* __restrict__ r2, int no, int nv, int tiles_per_dim) {
    __shared__ T tile[32][33];
    const long long ij = blockIdx.y;
    const int i = (int)(ij / no), j = (int)(ij % no);
    const int ta = blockIdx.x / tiles_per_dim, tb = blockIdx.x % tiles_per_dim;
    // canonical representative of the orbit {(i,j,ta,tb), (j,i,tb,ta)}
    if (i > j || (i == j && ta > tb)) return;
    const bool self = (i == j && ta == tb);
    T* P = r2 + ((long long)i * no + j) * nv * nv;   // [a][b]
    T* Q = r2 + ((long long)j * no + i) * nv * nv;   // partner, element [b][a]
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    // tile[r][c] = Q[b0 + r][a0 + c]
    for (int r = ty; r < 32; r += 8) {
        const int b = tb * 32 + r, a = ta * 32 + tx;
        tile[r][tx] = (a < nv && b < nv) ? Q[(long long)b * nv + a] : T(0);
    }
    __syncthreads();
    // s[a][b] = P[a][b] + Q[b][a]; write P now, stage s for the partner write
    T s_reg[4];
    int cnt = 0;
    for (int r = ty; r < 32; r += 8, ++cnt) {
        const int a = ta * 32 + r, b = tb * 32 + tx;
        T s = T(0);
        if (a < nv && b < nv) {
            s = P[(long long)a * nv + b] + tile[tx][r];
        }
        s_reg[cnt] = s;
    }
    __syncthreads();
    cnt = 0;
    for (int r = ty; r < 32; r += 8, ++cnt) {
        const int a = ta * 32 + r, b = tb * 32 + tx;
        if (a < nv && b < nv) P[(long long)a * nv + b] = s_reg[cnt];
        tile[r][tx] = s_reg[cnt];        // tile[a_local][b_local] = s
    }
    __syncthreads();
    if (self) return;
    for (int r = ty; r < 32; r += 8) {
        const int b = tb * 32 + r, a = ta * 32 + tx;
        if (a < nv && b < nv) Q[(long long)b * nv + a] = tile[tx][r];
    }
}

// ---- permute4: out[d0,d1,d2,d3] = alpha*in[strided] + beta*out ---------------------------------
template <typename T>
__global__ void permute4_kernel(T* __restrict__ out, const T* __restrict__ in, long long n0, long long n1,
                                long long n2, long long n3, long long s0, long long s1, long long s2, long long s3,
                                double alpha, double beta) {
    const long long total = n0 * n1 * n2 * n3;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        long long r = idx;
        const long long i3 = r % n3; r /= n3;
        const long long i2 = r % n2; r /= n2;
        const long long i1 = r % n1; r /= n1;
        const long long i0 = r;
        double v = alpha * (double)in[i0 * s0 + i1 * s1 + i2 * s2 + i3 * s3];
        if (beta != 0.0) v += beta * (double)out[idx];
        out[idx] = (T)v;
    }
}

// ---- amplitude update + squared norm of the increment ------------------------------------------
template <typename T>
__global__ void update_t2_kernel(T* __restrict__ t2, const T* __restrict__ r2, const T* __restrict__ eps, int no,
                                 int nv, int symmetrize, double* __restrict__ partials) {
    const long long nvv = (long long)nv * nv, total = (long long)no * no * nvv;
    double acc = 0.0;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long ij = idx / nvv, ab = idx - ij * nvv;
        const int i = (int)(ij / no), j = (int)(ij - (long long)i * no);
        const int a = (int)(ab / nv), b = (int)(ab - (long long)a * nv);
        double r = (double)r2[idx];
        if (symmetrize) r += (double)r2[((long long)j * no + i) * nvv + (long long)b * nv + a];
        const double d = (double)eps[i] + (double)eps[j] - (double)eps[no + a] - (double)eps[no + b];
        const double inc = (double)(T)(r / d);
        t2[idx] = (T)((double)t2[idx] + inc);
        acc += inc * inc;
    }
    acc = block_sum(acc);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

template <typename T>
__global__ void update_t1_kernel(T* __restrict__ t1, const T* __restrict__ r1, const T* __restrict__ eps, int no,
                                 int nv, double* __restrict__ partials) {
    const long long total = (long long)no * nv;
    double acc = 0.0;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(idx / nv), a = (int)(idx - (long long)i * nv);
        const double d = (double)eps[i] - (double)eps[no + a];
        const double inc = (double)(T)((double)r1[idx] / d);
        t1[idx] = (T)((double)t1[idx] + inc);
        acc += inc * inc;
    }
    acc = block_sum(acc);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

// ---- energy --------------------------------------------------------------------------------------
// One CTA per 32x32 (a,b) tile of one (i,j): E += (t2[ijab] + t1[ia] t1[jb]) (2 g[ijab] - g[ijba])
template <typename T>
__global__ void energy_kernel(const T* __restrict__ fov, long long ld_fov, const T* __restrict__ t1,
                              const T* __restrict__ t2, const T* __restrict__ g, int no, int nv, int tiles_per_dim,
                              double* __restrict__ partials) {
    __shared__ T tile[32][33];
    const long long ij = blockIdx.y;
    const int i = (int)(ij / no), j = (int)(ij % no);
    const int ta = blockIdx.x / tiles_per_dim, tb = blockIdx.x % tiles_per_dim;
    const T* gij = g + ij * (long long)nv * nv;
    const T* tij = t2 + ij * (long long)nv * nv;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int b = tb * 32 + r, a = ta * 32 + tx;
        tile[r][tx] = (a < nv && b < nv) ? gij[(long long)b * nv + a] : T(0);
    }
    __syncthreads();
    double acc = 0.0;
    for (int r = ty; r < 32; r += 8) {
        const int a = ta * 32 + r, b = tb * 32 + tx;
        if (a < nv && b < nv) {
            const double tau = (double)tij[(long long)a * nv + b] +
                               (double)t1[(long long)i * nv + a] * (double)t1[(long long)j * nv + b];
            const double L = 2.0 * (double)gij[(long long)a * nv + b] - (double)tile[tx][r];
            acc += tau * L;
        }
    }
    // singles part 2 F_ia t_ia: added once, by the (i, j=0, tile 0) CTAs, spread over threads
    if (j == 0 && blockIdx.x == 0) {
        for (int a = threadIdx.x; a < nv; a += blockDim.x)
            acc += 2.0 * (double)fov[(long long)i * ld_fov + a] * (double)t1[(long long)i * nv + a];
    }
    acc = block_sum(acc);
    if (threadIdx.x == 0) partials[(long long)blockIdx.y * gridDim.x + blockIdx.x] = acc;
}

// ---- DIIS vector algebra ---------------------------------------------------------------------
template <typename T>
__global__ void dot_kernel(const T* __restrict__ x, const T* __restrict__ y, long long n, double* __restrict__ partials) {
    double acc = 0.0;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n;
         idx += (long long)gridDim.x * blockDim.x)
        acc += (double)x[idx] * (double)y[idx];
    acc = block_sum(acc);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

template <typename T>
__global__ void axpby_kernel(T* __restrict__ out, double a, const T* __restrict__ x, double b, const T* __restrict__ y,
                             long long n) {
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n;
         idx += (long long)gridDim.x * blockDim.x) {
        double v = a * (double)x[idx];
        if (y != nullptr) v += b * (double)y[idx];
        out[idx] = (T)v;
    }
}

constexpr int MAX_LINCOMB = 16;
template <typename T>
struct LincombArgs {
    const T* v[MAX_LINCOMB];
    double c[MAX_LINCOMB];
    int n;
};

template <typename T>
__global__ void lincomb_kernel(T* __restrict__ out, const __grid_constant__ LincombArgs<T> args, long long n) {
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n;
         idx += (long long)gridDim.x * blockDim.x) {
        double v = 0.0;
        for (int k = 0; k < args.n; ++k) v += args.c[k] * (double)args.v[k][idx];
        out[idx] = (T)v;
    }
}

int grid_for(long long n, int threads, int cap) {
    long long b = (n + threads - 1) / threads;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

// sum `n` partials per output on the device, copy `nout` doubles to the pinned host slot and wait
int finish_reduction(pccb_ctx* ctx, int n, int nout, double* host_out) {
    double* dev_out = ctx->partial_dev + ctx->partial_cap - 64;   // last 64 doubles hold results
    final_sum_kernel<<<1, RED_THREADS, 0, ctx->stream>>>(ctx->partial_dev, n, nout, dev_out);
    PCCB_CHECK_LAUNCH(ctx, "final_sum_kernel");
    PCCB_CUDA(ctx, cudaMemcpyAsync(ctx->result_host, dev_out, sizeof(double) * nout, cudaMemcpyDeviceToHost,
                                   ctx->stream));
    PCCB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < nout; ++i) host_out[i] = ctx->result_host[i];
    return PCCB_OK;
}

template <typename T>
int pair_combine_t(pccb_ctx* ctx, void* out, const void* X, double c0, double c1, const void* u, const void* w,
                   double c2, int64_t no, int64_t nv) {
    const int tpd = (int)((nv + 31) / 32);
    dim3 grid(tpd * tpd, (unsigned)(no * no));
    pair_combine_kernel<T><<<grid, 256, 0, ctx->stream>>>((T*)out, (const T*)X, c0, c1, (const T*)u, (const T*)w, c2,
                                                          (int)no, (int)nv, tpd);
    PCCB_CHECK_LAUNCH(ctx, "pair_combine_kernel");
    return PCCB_OK;
}

template <typename T>
int update_t(pccb_ctx* ctx, void* t1, void* t2, const void* r1, const void* r2, const void* eps, int64_t no,
             int64_t nv, int symmetrize, double* rms2_host) {
    const long long n2 = no * no * nv * nv, n1 = no * nv;
    const int g2 = grid_for(n2, RED_THREADS, MAX_RED_BLOCKS), g1 = grid_for(n1, RED_THREADS, 64);
    int rc = pccb_ensure_partials(ctx, (size_t)g1 + g2 + 64);
    if (rc) return rc;
    update_t2_kernel<T><<<g2, RED_THREADS, 0, ctx->stream>>>((T*)t2, (const T*)r2, (const T*)eps, (int)no, (int)nv,
                                                             symmetrize, ctx->partial_dev);
    PCCB_CHECK_LAUNCH(ctx, "update_t2_kernel");
    update_t1_kernel<T><<<g1, RED_THREADS, 0, ctx->stream>>>((T*)t1, (const T*)r1, (const T*)eps, (int)no, (int)nv,
                                                             ctx->partial_dev + g2);
    PCCB_CHECK_LAUNCH(ctx, "update_t1_kernel");
    return finish_reduction(ctx, g1 + g2, 1, rms2_host);
}

template <typename T>
int energy_t(pccb_ctx* ctx, const void* fov, int64_t ld_fov, const void* t1, const void* t2, const void* g,
             int64_t no, int64_t nv, double* e_host) {
    const int tpd = (int)((nv + 31) / 32);
    dim3 grid(tpd * tpd, (unsigned)(no * no));
    const size_t nb = (size_t)grid.x * grid.y;
    int rc = pccb_ensure_partials(ctx, nb + 64);
    if (rc) return rc;
    energy_kernel<T><<<grid, 256, 0, ctx->stream>>>((const T*)fov, ld_fov, (const T*)t1, (const T*)t2, (const T*)g,
                                                    (int)no, (int)nv, tpd, ctx->partial_dev);
    PCCB_CHECK_LAUNCH(ctx, "energy_kernel");
    return finish_reduction(ctx, (int)nb, 1, e_host);
}

}  // namespace

int pccb_ensure_partials(pccb_ctx* ctx, size_t ndoubles) {
    if (ndoubles + 64 <= ctx->partial_cap) return PCCB_OK;
    if (ctx->partial_dev) cudaFree(ctx->partial_dev);
    ctx->partial_dev = nullptr;
    size_t cap = ndoubles + 64;
    if (cap < (1u << 16)) cap = 1u << 16;
    PCCB_CUDA(ctx, cudaMalloc(&ctx->partial_dev, cap * sizeof(double)));
    ctx->partial_cap = cap;
    return PCCB_OK;
}

#define DISPATCH_DTYPE(ctx, dtype, expr_f64, expr_f32)                          \
    do {                                                                        \
        if ((dtype) == PCCB_F64) return (expr_f64);                             \
        if ((dtype) == PCCB_F32) return (expr_f32);                             \
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", (int)(dtype));         \
    } while (0)

extern "C" int pccb_pair_combine(pccb_ctx* ctx, int dtype, void* out_dev, const void* X_dev, double c0, double c1,
                                 const void* u_dev, const void* w_dev, double c2, int64_t no, int64_t nv) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, out_dev && X_dev && no > 0 && nv > 0, "pccb_pair_combine: bad arguments");
    PCCB_REQUIRE(ctx, (u_dev == nullptr) == (w_dev == nullptr), "pccb_pair_combine: u and w must both be given");
    PCCB_REQUIRE(ctx, out_dev != X_dev || c1 == 0.0, "pccb_pair_combine: in-place only without the transposed term");
    PCCB_REQUIRE(ctx, no * no <= 65535, "pccb_pair_combine: no too large");
    DISPATCH_DTYPE(ctx, dtype, pair_combine_t<double>(ctx, out_dev, X_dev, c0, c1, u_dev, w_dev, c2, no, nv),
                   pair_combine_t<float>(ctx, out_dev, X_dev, c0, c1, u_dev, w_dev, c2, no, nv));
}

extern "C" int pccb_pair_symmetrize(pccb_ctx* ctx, int dtype, void* r2_dev, int64_t no, int64_t nv) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, r2_dev && no > 0 && nv > 0 && no * no <= 65535, "pccb_pair_symmetrize: bad arguments");
    const int tpd = (int)((nv + 31) / 32);
    dim3 grid(tpd * tpd, (unsigned)(no * no));
    if (dtype == PCCB_F64)
        pair_symmetrize_kernel<double><<<grid, 256, 0, ctx->stream>>>((double*)r2_dev, (int)no, (int)nv, tpd);
    else if (dtype == PCCB_F32)
        pair_symmetrize_kernel<float><<<grid, 256, 0, ctx->stream>>>((float*)r2_dev, (int)no, (int)nv, tpd);
    else
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    PCCB_CHECK_LAUNCH(ctx, "pair_symmetrize_kernel");
    return PCCB_OK;
}

extern "C" int pccb_permute4(pccb_ctx* ctx, int dtype, void* out_dev, const void* in_dev, const int64_t shape[4],
                             const int64_t in_strides[4], double alpha, double beta) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, out_dev && in_dev && shape && in_strides, "pccb_permute4: null argument");
    const long long total = shape[0] * shape[1] * shape[2] * shape[3];
    PCCB_REQUIRE(ctx, total > 0, "pccb_permute4: empty shape");
    const int g = grid_for(total, 256, 148 * 16);
    if (dtype == PCCB_F64)
        permute4_kernel<double><<<g, 256, 0, ctx->stream>>>((double*)out_dev, (const double*)in_dev, shape[0],
                                                            shape[1], shape[2], shape[3], in_strides[0],
                                                            in_strides[1], in_strides[2], in_strides[3], alpha, beta);
    else if (dtype == PCCB_F32)
        permute4_kernel<float><<<g, 256, 0, ctx->stream>>>((float*)out_dev, (const float*)in_dev, shape[0], shape[1],
                                                           shape[2], shape[3], in_strides[0], in_strides[1],
                                                           in_strides[2], in_strides[3], alpha, beta);
    else
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    PCCB_CHECK_LAUNCH(ctx, "permute4_kernel");
    return PCCB_OK;
}

extern "C" int pccb_update_amplitudes(pccb_ctx* ctx, int dtype, void* t1_dev, void* t2_dev, const void* r1_dev,
                                      const void* r2_dev, const void* eps_dev, int64_t no, int64_t nv,
                                      int symmetrize, double* rms2_host) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, t1_dev && t2_dev && r1_dev && r2_dev && eps_dev && rms2_host && no > 0 && nv > 0,
                 "pccb_update_amplitudes: bad arguments");
    DISPATCH_DTYPE(ctx, dtype,
                   update_t<double>(ctx, t1_dev, t2_dev, r1_dev, r2_dev, eps_dev, no, nv, symmetrize, rms2_host),
                   update_t<float>(ctx, t1_dev, t2_dev, r1_dev, r2_dev, eps_dev, no, nv, symmetrize, rms2_host));
}

extern "C" int pccb_cc_energy(pccb_ctx* ctx, int dtype, const void* fov_dev, int64_t ld_fov, const void* t1_dev,
                              const void* t2_dev, const void* oovv_dev, int64_t no, int64_t nv, double* e_host) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, fov_dev && t1_dev && t2_dev && oovv_dev && e_host && no > 0 && nv > 0 && no * no <= 65535,
                 "pccb_cc_energy: bad arguments");
    DISPATCH_DTYPE(ctx, dtype, energy_t<double>(ctx, fov_dev, ld_fov, t1_dev, t2_dev, oovv_dev, no, nv, e_host),
                   energy_t<float>(ctx, fov_dev, ld_fov, t1_dev, t2_dev, oovv_dev, no, nv, e_host));
}

extern "C" int pccb_dot(pccb_ctx* ctx, int dtype, const void* x_dev, const void* y_dev, int64_t n, double* out_host) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, x_dev && y_dev && out_host && n > 0, "pccb_dot: bad arguments");
    const int g = grid_for(n, RED_THREADS, MAX_RED_BLOCKS);
    int rc = pccb_ensure_partials(ctx, (size_t)g + 64);
    if (rc) return rc;
    if (dtype == PCCB_F64)
        dot_kernel<double><<<g, RED_THREADS, 0, ctx->stream>>>((const double*)x_dev, (const double*)y_dev, n,
                                                               ctx->partial_dev);
    else if (dtype == PCCB_F32)
        dot_kernel<float><<<g, RED_THREADS, 0, ctx->stream>>>((const float*)x_dev, (const float*)y_dev, n,
                                                              ctx->partial_dev);
    else
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    PCCB_CHECK_LAUNCH(ctx, "dot_kernel");
    return finish_reduction(ctx, g, 1, out_host);
}

extern "C" int pccb_axpby(pccb_ctx* ctx, int dtype, void* out_dev, double a, const void* x_dev, double b,
                          const void* y_dev, int64_t n) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, out_dev && x_dev && n > 0, "pccb_axpby: bad arguments");
    const int g = grid_for(n, 256, 148 * 16);
    if (dtype == PCCB_F64)
        axpby_kernel<double><<<g, 256, 0, ctx->stream>>>((double*)out_dev, a, (const double*)x_dev, b,
                                                         (const double*)y_dev, n);
    else if (dtype == PCCB_F32)
        axpby_kernel<float><<<g, 256, 0, ctx->stream>>>((float*)out_dev, a, (const float*)x_dev, b,
                                                        (const float*)y_dev, n);
    else
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    PCCB_CHECK_LAUNCH(ctx, "axpby_kernel");
    return PCCB_OK;
}

extern "C" int pccb_lincomb(pccb_ctx* ctx, int dtype, void* out_dev, const void* const* vecs, const double* coef,
                            int nvec, int64_t n) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, out_dev && vecs && coef && nvec >= 1 && nvec <= MAX_LINCOMB && n > 0,
                 "pccb_lincomb: bad arguments (nvec=%d)", nvec);
    const int g = grid_for(n, 256, 148 * 16);
    if (dtype == PCCB_F64) {
        LincombArgs<double> a;
        a.n = nvec;
        for (int k = 0; k < nvec; ++k) { a.v[k] = (const double*)vecs[k]; a.c[k] = coef[k]; }
        lincomb_kernel<double><<<g, 256, 0, ctx->stream>>>((double*)out_dev, a, n);
    } else if (dtype == PCCB_F32) {
        LincombArgs<float> a;
        a.n = nvec;
        for (int k = 0; k < nvec; ++k) { a.v[k] = (const float*)vecs[k]; a.c[k] = coef[k]; }
        lincomb_kernel<float><<<g, 256, 0, ctx->stream>>>((float*)out_dev, a, n);
    } else {
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    }
    PCCB_CHECK_LAUNCH(ctx, "lincomb_kernel");
    return PCCB_OK;
}
