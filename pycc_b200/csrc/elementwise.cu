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
// One CTA (256 threads, 32 x 8) per 64 x 64 (a,b) tile of one (i,j) pair; the transposed read goes through shared
// memory so both global reads are coalesced.  Every thread has its 16 elements of each stream in flight before the
// barrier: with 32 x 32 tiles (4 per thread) these kernels were latency-bound at 0.54 - 0.63 of the HBM roofline.
constexpr int PT = 64;                    // pair-kernel tile edge
#define PAIR_K_R(k) (ty + 8 * ((k) >> 1))
#define PAIR_K_C(k) (tx + 32 * ((k) & 1))

template <typename T>
__global__ void __launch_bounds__(256) pair_combine_kernel(T* __restrict__ out, const T* __restrict__ X, double c0, double c1,
                                                           const T* __restrict__ u, const T* __restrict__ w, double c2, int no,
                                                           int nv, int tiles_per_dim) {
    __shared__ T tile[PT][PT + 1];
    const long long ij = blockIdx.y;
    const int ta = blockIdx.x / tiles_per_dim, tb = blockIdx.x % tiles_per_dim;
    const int i = (int)(ij / no), j = (int)(ij % no);
    const T* Xij = X + ij * (long long)nv * nv;
    T* Oij = out + ij * (long long)nv * nv;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
    T xt[16], xd[16];
    const bool want_t = c1 != 0.0, want_d = c0 != 0.0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {        // tile[r][c] = X[b0 + r][a0 + c]  (rows are b, columns a)
        const int b = tb * PT + PAIR_K_R(k), a = ta * PT + PAIR_K_C(k);
        xt[k] = (want_t && a < nv && b < nv) ? Xij[(long long)b * nv + a] : T(0);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int a = ta * PT + PAIR_K_R(k), b = tb * PT + PAIR_K_C(k);
        xd[k] = (want_d && a < nv && b < nv) ? Xij[(long long)a * nv + b] : T(0);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) tile[PAIR_K_R(k)][PAIR_K_C(k)] = xt[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int r = PAIR_K_R(k), c = PAIR_K_C(k);
        const int a = ta * PT + r, b = tb * PT + c;
        if (a < nv && b < nv) {
            double v = 0.0;
            if (want_d) v += c0 * (double)xd[k];
            if (want_t) v += c1 * (double)tile[c][r];
            if (u != nullptr) v += c2 * (double)u[(long long)i * nv + a] * (double)w[(long long)j * nv + b];
            Oij[(long long)a * nv + b] = (T)v;
        }
    }
}

// ---- pair_symmetrize: r2[ijab] += r2[jiba] in place --------------------------------------------
// CTA handles the tile pair {(i,j; A,B), (j,i; B,A)}; only canonical pairs do work.
template <typename T>
__global__ void __launch_bounds__(256) pair_symmetrize_kernel(T* __restrict__ r2, int no, int nv, int tiles_per_dim) {
    __shared__ T tile[PT][PT + 1];
    const long long ij = blockIdx.y;
    const int i = (int)(ij / no), j = (int)(ij % no);
    const int ta = blockIdx.x / tiles_per_dim, tb = blockIdx.x % tiles_per_dim;
    // canonical representative of the orbit {(i,j,ta,tb), (j,i,tb,ta)}
    if (i > j || (i == j && ta > tb)) return;
    const bool self = (i == j && ta == tb);
    T* P = r2 + ((long long)i * no + j) * nv * nv;   // [a][b]
    T* Q = r2 + ((long long)j * no + i) * nv * nv;   // partner, element [b][a]
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    T q[16], s[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {        // tile[r][c] = Q[b0 + r][a0 + c]
        const int b = tb * PT + PAIR_K_R(k), a = ta * PT + PAIR_K_C(k);
        q[k] = (a < nv && b < nv) ? Q[(long long)b * nv + a] : T(0);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int a = ta * PT + PAIR_K_R(k), b = tb * PT + PAIR_K_C(k);
        s[k] = (a < nv && b < nv) ? P[(long long)a * nv + b] : T(0);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) tile[PAIR_K_R(k)][PAIR_K_C(k)] = q[k];
    __syncthreads();
    // s[a][b] = P[a][b] + Q[b][a]; write P now, stage s for the partner write
#pragma unroll
    for (int k = 0; k < 16; ++k) s[k] += tile[PAIR_K_C(k)][PAIR_K_R(k)];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int r = PAIR_K_R(k), c = PAIR_K_C(k);
        const int a = ta * PT + r, b = tb * PT + c;
        if (a < nv && b < nv) P[(long long)a * nv + b] = s[k];
        tile[r][c] = s[k];        // tile[a_local][b_local] = s
    }
    __syncthreads();
    if (self) return;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int r = PAIR_K_R(k), c = PAIR_K_C(k);
        const int b = tb * PT + r, a = ta * PT + c;
        if (a < nv && b < nv) Q[(long long)b * nv + a] = tile[c][r];
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

// Two faster forms of the same operation (pccb_permute4 picks one; the kernel above keeps the shapes they do not take):
//  * rows: the input's contiguous index is also the output's last one.  One warp per output row, lanes along it; the
//    row index is decomposed once per row instead of four 64-bit divisions per element (the generic kernel ran at
//    0.46 of the HBM roofline on [i,a,m,e] <- [i,m,a,e] although both sides are coalesced).
//  * tiles: the input's contiguous index q is another output index.  64 x 64 tiles over (q, last) through shared
//    memory, read with lanes along q and written with lanes along the last index; 16 loads per thread are in flight
//    before the barrier (a 32 x 32 tile leaves 4 per thread: latency-bound).
template <typename T>
__global__ void __launch_bounds__(256) permute4_rows_kernel(T* __restrict__ out, const T* __restrict__ in, unsigned n1, unsigned n2,
                                                            unsigned n3, long long s0, long long s1, long long s2, long long s3,
                                                            unsigned rows, double alpha, double beta) {
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (unsigned row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
        const unsigned i2 = row % n2, r = row / n2, i1 = r % n1, i0 = r / n1;
        const T* src = in + i0 * s0 + i1 * s1 + i2 * s2;
        T* dst = out + (long long)row * n3;
        for (unsigned c = lane; c < n3; c += 128) {
            T x[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) x[u] = (c + 32 * u < n3) ? src[(long long)(c + 32 * u) * s3] : T(0);
            if (beta != 0.0) {
                T y[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) y[u] = (c + 32 * u < n3) ? dst[c + 32 * u] : T(0);
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (c + 32 * u < n3) dst[c + 32 * u] = (T)(alpha * (double)x[u] + beta * (double)y[u]);
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (c + 32 * u < n3) dst[c + 32 * u] = (T)(alpha * (double)x[u]);
            }
        }
    }
}

// q: the output dimension (0..2) along which the input is contiguous; u, w: the two remaining output dimensions (grid y, z)
template <typename T>
__global__ void __launch_bounds__(256) permute4_tiles_kernel(T* __restrict__ out, const T* __restrict__ in, unsigned nq, unsigned n3,
                                                             unsigned tq, long long sq, long long s3, long long su, long long sw,
                                                             long long oq, long long ou, long long ow, double alpha, double beta) {
    __shared__ T tile[64][65];
    const unsigned tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
    const unsigned q0 = (blockIdx.x % tq) * 64, c0 = (blockIdx.x / tq) * 64;
    const T* src = in + blockIdx.y * su + blockIdx.z * sw;
    T* dst = out + blockIdx.y * ou + blockIdx.z * ow;
    T x[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {            // row r of the tile = last index, columns = q
        const unsigned r = ty + 8 * (k >> 1), iq = q0 + tx + 32 * (k & 1), i3 = c0 + r;
        x[k] = (iq < nq && i3 < n3) ? src[(long long)iq * sq + (long long)i3 * s3] : T(0);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) tile[ty + 8 * (k >> 1)][tx + 32 * (k & 1)] = x[k];
    __syncthreads();
    if (beta != 0.0) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const unsigned r = ty + 8 * (k >> 1), iq = q0 + r, i3 = c0 + tx + 32 * (k & 1);
            x[k] = (iq < nq && i3 < n3) ? dst[(long long)iq * oq + i3] : T(0);
        }
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const unsigned r = ty + 8 * (k >> 1), iq = q0 + r, i3 = c0 + tx + 32 * (k & 1);
        if (iq < nq && i3 < n3) {
            double v = alpha * (double)tile[tx + 32 * (k & 1)][r];
            if (beta != 0.0) v += beta * (double)x[k];
            dst[(long long)iq * oq + i3] = (T)v;
        }
    }
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
constexpr int MAX_LINCOMB = 16;
template <typename T>
struct LincombArgs {
    const T* v[MAX_LINCOMB];
    double c[MAX_LINCOMB];
    int n;
};


// ---- 16-byte vectorised streaming kernels (DIIS algebra) ---------------------------------------------
// VEC elements of T make one 16-byte access (2 doubles / 4 floats); the host picks VEC = 1 when a pointer
// is not 16-byte aligned.  The tail (n % VEC) is handled by scalar accesses of the last threads.
template <typename T, int VEC> struct alignas(sizeof(T) * VEC) Pack { T v[VEC]; };

template <typename T, int VEC>
__global__ void dot_vec_kernel(const T* __restrict__ x, const T* __restrict__ y, long long n, double* __restrict__ partials) {
    using P = Pack<T, VEC>;
    const long long nvec = n / VEC, stride = (long long)gridDim.x * blockDim.x;
    double acc = 0.0;
    const P* xp = reinterpret_cast<const P*>(x);
    const P* yp = reinterpret_cast<const P*>(y);
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += stride) {
        const P a = xp[q], b = yp[q];
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc += (double)a.v[e] * (double)b.v[e];
    }
    for (long long idx = nvec * VEC + (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride)
        acc += (double)x[idx] * (double)y[idx];
    acc = block_sum(acc);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

template <typename T, int VEC>
__global__ void axpby_vec_kernel(T* __restrict__ out, double a, const T* __restrict__ x, double b, const T* __restrict__ y,
                                 long long n) {
    using P = Pack<T, VEC>;
    const long long nvec = n / VEC, stride = (long long)gridDim.x * blockDim.x;
    const P* xp = reinterpret_cast<const P*>(x);
    const P* yp = reinterpret_cast<const P*>(y);
    P* op = reinterpret_cast<P*>(out);
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += stride) {
        const P u = xp[q];
        P r;
        if (y != nullptr) {
            const P w = yp[q];
#pragma unroll
            for (int e = 0; e < VEC; ++e) r.v[e] = (T)(a * (double)u.v[e] + b * (double)w.v[e]);
        } else {
#pragma unroll
            for (int e = 0; e < VEC; ++e) r.v[e] = (T)(a * (double)u.v[e]);
        }
        op[q] = r;
    }
    for (long long idx = nvec * VEC + (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride) {
        double v = a * (double)x[idx];
        if (y != nullptr) v += b * (double)y[idx];
        out[idx] = (T)v;
    }
}

template <typename T, int VEC>
__global__ void lincomb_vec_kernel(T* __restrict__ out, const __grid_constant__ LincombArgs<T> args, long long n) {
    using P = Pack<T, VEC>;
    const long long nvec = n / VEC, stride = (long long)gridDim.x * blockDim.x;
    P* op = reinterpret_cast<P*>(out);
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += stride) {
        double acc[VEC];
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[e] = 0.0;
        for (int k = 0; k < args.n; ++k) {
            const P u = reinterpret_cast<const P*>(args.v[k])[q];
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[e] += args.c[k] * (double)u.v[e];
        }
        P r;
#pragma unroll
        for (int e = 0; e < VEC; ++e) r.v[e] = (T)acc[e];
        op[q] = r;
    }
    for (long long idx = nvec * VEC + (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride) {
        double v = 0.0;
        for (int k = 0; k < args.n; ++k) v += args.c[k] * (double)args.v[k][idx];
        out[idx] = (T)v;
    }
}

// m dot products <x, y_k> in one pass: x is read once, partials[k * gridDim.x + block] (DIIS: the new row of the
// Gram matrix, utils.py:108-112, in one launch and one host synchronisation instead of m)
template <typename T, int VEC>
__global__ void dot_many_kernel(const T* __restrict__ x, const __grid_constant__ LincombArgs<T> ys, long long n,
                                double* __restrict__ partials) {
    using P = Pack<T, VEC>;
    const long long nvec = n / VEC, stride = (long long)gridDim.x * blockDim.x;
    double acc[MAX_LINCOMB];
#pragma unroll
    for (int k = 0; k < MAX_LINCOMB; ++k) acc[k] = 0.0;
    const P* xp = reinterpret_cast<const P*>(x);
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += stride) {
        const P a = xp[q];
#pragma unroll
        for (int k = 0; k < MAX_LINCOMB; ++k)
            if (k < ys.n) {
                const P b = reinterpret_cast<const P*>(ys.v[k])[q];
#pragma unroll
                for (int e = 0; e < VEC; ++e) acc[k] += (double)a.v[e] * (double)b.v[e];
            }
    }
    for (long long idx = nvec * VEC + (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride)
#pragma unroll
        for (int k = 0; k < MAX_LINCOMB; ++k)
            if (k < ys.n) acc[k] += (double)x[idx] * (double)ys.v[k][idx];
#pragma unroll
    for (int k = 0; k < MAX_LINCOMB; ++k)
        if (k < ys.n) {          // uniform over the block
            const double r = block_sum(acc[k]);
            if (threadIdx.x == 0) partials[(size_t)k * gridDim.x + blockIdx.x] = r;
        }
}

__host__ bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---- fused amplitude update + increment norm + energy ----------------------------------------------
// One CTA per 32x32 (a,b) tile of one (i,j) block, for the blocks ij in [ij_lo, ij_hi) (a rank's share of
// the update when several GPUs split it; the whole range on one GPU):
//     r      = r2[ijab] (+ r2[jiba] when `symmetrize`: the partner tile is read coalesced and transposed
//              through shared memory, ccwfn.py:621)
//     t2    += r / D_ijab,  D from the Fock diagonal                                  (ccwfn.py:288)
//     rms2  += (r / D)^2                                                              (ccwfn.py:289-290)
//     E     += (t2_new[ijab] + t1[ia] t1[jb]) (2 <ij|ab> - <ij|ba>)   when `g` is given (ccwfn.py:717-718)
// t2 is in registers when the energy needs it, so fusing saves one full read of t2 per iteration.  All
// global accesses are two-element vectors along b (16 bytes in FP64), 256-byte runs per tile row.
template <typename T>
__global__ void __launch_bounds__(256) update_energy_kernel(T* __restrict__ t2, const T* __restrict__ r2,
                                                            const T* __restrict__ eps, const T* __restrict__ t1,
                                                            const T* __restrict__ g, int no, int nv, int tiles_per_dim,
                                                            int ij_lo, int symmetrize, double* __restrict__ part_rms,
                                                            double* __restrict__ part_e) {
    __shared__ T tr[32][33];      // r2[ji] tile, [b][a]
    __shared__ T tg[32][33];      // g[ij] tile at (b, a)
    const int ij = ij_lo + blockIdx.y;
    const int i = ij / no, j = ij - i * no;
    const int ta = blockIdx.x / tiles_per_dim, tb = blockIdx.x % tiles_per_dim;
    const long long nvv = (long long)nv * nv;
    const long long Pb = (long long)ij * nvv, Qb = ((long long)j * no + i) * nvv;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;       // 16 x 16, two columns per thread
    const bool vec = (nv & 1) == 0;
    const bool want_e = g != nullptr;
    // transposed inputs: rows b of the tile (tb), columns a (ta)
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int b = tb * 32 + ty + 16 * p, a0 = ta * 32 + 2 * tx;
        T x0 = T(0), x1 = T(0), y0 = T(0), y1 = T(0);
        if (b < nv && a0 < nv) {
            const long long off = (long long)b * nv + a0;
            if (vec) {
                if (symmetrize) { const Pack<T, 2> q = *reinterpret_cast<const Pack<T, 2>*>(r2 + Qb + off); x0 = q.v[0]; x1 = q.v[1]; }
                if (want_e) { const Pack<T, 2> q = *reinterpret_cast<const Pack<T, 2>*>(g + Pb + off); y0 = q.v[0]; y1 = q.v[1]; }
            } else {
                if (symmetrize) { x0 = r2[Qb + off]; if (a0 + 1 < nv) x1 = r2[Qb + off + 1]; }
                if (want_e) { y0 = g[Pb + off]; if (a0 + 1 < nv) y1 = g[Pb + off + 1]; }
            }
        }
        tr[ty + 16 * p][2 * tx] = x0; tr[ty + 16 * p][2 * tx + 1] = x1;
        tg[ty + 16 * p][2 * tx] = y0; tg[ty + 16 * p][2 * tx + 1] = y1;
    }
    // the direct operands are requested BEFORE the barrier, so that one DRAM round trip covers all five streams
    T rrp[2][2], ttp[2][2], ggp[2][2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int a = ta * 32 + ty + 16 * p, b0 = tb * 32 + 2 * tx;
        rrp[p][0] = rrp[p][1] = ttp[p][0] = ttp[p][1] = ggp[p][0] = ggp[p][1] = T(0);
        if (a < nv && b0 < nv) {
            const long long off = (long long)a * nv + b0;
            const bool two = b0 + 1 < nv;
            if (vec) {
                const Pack<T, 2> q = *reinterpret_cast<const Pack<T, 2>*>(r2 + Pb + off); rrp[p][0] = q.v[0]; rrp[p][1] = q.v[1];
                const Pack<T, 2> u = *reinterpret_cast<const Pack<T, 2>*>(t2 + Pb + off); ttp[p][0] = u.v[0]; ttp[p][1] = u.v[1];
                if (want_e) { const Pack<T, 2> w = *reinterpret_cast<const Pack<T, 2>*>(g + Pb + off); ggp[p][0] = w.v[0]; ggp[p][1] = w.v[1]; }
            } else {
                rrp[p][0] = r2[Pb + off]; ttp[p][0] = t2[Pb + off];
                if (want_e) ggp[p][0] = g[Pb + off];
                if (two) { rrp[p][1] = r2[Pb + off + 1]; ttp[p][1] = t2[Pb + off + 1]; if (want_e) ggp[p][1] = g[Pb + off + 1]; }
            }
        }
    }
    __syncthreads();
    const double ei = (double)eps[i] + (double)eps[j];
    double acc_r = 0.0, acc_e = 0.0;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int al = ty + 16 * p, a = ta * 32 + al, b0 = tb * 32 + 2 * tx;
        if (a < nv && b0 < nv) {
            const long long off = (long long)a * nv + b0;
            const bool two = b0 + 1 < nv;
            const T* rr = rrp[p];
            const T* tt = ttp[p];
            const T* gg = ggp[p];
            const double ea = ei - (double)eps[no + a];
            const double t1a = want_e ? (double)t1[(long long)i * nv + a] : 0.0;
            T out[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int b = b0 + e;
                double r = (double)rr[e] + (double)tr[2 * tx + e][al];
                const double d = ea - (double)eps[no + (b < nv ? b : 0)];
                const double inc = (double)(T)(r / d);
                const double tn = (double)tt[e] + inc;
                out[e] = (T)tn;
                if (b < nv) {
                    acc_r += inc * inc;
                    if (want_e)
                        acc_e += ((double)out[e] + t1a * (double)t1[(long long)j * nv + b]) *
                                 (2.0 * (double)gg[e] - (double)tg[2 * tx + e][al]);
                }
            }
            if (vec) {
                Pack<T, 2> q; q.v[0] = out[0]; q.v[1] = out[1];
                *reinterpret_cast<Pack<T, 2>*>(t2 + Pb + off) = q;
            } else {
                t2[Pb + off] = out[0];
                if (two) t2[Pb + off + 1] = out[1];
            }
        }
    }
    const long long slot = (long long)blockIdx.y * gridDim.x + blockIdx.x;
    acc_r = block_sum(acc_r);
    if (threadIdx.x == 0) part_rms[slot] = acc_r;
    acc_e = block_sum(acc_e);
    if (threadIdx.x == 0) part_e[slot] = acc_e;
}

// singles: t1 += r1/D, increment norm and 2 sum F_ia t_ia of the NEW t1 (one small block grid)
template <typename T>
__global__ void update_t1_energy_kernel(T* __restrict__ t1, const T* __restrict__ r1, const T* __restrict__ eps,
                                        const T* __restrict__ fov, long long ld_fov, int no, int nv,
                                        double* __restrict__ part_rms, double* __restrict__ part_e) {
    const long long total = (long long)no * nv;
    double acc = 0.0, ace = 0.0;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(idx / nv), a = (int)(idx - (long long)i * nv);
        const double d = (double)eps[i] - (double)eps[no + a];
        const double inc = (double)(T)((double)r1[idx] / d);
        const T tn = (T)((double)t1[idx] + inc);
        t1[idx] = tn;
        acc += inc * inc;
        if (fov != nullptr) ace += 2.0 * (double)fov[(long long)i * ld_fov + a] * (double)tn;
    }
    acc = block_sum(acc);
    if (threadIdx.x == 0) part_rms[blockIdx.x] = acc;
    ace = block_sum(ace);
    if (threadIdx.x == 0) part_e[blockIdx.x] = ace;
}

// ---- fused per-iteration preparation of the amplitude-dependent operands ---------------------------
// One pass over t2 (each tile and its (a,b)-transposed partner) writes every o^2 v^2 operand the planned
// contractions read: the three tau variants and 2 t2 - t2^T (build_tau, ccwfn.py:408-410, :554), t1 x t1
// (CC2), and the plain-matrix layouts of the ring-term operands (engine.STATEMENTS_SP):
//     t2sP[i,a,m,e] = (2 t2 - t2^T)[i,m,a,e]    t2P[i,a,m,e] = t2[i,m,a,e]    t2Q[j,a,m,e] = t2[m,j,a,e]
//     tauWPn[j,B,n,f] = -(t2/2 + t1 t1)[j,n,f,B]   t2Rh[j,B,n,f] = t2[n,j,f,B] / 2     (B in [vlo, vlo+vl))
// 11 tensor passes instead of the 22 of four pair_combine + five permute4 launches.
struct PrepareOut {
    void *tau, *tauh, *tauW, *t2s, *t11, *t2sP, *t2P, *t2Q, *tauWPn, *t2Rh;
};

template <typename T>
__global__ void __launch_bounds__(256) prepare_kernel(const T* __restrict__ t1, const T* __restrict__ t2, int no, int nv,
                                                      int vlo, int vl, int tiles_per_dim, const __grid_constant__ PrepareOut o) {
    __shared__ T tile[32][33];
    __shared__ T tile2[32][33];
    const int ij = blockIdx.y, i = ij / no, j = ij - i * no;
    const int ta = blockIdx.x / tiles_per_dim, tb = blockIdx.x % tiles_per_dim;
    const long long nvv = (long long)nv * nv, Pb = (long long)ij * nvv;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const bool vec = (nv & 1) == 0;
    const bool need_t = o.t2s != nullptr || o.t2sP != nullptr;
    if (need_t) {       // tile[b_local][a_local] = t2[ij][b][a]
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const int b = tb * 32 + ty + 16 * p, a0 = ta * 32 + 2 * tx;
            T x0 = T(0), x1 = T(0);
            if (b < nv && a0 < nv) {
                const long long off = Pb + (long long)b * nv + a0;
                if (vec) { const Pack<T, 2> q = *reinterpret_cast<const Pack<T, 2>*>(t2 + off); x0 = q.v[0]; x1 = q.v[1]; }
                else { x0 = t2[off]; if (a0 + 1 < nv) x1 = t2[off + 1]; }
            }
            tile[ty + 16 * p][2 * tx] = x0; tile[ty + 16 * p][2 * tx + 1] = x1;
        }
        __syncthreads();
    }
    auto put2 = [&](void* base, long long off, double v0, double v1, bool two) {
        T* q = static_cast<T*>(base) + off;
        if (vec) { Pack<T, 2> w; w.v[0] = (T)v0; w.v[1] = (T)v1; *reinterpret_cast<Pack<T, 2>*>(q) = w; }
        else { q[0] = (T)v0; if (two) q[1] = (T)v1; }
    };
    double xw[2][2], xx[2][2];     // tauW and t2 of this thread's elements, kept for the transposed writes
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int al = ty + 16 * p, a = ta * 32 + al, b0 = tb * 32 + 2 * tx;
        xw[p][0] = xw[p][1] = xx[p][0] = xx[p][1] = 0.0;
        if (a < nv && b0 < nv) {
            const bool two = b0 + 1 < nv;
            const long long off = Pb + (long long)a * nv + b0;
            double x[2] = {0.0, 0.0};
            if (vec) { const Pack<T, 2> q = *reinterpret_cast<const Pack<T, 2>*>(t2 + off); x[0] = (double)q.v[0]; x[1] = (double)q.v[1]; }
            else { x[0] = (double)t2[off]; if (two) x[1] = (double)t2[off + 1]; }
            const double ua = (double)t1[(long long)i * nv + a];
            const double pr[2] = {ua * (double)t1[(long long)j * nv + b0], two ? ua * (double)t1[(long long)j * nv + b0 + 1] : 0.0};
            const double s[2] = {2.0 * x[0] - (need_t ? (double)tile[2 * tx][al] : 0.0),
                                 2.0 * x[1] - (need_t ? (double)tile[2 * tx + 1][al] : 0.0)};
            if (o.tau) put2(o.tau, off, x[0] + pr[0], x[1] + pr[1], two);
            if (o.tauh) put2(o.tauh, off, x[0] + 0.5 * pr[0], x[1] + 0.5 * pr[1], two);
            if (o.tauW) put2(o.tauW, off, 0.5 * x[0] + pr[0], 0.5 * x[1] + pr[1], two);
            if (o.t2s) put2(o.t2s, off, s[0], s[1], two);
            if (o.t11) put2(o.t11, off, pr[0], pr[1], two);
            // row-permuted copies: the run along b stays contiguous
            const long long offP = (((long long)i * nv + a) * no + j) * nv + b0;       // [i][a][j][b]
            if (o.t2sP) put2(o.t2sP, offP, s[0], s[1], two);
            if (o.t2P) put2(o.t2P, offP, x[0], x[1], two);
            if (o.t2Q) put2(o.t2Q, (((long long)j * nv + a) * no + i) * nv + b0, x[0], x[1], two);   // [j][a][i][b]
            xw[p][0] = 0.5 * x[0] + pr[0]; xw[p][1] = 0.5 * x[1] + pr[1];
            xx[p][0] = x[0]; xx[p][1] = x[1];
        }
    }
    if (o.tauWPn == nullptr && o.t2Rh == nullptr) return;
    if (tb * 32 >= vlo + vl || tb * 32 + 32 <= vlo) return;      // no column B of this tile is in the slice
    __syncthreads();
#pragma unroll
    for (int p = 0; p < 2; ++p) {       // tile[a_local][b_local] = tauW, tile2 = t2
        tile[ty + 16 * p][2 * tx] = (T)xw[p][0]; tile[ty + 16 * p][2 * tx + 1] = (T)xw[p][1];
        tile2[ty + 16 * p][2 * tx] = (T)xx[p][0]; tile2[ty + 16 * p][2 * tx + 1] = (T)xx[p][1];
    }
    __syncthreads();
#pragma unroll
    for (int p = 0; p < 2; ++p) {       // rows b (slice only), two columns a per thread
        const int bl = ty + 16 * p, b = tb * 32 + bl, a0 = ta * 32 + 2 * tx;
        if (b < nv && b >= vlo && b < vlo + vl && a0 < nv) {
            const bool two = a0 + 1 < nv;
            // tauWPn[i][B][j][a] = -tauW[i,j,a,B];  t2Rh[j][B][i][a] = t2[i,j,a,B] / 2
            if (o.tauWPn) put2(o.tauWPn, (((long long)i * vl + (b - vlo)) * no + j) * nv + a0,
                               -(double)tile[2 * tx][bl], -(double)tile[2 * tx + 1][bl], two);
            if (o.t2Rh) put2(o.t2Rh, (((long long)j * vl + (b - vlo)) * no + i) * nv + a0,
                             0.5 * (double)tile2[2 * tx][bl], 0.5 * (double)tile2[2 * tx + 1][bl], two);
        }
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
    const int tpd = (int)((nv + PT - 1) / PT);
    dim3 grid(tpd * tpd, (unsigned)(no * no));
    pair_combine_kernel<T><<<grid, 256, 0, ctx->stream>>>((T*)out, (const T*)X, c0, c1, (const T*)u, (const T*)w, c2,
                                                          (int)no, (int)nv, tpd);
    PCCB_CHECK_LAUNCH(ctx, "pair_combine_kernel");
    return PCCB_OK;
}

// t1 / t2 update (+ optional fused energy) for the (i,j) blocks [ij_lo, ij_hi); out_host[0] = squared
// increment norm, out_host[1] = energy of the new amplitudes over those blocks (0 without `g`).  `with_t1`:
// also update t1 and count its norm / singles energy (one rank does when the blocks are split over ranks).
template <typename T>
int update_energy_t(pccb_ctx* ctx, void* t1, void* t2, const void* r1, const void* r2, const void* eps, const void* fov,
                    int64_t ld_fov, const void* g, int64_t no, int64_t nv, int symmetrize, int64_t ij_lo, int64_t ij_hi,
                    int with_t1, int count_t1, double* out_host) {
    const int tpd = (int)((nv + 31) / 32);
    const long long nblk = ij_hi - ij_lo;
    const long long g2 = nblk * tpd * tpd;
    const int g1 = with_t1 ? grid_for(no * nv, RED_THREADS, 64) : 0;
    const long long n = g2 + g1;
    int rc = pccb_ensure_partials(ctx, (size_t)(2 * n) + 64);
    if (rc) return rc;
    double* prms = ctx->partial_dev;
    double* pe = ctx->partial_dev + n;
    if (with_t1) {
        // t1 first: the doubles energy reads the NEW singles
        update_t1_energy_kernel<T><<<g1, RED_THREADS, 0, ctx->stream>>>((T*)t1, (const T*)r1, (const T*)eps,
                                                                        g ? (const T*)fov : nullptr, ld_fov, (int)no, (int)nv,
                                                                        prms + g2, pe + g2);
        PCCB_CHECK_LAUNCH(ctx, "update_t1_energy_kernel");
        if (!count_t1) PCCB_CUDA(ctx, cudaMemsetAsync(prms + g2, 0, sizeof(double) * g1, ctx->stream));
        if (!count_t1) PCCB_CUDA(ctx, cudaMemsetAsync(pe + g2, 0, sizeof(double) * g1, ctx->stream));
    }
    if (nblk > 0) {
        dim3 grid((unsigned)(tpd * tpd), (unsigned)nblk);
        update_energy_kernel<T><<<grid, 256, 0, ctx->stream>>>((T*)t2, (const T*)r2, (const T*)eps, (const T*)t1,
                                                               (const T*)g, (int)no, (int)nv, tpd, (int)ij_lo, symmetrize,
                                                               prms, pe);
        PCCB_CHECK_LAUNCH(ctx, "update_energy_kernel");
    }
    if (n == 0) { out_host[0] = out_host[1] = 0.0; return PCCB_OK; }
    return finish_reduction(ctx, (int)n, 2, out_host);
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
    const int tpd = (int)((nv + PT - 1) / PT);
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
    PCCB_REQUIRE(ctx, dtype == PCCB_F64 || dtype == PCCB_F32, "pccb_permute4: unknown dtype");
    // the input's contiguous index: smallest stride among the indices of extent > 1
    int q = -1;
    for (int d = 0; d < 4; ++d)
        if (shape[d] > 1 && in_strides[d] > 0 && (q < 0 || in_strides[d] < in_strides[q])) q = d;
    static const bool fast = [] { const char* e = getenv("PCCB_PERMUTE_FAST"); return e ? atoi(e) != 0 : true; }();
    const long long rows = shape[0] * shape[1] * shape[2];
    if (fast && q == 3 && rows < (1ll << 31) && shape[3] < (1ll << 31)) {
        const int gr = (int)((rows + 7) / 8 < 148 * 32 ? (rows + 7) / 8 : 148 * 32);
        if (dtype == PCCB_F64)
            permute4_rows_kernel<double><<<gr, 256, 0, ctx->stream>>>((double*)out_dev, (const double*)in_dev, (unsigned)shape[1],
                                                                      (unsigned)shape[2], (unsigned)shape[3], in_strides[0], in_strides[1],
                                                                      in_strides[2], in_strides[3], (unsigned)rows, alpha, beta);
        else
            permute4_rows_kernel<float><<<gr, 256, 0, ctx->stream>>>((float*)out_dev, (const float*)in_dev, (unsigned)shape[1],
                                                                     (unsigned)shape[2], (unsigned)shape[3], in_strides[0], in_strides[1],
                                                                     in_strides[2], in_strides[3], (unsigned)rows, alpha, beta);
        PCCB_CHECK_LAUNCH(ctx, "permute4_rows_kernel");
        return PCCB_OK;
    }
    if (fast && q >= 0 && q < 3) {
        int uw[2], k = 0;
        for (int d = 0; d < 3; ++d)
            if (d != q) uw[k++] = d;
        const long long ostr[4] = {shape[1] * shape[2] * shape[3], shape[2] * shape[3], shape[3], 1};
        const long long tq = (shape[q] + 63) / 64, t3 = (shape[3] + 63) / 64;
        if (shape[uw[0]] <= 65535 && shape[uw[1]] <= 65535 && tq * t3 < (1ll << 31)) {
            const dim3 grid((unsigned)(tq * t3), (unsigned)shape[uw[0]], (unsigned)shape[uw[1]]);
            if (dtype == PCCB_F64)
                permute4_tiles_kernel<double><<<grid, 256, 0, ctx->stream>>>(
                    (double*)out_dev, (const double*)in_dev, (unsigned)shape[q], (unsigned)shape[3], (unsigned)tq, in_strides[q],
                    in_strides[3], in_strides[uw[0]], in_strides[uw[1]], ostr[q], ostr[uw[0]], ostr[uw[1]], alpha, beta);
            else
                permute4_tiles_kernel<float><<<grid, 256, 0, ctx->stream>>>(
                    (float*)out_dev, (const float*)in_dev, (unsigned)shape[q], (unsigned)shape[3], (unsigned)tq, in_strides[q],
                    in_strides[3], in_strides[uw[0]], in_strides[uw[1]], ostr[q], ostr[uw[0]], ostr[uw[1]], alpha, beta);
            PCCB_CHECK_LAUNCH(ctx, "permute4_tiles_kernel");
            return PCCB_OK;
        }
    }
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
    PCCB_REQUIRE(ctx, t1_dev && t2_dev && r1_dev && r2_dev && eps_dev && rms2_host && no > 0 && nv > 0 && no * no <= 65535,
                 "pccb_update_amplitudes: bad arguments");
    double out[2] = {0.0, 0.0};
    int rc;
    if (dtype == PCCB_F64)
        rc = update_energy_t<double>(ctx, t1_dev, t2_dev, r1_dev, r2_dev, eps_dev, nullptr, 0, nullptr, no, nv, symmetrize, 0,
                                     no * no, 1, 1, out);
    else if (dtype == PCCB_F32)
        rc = update_energy_t<float>(ctx, t1_dev, t2_dev, r1_dev, r2_dev, eps_dev, nullptr, 0, nullptr, no, nv, symmetrize, 0,
                                    no * no, 1, 1, out);
    else
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    *rms2_host = out[0];
    return rc;
}

extern "C" int pccb_update_energy(pccb_ctx* ctx, int dtype, void* t1_dev, void* t2_dev, const void* r1_dev,
                                  const void* r2_dev, const void* eps_dev, const void* fov_dev, int64_t ld_fov,
                                  const void* oovv_dev, int64_t no, int64_t nv, int symmetrize, int64_t ij_lo,
                                  int64_t ij_hi, int count_t1, double* out_host) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, t1_dev && t2_dev && r1_dev && r2_dev && eps_dev && fov_dev && oovv_dev && out_host && no > 0 && nv > 0 &&
                          no * no <= 65535 && ij_lo >= 0 && ij_lo <= ij_hi && ij_hi <= no * no,
                 "pccb_update_energy: bad arguments");
    DISPATCH_DTYPE(ctx, dtype,
                   update_energy_t<double>(ctx, t1_dev, t2_dev, r1_dev, r2_dev, eps_dev, fov_dev, ld_fov, oovv_dev, no, nv,
                                           symmetrize, ij_lo, ij_hi, 1, count_t1, out_host),
                   update_energy_t<float>(ctx, t1_dev, t2_dev, r1_dev, r2_dev, eps_dev, fov_dev, ld_fov, oovv_dev, no, nv,
                                          symmetrize, ij_lo, ij_hi, 1, count_t1, out_host));
}

extern "C" int pccb_prepare_amplitudes(pccb_ctx* ctx, int dtype, const void* t1_dev, const void* t2_dev, int64_t no,
                                       int64_t nv, int64_t vlo, int64_t vl, const pccb_prepare_outputs* out) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, t1_dev && t2_dev && out && no > 0 && nv > 0 && no * no <= 65535 && vlo >= 0 && vl >= 0 && vlo + vl <= nv,
                 "pccb_prepare_amplitudes: bad arguments");
    PrepareOut o;
    o.tau = out->tau; o.tauh = out->tauh; o.tauW = out->tauW; o.t2s = out->t2s; o.t11 = out->t11;
    o.t2sP = out->t2sP; o.t2P = out->t2P; o.t2Q = out->t2Q; o.tauWPn = out->tauWPn; o.t2Rh = out->t2Rh;
    const int tpd = (int)((nv + 31) / 32);
    dim3 grid((unsigned)(tpd * tpd), (unsigned)(no * no));
    if (dtype == PCCB_F64)
        prepare_kernel<double><<<grid, 256, 0, ctx->stream>>>((const double*)t1_dev, (const double*)t2_dev, (int)no, (int)nv,
                                                              (int)vlo, (int)vl, tpd, o);
    else if (dtype == PCCB_F32)
        prepare_kernel<float><<<grid, 256, 0, ctx->stream>>>((const float*)t1_dev, (const float*)t2_dev, (int)no, (int)nv,
                                                             (int)vlo, (int)vl, tpd, o);
    else
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    PCCB_CHECK_LAUNCH(ctx, "prepare_kernel");
    return PCCB_OK;
}

extern "C" int pccb_cc_energy(pccb_ctx* ctx, int dtype, const void* fov_dev, int64_t ld_fov, const void* t1_dev,
                              const void* t2_dev, const void* oovv_dev, int64_t no, int64_t nv, double* e_host) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, fov_dev && t1_dev && t2_dev && oovv_dev && e_host && no > 0 && nv > 0 && no * no <= 65535,
                 "pccb_cc_energy: bad arguments");
    DISPATCH_DTYPE(ctx, dtype, energy_t<double>(ctx, fov_dev, ld_fov, t1_dev, t2_dev, oovv_dev, no, nv, e_host),
                   energy_t<float>(ctx, fov_dev, ld_fov, t1_dev, t2_dev, oovv_dev, no, nv, e_host));
}

template <typename T>
int dot_t(pccb_ctx* ctx, const void* x, const void* y, int64_t n, double* out_host) {
    constexpr int V = 16 / sizeof(T);
    const int g = grid_for((n + V - 1) / V, RED_THREADS, MAX_RED_BLOCKS);
    int rc = pccb_ensure_partials(ctx, (size_t)g + 64);
    if (rc) return rc;
    if (aligned16(x) && aligned16(y))
        dot_vec_kernel<T, V><<<g, RED_THREADS, 0, ctx->stream>>>((const T*)x, (const T*)y, n, ctx->partial_dev);
    else
        dot_vec_kernel<T, 1><<<g, RED_THREADS, 0, ctx->stream>>>((const T*)x, (const T*)y, n, ctx->partial_dev);
    PCCB_CHECK_LAUNCH(ctx, "dot_kernel");
    return finish_reduction(ctx, g, 1, out_host);
}

template <typename T>
int axpby_t(pccb_ctx* ctx, void* out, double a, const void* x, double b, const void* y, int64_t n) {
    constexpr int V = 16 / sizeof(T);
    const int g = grid_for((n + V - 1) / V, 256, 148 * 8);
    if (aligned16(out) && aligned16(x) && aligned16(y))
        axpby_vec_kernel<T, V><<<g, 256, 0, ctx->stream>>>((T*)out, a, (const T*)x, b, (const T*)y, n);
    else
        axpby_vec_kernel<T, 1><<<g, 256, 0, ctx->stream>>>((T*)out, a, (const T*)x, b, (const T*)y, n);
    PCCB_CHECK_LAUNCH(ctx, "axpby_kernel");
    return PCCB_OK;
}

template <typename T>
int lincomb_t(pccb_ctx* ctx, void* out, const void* const* vecs, const double* coef, int nvec, int64_t n) {
    constexpr int V = 16 / sizeof(T);
    LincombArgs<T> a;
    a.n = nvec;
    bool al = aligned16(out);
    for (int k = 0; k < nvec; ++k) { a.v[k] = (const T*)vecs[k]; a.c[k] = coef[k]; al = al && aligned16(vecs[k]); }
    const int g = grid_for((n + V - 1) / V, 256, 148 * 8);
    if (al) lincomb_vec_kernel<T, V><<<g, 256, 0, ctx->stream>>>((T*)out, a, n);
    else lincomb_vec_kernel<T, 1><<<g, 256, 0, ctx->stream>>>((T*)out, a, n);
    PCCB_CHECK_LAUNCH(ctx, "lincomb_kernel");
    return PCCB_OK;
}

template <typename T>
int dot_many_t(pccb_ctx* ctx, const void* x, const void* const* ys, int m, int64_t n, double* out_host) {
    constexpr int V = 16 / sizeof(T);
    const int g = grid_for((n + V - 1) / V, RED_THREADS, 148 * 4);
    int rc = pccb_ensure_partials(ctx, (size_t)g * m + 64);
    if (rc) return rc;
    LincombArgs<T> a;
    a.n = m;
    bool al = aligned16(x);
    for (int k = 0; k < m; ++k) { a.v[k] = (const T*)ys[k]; a.c[k] = 0.0; al = al && aligned16(ys[k]); }
    if (al) dot_many_kernel<T, V><<<g, RED_THREADS, 0, ctx->stream>>>((const T*)x, a, n, ctx->partial_dev);
    else dot_many_kernel<T, 1><<<g, RED_THREADS, 0, ctx->stream>>>((const T*)x, a, n, ctx->partial_dev);
    PCCB_CHECK_LAUNCH(ctx, "dot_many_kernel");
    return finish_reduction(ctx, g, m, out_host);
}

extern "C" int pccb_dot_many(pccb_ctx* ctx, int dtype, const void* x_dev, const void* const* ys_dev_hostarray, int m,
                             int64_t n, double* out_host) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, x_dev && ys_dev_hostarray && out_host && m >= 1 && m <= MAX_LINCOMB && n > 0,
                 "pccb_dot_many: bad arguments (m=%d)", m);
    DISPATCH_DTYPE(ctx, dtype, dot_many_t<double>(ctx, x_dev, ys_dev_hostarray, m, n, out_host),
                   dot_many_t<float>(ctx, x_dev, ys_dev_hostarray, m, n, out_host));
}

extern "C" int pccb_dot(pccb_ctx* ctx, int dtype, const void* x_dev, const void* y_dev, int64_t n, double* out_host) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, x_dev && y_dev && out_host && n > 0, "pccb_dot: bad arguments");
    DISPATCH_DTYPE(ctx, dtype, dot_t<double>(ctx, x_dev, y_dev, n, out_host), dot_t<float>(ctx, x_dev, y_dev, n, out_host));
}

extern "C" int pccb_axpby(pccb_ctx* ctx, int dtype, void* out_dev, double a, const void* x_dev, double b,
                          const void* y_dev, int64_t n) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, out_dev && x_dev && n > 0, "pccb_axpby: bad arguments");
    DISPATCH_DTYPE(ctx, dtype, axpby_t<double>(ctx, out_dev, a, x_dev, b, y_dev, n),
                   axpby_t<float>(ctx, out_dev, a, x_dev, b, y_dev, n));
}

extern "C" int pccb_lincomb(pccb_ctx* ctx, int dtype, void* out_dev, const void* const* vecs, const double* coef,
                            int nvec, int64_t n) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, out_dev && vecs && coef && nvec >= 1 && nvec <= MAX_LINCOMB && n > 0,
                 "pccb_lincomb: bad arguments (nvec=%d)", nvec);
    DISPATCH_DTYPE(ctx, dtype, lincomb_t<double>(ctx, out_dev, vecs, coef, nvec, n),
                   lincomb_t<float>(ctx, out_dev, vecs, coef, nvec, n));
}

// ---- pccb_cc_energy_static: the energy reduction with NO allocation anywhere --------------------------
// (No counterpart in the reference; added for callers that cannot
//  call cudaMalloc: CUDA-graph capture, tight serving loops, harnesses that forbid it.)
// pccb_cc_energy keeps its per-CTA partial sums in ctx->partial_dev, which pccb_ensure_partials
// cudaMalloc's / cudaFree's on demand, and returns through a pinned host slot after a stream
// synchronise.  This variant runs the SAME two kernels (energy_kernel, final_sum_kernel) but
//   * the partial sums live in the file-scope array below: PCCB_STATIC_SCRATCH_DOUBLES doubles = 1 MiB,
//     held by every GPU that loads the library until the process exits;
//   * the result goes to a caller-supplied DEVICE double: no host copy, no synchronise.
// In host code `g_static_scratch` is a host-side stand-in, not a device address: its device address is
// looked up once per device with cudaGetSymbolAddress (that device current) and that pointer is what
// the kernels receive.  Every call shares the array and it keeps the previous call's contents; a call
// overwrites every slot it later sums (one per CTA), so nothing needs clearing, but two calls that
// overlap on the GPU (different streams, or concurrent graph branches) race on it.
constexpr size_t PCCB_STATIC_SCRATCH_DOUBLES = size_t(1) << 17;
static __device__ __align__(256) double g_static_scratch[PCCB_STATIC_SCRATCH_DOUBLES];

namespace {

constexpr int MAX_STATIC_DEVICES = 64;
double* g_static_scratch_ptr[MAX_STATIC_DEVICES] = {};     // device address of g_static_scratch, per device

// set-up, once per device: the device address of the scratch array (never an allocation)
int static_scratch_for(pccb_ctx* ctx, double** out) {
    const int dev = ctx->device;
    PCCB_REQUIRE(ctx, dev >= 0 && dev < MAX_STATIC_DEVICES, "pccb_cc_energy_static: device index %d out of range", dev);
    if (!g_static_scratch_ptr[dev]) {
        int cur = -1;
        PCCB_CUDA(ctx, cudaGetDevice(&cur));
        if (cur != dev) PCCB_CUDA(ctx, cudaSetDevice(dev));
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (cur != dev && cur >= 0) cudaSetDevice(cur);
        if (e != cudaSuccess || !p)
            PCCB_FAIL(ctx, PCCB_ERR_CUDA, "cudaGetSymbolAddress(g_static_scratch) failed on device %d: %s", dev,
                      cudaGetErrorString(e));
        g_static_scratch_ptr[dev] = static_cast<double*>(p);
    }
    *out = g_static_scratch_ptr[dev];
    return PCCB_OK;
}

template <typename T>
int energy_static_t(pccb_ctx* ctx, const void* fov, int64_t ld_fov, const void* t1, const void* t2, const void* g,
                    int64_t no, int64_t nv, double* scratch, double* e_dev) {
    const int tpd = (int)((nv + 31) / 32);
    dim3 grid(tpd * tpd, (unsigned)(no * no));
    const size_t nb = (size_t)grid.x * grid.y;
    energy_kernel<T><<<grid, 256, 0, ctx->stream>>>((const T*)fov, ld_fov, (const T*)t1, (const T*)t2, (const T*)g,
                                                    (int)no, (int)nv, tpd, scratch);
    PCCB_CHECK_LAUNCH(ctx, "energy_kernel");
    final_sum_kernel<<<1, RED_THREADS, 0, ctx->stream>>>(scratch, (int)nb, 1, e_dev);
    PCCB_CHECK_LAUNCH(ctx, "final_sum_kernel");
    return PCCB_OK;
}

}  // namespace

extern "C" int64_t pccb_cc_energy_static_max_partials(void) { return (int64_t)PCCB_STATIC_SCRATCH_DOUBLES; }

extern "C" int pccb_cc_energy_static(pccb_ctx* ctx, int dtype, const void* fov_dev, int64_t ld_fov, const void* t1_dev,
                                     const void* t2_dev, const void* oovv_dev, int64_t no, int64_t nv, double* e_dev) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, fov_dev && t1_dev && t2_dev && oovv_dev && e_dev && no > 0 && nv > 0 && no * no <= 65535,
                 "pccb_cc_energy_static: bad arguments");
    const int64_t tpd = (nv + 31) / 32;
    const int64_t need = tpd * tpd * no * no;
    PCCB_REQUIRE(ctx, need <= (int64_t)PCCB_STATIC_SCRATCH_DOUBLES,
                 "pccb_cc_energy_static: no=%lld nv=%lld needs %lld partial sums, the static scratch holds %lld "
                 "(ceil(nv/32)^2 * no^2 must not exceed it); use pccb_cc_energy",
                 (long long)no, (long long)nv, (long long)need, (long long)PCCB_STATIC_SCRATCH_DOUBLES);
    double* scratch = nullptr;
    int rc = static_scratch_for(ctx, &scratch);
    if (rc) return rc;
    DISPATCH_DTYPE(ctx, dtype,
                   energy_static_t<double>(ctx, fov_dev, ld_fov, t1_dev, t2_dev, oovv_dev, no, nv, scratch, e_dev),
                   energy_static_t<float>(ctx, fov_dev, ld_fov, t1_dev, t2_dev, oovv_dev, no, nv, scratch, e_dev));
}
