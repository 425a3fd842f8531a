// pccb_contract: fused, strided, batched FP64 tensor contraction for sm_100a.
//
// Replaces opt_einsum.contract behind cc_contract (pycc/utils.py:162-187) for every pairwise
// contraction of the CCSD residual (pycc/ccwfn.py:413-622) and of the (T) W intermediate
// (pycc/cctriples.py:10-22), *including* the operand/result permutations that opt_einsum
// performs as separate transpose copies.
//
// Hardware mapping (measured on this pool's B200, profiles/r01_peaks.json):
//   * FP64 has no tcgen05/TMEM path on Blackwell; the FP64 tensor instruction is the
//     warp-synchronous DMMA.8x8x4 (PTX mma.sync.m8n8k4.f64), one issue per 16 cycles per SM
//     sub-partition = 37.0 TFLOP/s per GPU; cuBLAS DGEMM reaches 35.5.
//   * CTA tile 128x128x16, 8 warps as 2(M) x 4(N), warp tile 64x32: 32 DMMA per 12 LDS.64, so
//     the tensor pipe, not shared memory or issue slots, is the limiter.
//   * operand tiles are staged with cp.async (LDGSTS, zero-fill for edges) through a 4-stage
//     shared-memory ring; because a DMMA occupies the pipe for 16 cycles the per-element
//     address arithmetic of compound (two-level) einsum indices is free, which is what lets a
//     tile walk e.g. the (i,a) x (m,e) view of t2[i,m,a,e] without any transpose in HBM.
//   * shared tiles keep the global-contiguous direction contiguous ("K-major" [row][k] with
//     pitch 20 doubles, or "MN-major" [k][row] with pitch 132): both pitches are = 4 mod 16
//     eight-byte words, which makes the DMMA fragment loads bank-conflict free.
//   * small outputs (Fae, Fmi, Wmnij ...) use deterministic split-K through a caller-owned
//     workspace, reduced in fixed order together with the alpha/beta/gamma epilogue.
#include <climits>
#include <cstring>

#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16, NTHREADS = 256, STAGES = 4;
constexpr int PK = BK + 4;    // pitch of a K-major tile row   (doubles)
constexpr int PM = BM + 4;    // pitch of an MN-major tile row (doubles); BM == BN
constexpr int STAGE_ELEMS = BM * PK;   // 2560 doubles >= BK*PM = 2112
constexpr long long INVALID_ROW = LLONG_MIN;

struct DSeg {
    const double* A;
    const double* B;
    DIdx am, ak, bn, bk;
    long long a_batch, b_batch;
    double scale;
    unsigned K;
    int kt_begin;          // first global k-tile of this segment
    int a_mn_major, a_vec; // layout of the A tile in shared memory / cp.async width (doubles)
    int b_mn_major, b_vec;
};

struct DParams {
    unsigned M, N;
    int nseg, total_kt, splits, pad_;
    DSeg seg[PCCB_MAX_SEGMENTS];
    double* C;
    DIdx cm, cn;
    long long c_batch;
    const double* D;
    DIdx dm, dn;
    long long d_batch;
    double alpha, beta, gamma;
    double* ws;
};

__device__ __forceinline__ void cp_async16(unsigned dst, const void* src, int bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(bytes));
}
__device__ __forceinline__ void cp_async8(unsigned dst, const void* src, int bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(dst), "l"(src), "r"(bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
        : "+d"(c0), "+d"(c1)
        : "d"(a), "d"(b));
}

// Stage one operand tile (128 rows x 16 k) of the current segment into shared memory.
//   rows[]  : element offset of each tile row (INVALID_ROW beyond the edge), batch offset included
//   K-major : tile[row][k], pitch PK;   MN-major: tile[k][row], pitch PM
template <int VEC>
__device__ __forceinline__ void load_tile_kmajor(double* tile, const double* base, const long long* rows,
                                                 const DIdx& kidx, unsigned k0, unsigned K, int tid) {
    constexpr int CH = BK / VEC;            // chunks per row
    constexpr int RPP = NTHREADS / CH;      // rows per pass
    const int kc = tid % CH, r0 = tid / CH;
    const unsigned k = k0 + kc * VEC;
    int kbytes = 0;
    long long koff = 0;
    if (k < K) {
        koff = didx_off(kidx, k);
        kbytes = (VEC == 2 && k + 1 < K) ? 16 : 8;
    }
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(tile + kc * VEC);
#pragma unroll
    for (int p = 0; p < BM / RPP; ++p) {
        const int row = r0 + p * RPP;
        const long long ro = rows[row];
        const int bytes = (ro == INVALID_ROW) ? 0 : kbytes;
        const double* src = bytes ? base + ro + koff : base;
        const unsigned dst = sbase + row * (PK * 8);
        if (VEC == 2) cp_async16(dst, src, bytes);
        else cp_async8(dst, src, bytes);
    }
}

template <int VEC>
__device__ __forceinline__ void load_tile_mnmajor(double* tile, const double* base, const long long* rows,
                                                  const DIdx& kidx, unsigned k0, unsigned K, int tid) {
    constexpr int CH = BM / VEC;            // chunks per k-row
    constexpr int KPP = NTHREADS / CH;      // k-rows per pass
    const int mc = tid % CH, kr0 = tid / CH;
    const long long ro = rows[mc * VEC];
    int mbytes = 0;
    if (ro != INVALID_ROW) mbytes = (VEC == 2 && rows[mc * VEC + 1] != INVALID_ROW) ? 16 : 8;
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(tile + mc * VEC);
#pragma unroll
    for (int p = 0; p < BK / KPP; ++p) {
        const int kr = kr0 + p * KPP;
        const unsigned k = k0 + kr;
        const int bytes = (k < K) ? mbytes : 0;
        const double* src = bytes ? base + ro + didx_off(kidx, k) : base;
        const unsigned dst = sbase + kr * (PM * 8);
        if (VEC == 2) cp_async16(dst, src, bytes);
        else cp_async8(dst, src, bytes);
    }
}

__global__ void __launch_bounds__(NTHREADS, 1) contract_f64_kernel(const __grid_constant__ DParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sA = reinterpret_cast<double*>(smem_raw);
    double* sB = sA + STAGES * STAGE_ELEMS;
    long long* rowA = reinterpret_cast<long long*>(sB + STAGES * STAGE_ELEMS);
    long long* rowB = rowA + BM;
    int* sSeg = reinterpret_cast<int*>(rowB + BN);     // segment id of the tile held by each stage

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const unsigned m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int batch = blockIdx.z / p.splits, split = blockIdx.z % p.splits;
    const int kt_lo = (int)(((long long)p.total_kt * split) / p.splits);
    const int kt_hi = (int)(((long long)p.total_kt * (split + 1)) / p.splits);
    const int nkt = kt_hi - kt_lo;

    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    int cur_seg = -1;   // segment whose row tables are in shared memory (CTA-uniform)

    auto load_tile = [&](int kt, int st) {
        int s = 0;
#pragma unroll
        for (int q = 1; q < PCCB_MAX_SEGMENTS; ++q)
            if (q < p.nseg && kt >= p.seg[q].kt_begin) s = q;
        const DSeg& sg = p.seg[s];
        if (s != cur_seg) {   // uniform: (re)build the per-row offset tables for this segment
            __syncthreads();
            if (tid < BM) {
                const unsigned m = m0 + tid;
                rowA[tid] = (m < p.M) ? didx_off(sg.am, m) + (long long)batch * sg.a_batch : INVALID_ROW;
            } else {
                const unsigned n = n0 + (tid - BM);
                rowB[tid - BM] = (n < p.N) ? didx_off(sg.bn, n) + (long long)batch * sg.b_batch : INVALID_ROW;
            }
            __syncthreads();
            cur_seg = s;
        }
        const unsigned k0 = (unsigned)(kt - sg.kt_begin) * BK;
        double* ta = sA + st * STAGE_ELEMS;
        double* tb = sB + st * STAGE_ELEMS;
        if (sg.a_mn_major) {
            if (sg.a_vec == 2) load_tile_mnmajor<2>(ta, sg.A, rowA, sg.ak, k0, sg.K, tid);
            else load_tile_mnmajor<1>(ta, sg.A, rowA, sg.ak, k0, sg.K, tid);
        } else {
            if (sg.a_vec == 2) load_tile_kmajor<2>(ta, sg.A, rowA, sg.ak, k0, sg.K, tid);
            else load_tile_kmajor<1>(ta, sg.A, rowA, sg.ak, k0, sg.K, tid);
        }
        if (sg.b_mn_major) {
            if (sg.b_vec == 2) load_tile_mnmajor<2>(tb, sg.B, rowB, sg.bk, k0, sg.K, tid);
            else load_tile_mnmajor<1>(tb, sg.B, rowB, sg.bk, k0, sg.K, tid);
        } else {
            if (sg.b_vec == 2) load_tile_kmajor<2>(tb, sg.B, rowB, sg.bk, k0, sg.K, tid);
            else load_tile_kmajor<1>(tb, sg.B, rowB, sg.bk, k0, sg.K, tid);
        }
        if (tid == 0) sSeg[st] = s;
    };

#pragma unroll 1
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nkt) load_tile(kt_lo + s, s);
        cp_async_commit();
    }

    const int frow = lane >> 2, fk = lane & 3;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        const int nxt = it + STAGES - 1;
        if (nxt < nkt) load_tile(kt_lo + nxt, nxt % STAGES);
        cp_async_commit();

        const int st = it % STAGES;
        const DSeg& sg = p.seg[sSeg[st]];
        const double scale = sg.scale;
        const int a_rs = sg.a_mn_major ? 1 : PK, a_ks = sg.a_mn_major ? PM : 1;
        const int b_rs = sg.b_mn_major ? 1 : PK, b_ks = sg.b_mn_major ? PM : 1;
        const double* ap = sA + st * STAGE_ELEMS + (wm * 64 + frow) * a_rs + fk * a_ks;
        const double* bp = sB + st * STAGE_ELEMS + (wn * 32 + frow) * b_rs + fk * b_ks;
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            double a[8], b[4];
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = ap[i * 8 * a_rs + ks * 4 * a_ks];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bp[j * 8 * b_rs + ks * 4 * b_ks];
            if (scale != 1.0) {
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] *= scale;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    cp_async_wait<0>();

    // ---- epilogue ----------------------------------------------------------------------
    const unsigned colbase = n0 + wn * 32 + 2 * fk;
    if (p.splits > 1) {
        double* ws = p.ws + (size_t)blockIdx.z * p.M * p.N;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const unsigned row = m0 + wm * 64 + i * 8 + frow;
            if (row >= p.M) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const unsigned col = colbase + j * 8 + e;
                    if (col < p.N) ws[(size_t)row * p.N + col] = acc[i][j][e];
                }
        }
        return;
    }
    long long roff[8], droff[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const unsigned row = m0 + wm * 64 + i * 8 + frow;
        roff[i] = (row < p.M) ? didx_off(p.cm, row) + (long long)batch * p.c_batch : -1;
        droff[i] = (row < p.M && p.D) ? didx_off(p.dm, row) + (long long)batch * p.d_batch : 0;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const unsigned col = colbase + j * 8 + e;
            if (col >= p.N) continue;
            const long long co = didx_off(p.cn, col);
            const long long dco = p.D ? didx_off(p.dn, col) : 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (roff[i] < 0) continue;
                double v = p.alpha * acc[i][j][e];
                double* c = p.C + roff[i] + co;
                if (p.beta != 0.0) v += p.beta * (*c);
                if (p.D) v += p.gamma * p.D[droff[i] + dco];
                *c = v;
            }
        }
}

// fixed-order reduction of the split-K partial tiles + epilogue
__global__ void splitk_reduce_kernel(const __grid_constant__ DParams p, long long total) {
    const long long MN = (long long)p.M * p.N;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long b = idx / MN, r = idx - b * MN;
        const unsigned row = (unsigned)(r / p.N), col = (unsigned)(r - (long long)row * p.N);
        const double* w = p.ws + (size_t)b * p.splits * MN + r;
        double s = 0.0;
        for (int k = 0; k < p.splits; ++k) s += w[(size_t)k * MN];
        double v = p.alpha * s;
        double* c = p.C + didx_off(p.cm, row) + didx_off(p.cn, col) + b * p.c_batch;
        if (p.beta != 0.0) v += p.beta * (*c);
        if (p.D) v += p.gamma * p.D[didx_off(p.dm, row) + didx_off(p.dn, col) + b * p.d_batch];
        *c = v;
    }
}

// One thread per output element; the obviously-correct cross-check for the tile kernel
// (selected with pccb_set_contract_impl(ctx, PCCB_IMPL_NAIVE); used by tests only).
__global__ void contract_f64_naive_kernel(const __grid_constant__ DParams p, long long total) {
    const long long MN = (long long)p.M * p.N;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long b = idx / MN, r = idx - b * MN;
        const unsigned row = (unsigned)(r / p.N), col = (unsigned)(r - (long long)row * p.N);
        double s = 0.0;
        for (int q = 0; q < p.nseg; ++q) {
            const DSeg& sg = p.seg[q];
            const double* a = sg.A + didx_off(sg.am, row) + b * sg.a_batch;
            const double* bb = sg.B + didx_off(sg.bn, col) + b * sg.b_batch;
            double t = 0.0;
            for (unsigned k = 0; k < sg.K; ++k) t += a[didx_off(sg.ak, k)] * bb[didx_off(sg.bk, k)];
            s += sg.scale * t;
        }
        double v = p.alpha * s;
        double* c = p.C + didx_off(p.cm, row) + didx_off(p.cn, col) + b * p.c_batch;
        if (p.beta != 0.0) v += p.beta * (*c);
        if (p.D) v += p.gamma * p.D[didx_off(p.dm, row) + didx_off(p.dn, col) + b * p.d_batch];
        *c = v;
    }
}

constexpr size_t SMEM_BYTES = (size_t)2 * STAGES * STAGE_ELEMS * sizeof(double) + (BM + BN) * sizeof(long long) +
                              STAGES * sizeof(int) + 16;

bool even(long long x) { return (x & 1) == 0; }

// Decide shared-memory layout and cp.async width of one operand of one segment.
//   mn: index walking the tile rows (M or N), k: contraction index, extent_mn/extent_k: sizes.
void choose_layout(const void* base, const DIdx& mn, long long extent_mn, const DIdx& k, long long extent_k,
                   long long batch_stride, long long nbatch, int* mn_major, int* vec) {
    const bool base16 = (reinterpret_cast<uintptr_t>(base) & 15) == 0;
    const bool mn_flat = mn.div >= (unsigned)extent_mn, k_flat = k.div >= (unsigned)extent_k;
    const bool batch_ok = nbatch <= 1 || even(batch_stride);
    if (k.s_lo == 1) {   // K-contiguous: tile[row][k]
        *mn_major = 0;
        const bool ok = base16 && batch_ok && (k_flat || (even(k.div) && even(k.s_hi))) &&
                        (mn_flat ? true : even(mn.s_hi)) && even(mn.s_lo);
        *vec = ok ? 2 : 1;
        return;
    }
    if (mn.s_lo == 1) {  // row-contiguous: tile[k][row]
        *mn_major = 1;
        const bool ok = base16 && batch_ok && (mn_flat || (even(mn.div) && even(mn.s_hi))) &&
                        (k_flat ? true : even(k.s_hi)) && even(k.s_lo);
        *vec = ok ? 2 : 1;
        return;
    }
    *mn_major = 0;       // no contiguous direction: element-wise gather
    *vec = 1;
}

}  // namespace

extern "C" int pccb_contract(pccb_ctx* ctx, const pccb_contraction* d) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, d != nullptr, "pccb_contract: null descriptor");
    PCCB_REQUIRE(ctx, d->dtype == PCCB_F64, "pccb_contract: only PCCB_F64 is implemented (dtype=%d)", d->dtype);
    PCCB_REQUIRE(ctx, d->nseg >= 1 && d->nseg <= PCCB_MAX_SEGMENTS, "pccb_contract: nseg=%d out of range", d->nseg);
    PCCB_REQUIRE(ctx, d->M > 0 && d->N > 0 && d->batch > 0, "pccb_contract: empty output (M=%lld N=%lld batch=%lld)",
                 (long long)d->M, (long long)d->N, (long long)d->batch);
    PCCB_REQUIRE(ctx, d->M < (1ll << 31) && d->N < (1ll << 31), "pccb_contract: M/N too large");
    PCCB_REQUIRE(ctx, d->C_dev != nullptr, "pccb_contract: C_dev is null");

    DParams p;
    memset(&p, 0, sizeof(p));
    p.M = (unsigned)d->M;
    p.N = (unsigned)d->N;
    int kt = 0, ns = 0;
    for (int s = 0; s < d->nseg; ++s) {
        const pccb_segment& g = d->seg[s];
        PCCB_REQUIRE(ctx, g.K >= 0 && g.K < (1ll << 31), "pccb_contract: segment %d K out of range", s);
        if (g.K == 0) continue;
        PCCB_REQUIRE(ctx, g.A_dev && g.B_dev, "pccb_contract: segment %d has a null operand", s);
        DSeg& o = p.seg[ns++];
        o.A = static_cast<const double*>(g.A_dev);
        o.B = static_cast<const double*>(g.B_dev);
        o.am = to_didx(g.am, d->M);
        o.ak = to_didx(g.ak, g.K);
        o.bn = to_didx(g.bn, d->N);
        o.bk = to_didx(g.bk, g.K);
        o.a_batch = g.a_batch;
        o.b_batch = g.b_batch;
        o.scale = g.scale;
        o.K = (unsigned)g.K;
        o.kt_begin = kt;
        kt += (int)((g.K + BK - 1) / BK);
        choose_layout(o.A, o.am, d->M, o.ak, g.K, o.a_batch, d->batch, &o.a_mn_major, &o.a_vec);
        choose_layout(o.B, o.bn, d->N, o.bk, g.K, o.b_batch, d->batch, &o.b_mn_major, &o.b_vec);
    }
    p.nseg = ns;
    p.total_kt = kt;
    p.C = static_cast<double*>(d->C_dev);
    p.cm = to_didx(d->cm, d->M);
    p.cn = to_didx(d->cn, d->N);
    p.c_batch = d->c_batch;
    p.D = static_cast<const double*>(d->D_dev);
    p.dm = to_didx(d->dm, d->M);
    p.dn = to_didx(d->dn, d->N);
    p.d_batch = d->d_batch;
    p.alpha = d->alpha;
    p.beta = d->beta;
    p.gamma = d->D_dev ? d->gamma : 0.0;
    p.splits = 1;
    p.ws = static_cast<double*>(ctx->ws);

    const long long total = (long long)d->M * d->N * d->batch;
    if (ctx->impl == PCCB_IMPL_NAIVE || ns == 0) {
        // ns == 0: pure epilogue (alpha*0 + beta*C + gamma*D)
        const int threads = 256;
        long long blocks = (total + threads - 1) / threads;
        if (blocks > 148 * 64) blocks = 148 * 64;
        contract_f64_naive_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(p, total);
        PCCB_CHECK_LAUNCH(ctx, "contract_f64_naive_kernel");
        return PCCB_OK;
    }

    const long long gm = (d->M + BM - 1) / BM, gn = (d->N + BN - 1) / BN;
    PCCB_REQUIRE(ctx, gn <= 65535, "pccb_contract: N=%lld needs more than 65535 column tiles", (long long)d->N);
    const long long tiles = gm * gn * d->batch;
    int splits = 1;
    if (tiles < 2ll * ctx->sm_count && kt >= 16) {
        long long want = (2ll * ctx->sm_count + tiles - 1) / tiles;
        long long by_k = kt / 8;                       // at least 8 k-tiles (128 k) per split
        long long by_ws = ctx->ws ? (long long)(ctx->ws_bytes / ((size_t)total * sizeof(double))) : 1;
        long long sK = want < by_k ? want : by_k;
        if (sK > by_ws) sK = by_ws;
        if (sK > 1) splits = (int)sK;
    }
    PCCB_REQUIRE(ctx, (long long)d->batch * splits <= 65535, "pccb_contract: batch*splits=%lld exceeds grid.z",
                 (long long)d->batch * splits);
    p.splits = splits;

    static bool attr_set = false;
    if (!attr_set) {
        PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)SMEM_BYTES));
        attr_set = true;
    }
    dim3 grid((unsigned)gm, (unsigned)gn, (unsigned)(d->batch * splits));
    contract_f64_kernel<<<grid, NTHREADS, SMEM_BYTES, ctx->stream>>>(p);
    PCCB_CHECK_LAUNCH(ctx, "contract_f64_kernel");
    if (splits > 1) {
        const int threads = 256;
        long long blocks = (total + threads - 1) / threads;
        if (blocks > 148 * 32) blocks = 148 * 32;
        splitk_reduce_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(p, total);
        PCCB_CHECK_LAUNCH(ctx, "splitk_reduce_kernel");
    }
    return PCCB_OK;
}
