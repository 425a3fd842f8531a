// pccb_contract, FP32 operands (precision='SP' mode: FP32 contractions, FP64 energy / norm / DIIS).
//
// Same descriptor, same strided compound-index operand model and the same cp.async + mbarrier
// shared-memory ring as the FP64 kernel (contract.cu); what differs is the math pipe:
//   * true FP32 has no tensor-core instruction, and a single TF32 product (10-bit mantissa) does
//     not hold the 1e-6 Eh bar of the mixed-precision mode.  Each FP32 operand element x is
//     therefore split in registers into hi = tf32(x) and lo = tf32(x - hi) and a product is
//     evaluated as a_lo*b_hi + a_hi*b_lo + a_hi*b_hi on the tensor pipe
//     (mma.sync.m16n8k8.tf32, FP32 accumulate): FP32-level accuracy at 1/3 of the TF32
//     mma.sync rate (measured 277.6 TFLOP/s on this pool -> 92 TFLOP/s effective ceiling, vs
//     71 TFLOP/s FFMA peak and 64 TFLOP/s cuBLAS SGEMM; profiles/r01_peaks.json).
//   * CTA tile (32*MJ) x 128 x 32 floats, MJ in {1,2,3,4}; 8 warps as 2(M) x 4(N), warp tile
//     (16*MJ) x 32.  K-major tiles have pitch 36 floats (= 4 mod 32 banks), MN-major tiles
//     pitch rows+8 (= 8 or 24 mod 32): fragment loads are bank-conflict free.
#include <climits>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace {

constexpr int BN = 128, BK = 32, NTHREADS = 256, STAGES = 4;
constexpr int DIST = STAGES - 2;
constexpr int PK = BK + 4;

struct FSeg {
    const float* A;
    const float* B;
    DIdx am, ak, bn, bk;
    long long a_batch, b_batch;
    float scale;
    unsigned K;
    int kt_begin;
    int a_mn_major;
    int b_mn_major;
    int pad_;
};

struct FParams {
    unsigned M, N;
    int nseg, total_kt, splits, nz;
    FSeg seg[PCCB_MAX_SEGMENTS];
    float* C;
    DIdx cm, cn;
    long long c_batch;
    const float* D;
    DIdx dm, dn;
    long long d_batch;
    float alpha, beta, gamma;
    float* ws;
};

template <int BYTES>
__device__ __forceinline__ void cp_async(unsigned dst, const void* src, int bytes) {
    if (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
    else if (BYTES == 8)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cp_async_mbar_arrive(unsigned bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
// producer-side wait: the hardware may suspend the warp for up to ~1 us per try instead of spinning
// on the issue port the consumer warps need
__device__ __forceinline__ void mbar_wait_relaxed(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared.b64 p, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity), "r"(1000u)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ unsigned to_tf32(float x) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;\n" : "=r"(r) : "f"(x));
    return r;
}
// x = hi + lo exactly, hi = x with its 13 low mantissa bits cleared (what the tensor core reads when it
// is handed an FP32 register as a TF32 operand), lo = x - hi.  lo is passed to the MMA as it is: the
// hardware again reads its top 10 mantissa bits, so the pair carries ~21 bits of x for one LOP3 and
// one FADD per element (cvt.rna.tf32 is a multi-instruction sequence on sm_100a and made the consumer
// warps issue-bound: profiles/r01_ncu_ladder_f32_*).
__device__ __forceinline__ void split_tf32(float x, unsigned& hi, unsigned& lo) {
    hi = __float_as_uint(x) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// K-major tile[row][k] (pitch PK): a thread owns one k-chunk column and walks the rows
template <int VEC, int ROWS, int NT = NTHREADS>
__device__ __forceinline__ void load_tile_kmajor(unsigned tile_s, const float* base, const long long* rows, int rlim,
                                                 const DIdx& kidx, unsigned k0, unsigned K, int tid) {
    constexpr int CH = BK / VEC;
    constexpr int RPP = NT / CH;
    constexpr int PASSES = (ROWS + RPP - 1) / RPP;
    const int kc = tid % CH, r0 = tid / CH;
    const unsigned k = k0 + kc * VEC;
    int kbytes = 0;
    const float* bk = base;
    if (k < K) {
        bk = base + didx_off(kidx, k);
        const unsigned left = K - k;
        kbytes = 4 * (int)(left < (unsigned)VEC ? left : (unsigned)VEC);
    }
    const unsigned dst0 = tile_s + (r0 * PK + kc * VEC) * 4;
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
        const int row = r0 + ps * RPP;
        if (ROWS % RPP == 0 || row < ROWS) {
            const int bytes = (row < rlim) ? kbytes : 0;
            const float* src = bytes ? bk + rows[row] : base;
            cp_async<4 * VEC>(dst0 + ps * (RPP * PK * 4), src, bytes);
        }
    }
}

// MN-major tile[k][row] (pitch ROWS+8): a thread owns one row chunk and walks the k rows
template <int VEC, int ROWS, int NT = NTHREADS>
__device__ __forceinline__ void load_tile_mnmajor(unsigned tile_s, const float* base, const long long* rows, int rlim,
                                                  const DIdx& kidx, unsigned k0, unsigned K, int tid) {
    constexpr int PM = ROWS + 8;
    constexpr int CH = ROWS / VEC;
    constexpr int KPP = (NT / CH) < BK ? (NT / CH) : BK;
    constexpr int PASSES = (BK + KPP - 1) / KPP;
    const int mc = tid % CH, kr0 = tid / CH;
    const int row = mc * VEC;
    if (kr0 >= KPP) return;
    int mbytes = 0;
    if (row < rlim) {
        const int left = rlim - row;
        mbytes = 4 * (left < VEC ? left : VEC);
    }
    const float* bm = base + rows[row];
    const unsigned dst0 = tile_s + (kr0 * PM + row) * 4;
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
        const int kr = kr0 + ps * KPP;
        if (BK % KPP == 0 || kr < BK) {
            const unsigned k = k0 + kr;
            const int bytes = (k < K) ? mbytes : 0;
            const float* src = bytes ? bm + didx_off(kidx, k) : base;
            cp_async<4 * VEC>(dst0 + ps * (KPP * PM * 4), src, bytes);
        }
    }
}

// Epilogue of the mma.sync kernels: C = alpha * acc + beta * C + gamma * D over the thread's fragment (rows roff[i][h],
// columns colbase + 8 j + e).  The loads of a block of column groups are issued as plain predicated loads, all of them
// before the first value is used: written as `if (ok) v = beta * C[..]` every element became its own basic block
// (LDG, FMUL, BRA in the SASS) and a warp - which issues in order - paid one L2 / HBM round trip per element; the
// K = o launches of r_T2 (epilogue-bound) ran at 0.7 TB/s.
template <int MJ>
__device__ __forceinline__ void store_tile_f32(const FParams& p, float (&acc)[MJ][4][4], float alpha, const long long (&roff)[MJ][2],
                                               const long long (&droff)[MJ][2], unsigned colbase) {
    const bool use_beta = p.beta != 0.f, use_d = p.D != nullptr;
    constexpr int JG = MJ <= 2 ? 4 : 2;
#pragma unroll
    for (int j0 = 0; j0 < 4; j0 += JG) {
        long long co[JG][2];
        bool cok[JG][2];
        float add[JG][2][MJ][2];
#pragma unroll
        for (int jj = 0; jj < JG; ++jj)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const unsigned col = colbase + (j0 + jj) * 8 + e;
                cok[jj][e] = col < p.N;
                co[jj][e] = cok[jj][e] ? didx_off(p.cn, col) : 0;
#pragma unroll
                for (int i = 0; i < MJ; ++i)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                        add[jj][e][i][h] = 0.f;
            }
        // (masked elements read C[0] / D[0] - always mapped - instead of branching around the load; their result is never stored)
        if (use_beta) {
#pragma unroll
            for (int jj = 0; jj < JG; ++jj)
#pragma unroll
                for (int e = 0; e < 2; ++e)
#pragma unroll
                    for (int i = 0; i < MJ; ++i)
#pragma unroll
                        for (int h = 0; h < 2; ++h)
                            add[jj][e][i][h] = p.C[(cok[jj][e] && roff[i][h] >= 0) ? roff[i][h] + co[jj][e] : 0];
        }
#pragma unroll
        for (int jj = 0; jj < JG; ++jj)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int i = 0; i < MJ; ++i)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                        acc[i][j0 + jj][2 * h + e] = alpha * acc[i][j0 + jj][2 * h + e] + p.beta * add[jj][e][i][h];
        if (use_d) {
#pragma unroll
            for (int jj = 0; jj < JG; ++jj)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const unsigned col = colbase + (j0 + jj) * 8 + e;
                    const long long dco = cok[jj][e] ? didx_off(p.dn, col) : 0;
#pragma unroll
                    for (int i = 0; i < MJ; ++i)
#pragma unroll
                        for (int h = 0; h < 2; ++h)
                            add[jj][e][i][h] = p.D[(cok[jj][e] && roff[i][h] >= 0) ? droff[i][h] + dco : 0];
                }
#pragma unroll
            for (int jj = 0; jj < JG; ++jj)
#pragma unroll
                for (int e = 0; e < 2; ++e)
#pragma unroll
                    for (int i = 0; i < MJ; ++i)
#pragma unroll
                        for (int h = 0; h < 2; ++h) acc[i][j0 + jj][2 * h + e] += p.gamma * add[jj][e][i][h];
        }
#pragma unroll
        for (int jj = 0; jj < JG; ++jj)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int i = 0; i < MJ; ++i)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                        if (cok[jj][e] && roff[i][h] >= 0) p.C[roff[i][h] + co[jj][e]] = acc[i][j0 + jj][2 * h + e];
    }
}


template <int VEC, int MJ>
__global__ void __launch_bounds__(NTHREADS, 1) contract_f32_kernel(const __grid_constant__ FParams p) {
    constexpr int BM = 32 * MJ;
    constexpr int PMA = BM + 8, PMB = BN + 8;
    constexpr int A_STAGE = (BM * PK > BK * PMA) ? BM * PK : BK * PMA;   // floats
    constexpr int B_STAGE = (BN * PK > BK * PMB) ? BN * PK : BK * PMB;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sA = reinterpret_cast<float*>(smem_raw);
    float* sB = sA + STAGES * A_STAGE;
    long long* rowA = reinterpret_cast<long long*>(sB + STAGES * B_STAGE);
    long long* rowB = rowA + BM;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(rowB + BN);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const unsigned m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int batch = blockIdx.z / p.splits, split = blockIdx.z % p.splits;
    const int kt_lo = (int)(((long long)p.total_kt * split) / p.splits);
    const int kt_hi = (int)(((long long)p.total_kt * (split + 1)) / p.splits);
    const int nkt = kt_hi - kt_lo;
    const int mlim = (int)min((unsigned)BM, p.M - m0), nlim = (int)min((unsigned)BN, p.N - n0);

    const unsigned sA_s = (unsigned)__cvta_generic_to_shared(sA);
    const unsigned sB_s = (unsigned)__cvta_generic_to_shared(sB);
    const unsigned full_s = (unsigned)__cvta_generic_to_shared(bars);
    const unsigned empty_s = full_s + 8 * STAGES;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full_s + 8 * s, NTHREADS);
            mbar_init(empty_s + 8 * s, NTHREADS / 32);
        }
    }

    float acc[MJ][4][4];
#pragma unroll
    for (int i = 0; i < MJ; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;

    int cur_seg = -1;
    auto seg_of = [&](int kt) {
        int s = 0;
#pragma unroll
        for (int q = 1; q < PCCB_MAX_SEGMENTS; ++q)
            if (q < p.nseg && kt >= p.seg[q].kt_begin) s = q;
        return s;
    };
    auto set_rows = [&](int s) {
        const FSeg& sg = p.seg[s];
        __syncthreads();
        for (int r = tid; r < BM + BN; r += NTHREADS) {
            if (r < BM) {
                const unsigned m = m0 + r;
                rowA[r] = (m < p.M) ? didx_off(sg.am, m) + (long long)batch * sg.a_batch : 0;
            } else {
                const unsigned n = n0 + (r - BM);
                rowB[r - BM] = (n < p.N) ? didx_off(sg.bn, n) + (long long)batch * sg.b_batch : 0;
            }
        }
        __syncthreads();
        cur_seg = s;
    };
    auto issue_tile = [&](int kt, int st, int s) {
        const FSeg& sg = p.seg[s];
        const unsigned k0 = (unsigned)(kt - sg.kt_begin) * BK;
        const unsigned ta = sA_s + st * (A_STAGE * 4), tb = sB_s + st * (B_STAGE * 4);
        if (sg.a_mn_major) load_tile_mnmajor<VEC, BM>(ta, sg.A, rowA, mlim, sg.ak, k0, sg.K, tid);
        else load_tile_kmajor<VEC, BM>(ta, sg.A, rowA, mlim, sg.ak, k0, sg.K, tid);
        if (sg.b_mn_major) load_tile_mnmajor<VEC, BN>(tb, sg.B, rowB, nlim, sg.bk, k0, sg.K, tid);
        else load_tile_kmajor<VEC, BN>(tb, sg.B, rowB, nlim, sg.bk, k0, sg.K, tid);
        cp_async_mbar_arrive(full_s + 8 * st);
    };

#pragma unroll 1
    for (int s = 0; s < DIST; ++s) {
        if (s < nkt) {
            const int sg = seg_of(kt_lo + s);
            if (sg != cur_seg) set_rows(sg);
            issue_tile(kt_lo + s, s, sg);
        }
    }
    if (cur_seg < 0) __syncthreads();

    const int g = lane >> 2, t = lane & 3;
    int seg_acc = -1;
    float scale_acc = 1.f;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        const int st = it % STAGES;
        const int sgi = seg_of(kt_lo + it);
        if (sgi != seg_acc) {     // acc holds sum/scale_acc: re-express for the new segment's scale
            const float sc = p.seg[sgi].scale;
            if (seg_acc >= 0 && sc != scale_acc) {
                const float r = scale_acc / sc;
#pragma unroll
                for (int i = 0; i < MJ; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[i][j][e] *= r;
            }
            seg_acc = sgi;
            scale_acc = sc;
        }
        {   // refill: k-tile it + DIST goes into the stage that tile it + DIST - STAGES occupied
            const int nxt = it + DIST;
            if (nxt < nkt) {
                const int nseg_id = seg_of(kt_lo + nxt);
                if (nseg_id != cur_seg) set_rows(nseg_id);
                if (nxt >= STAGES) mbar_wait(empty_s + 8 * (nxt % STAGES), ((nxt - STAGES) / STAGES) & 1);
                issue_tile(kt_lo + nxt, nxt % STAGES, nseg_id);
            }
        }
        mbar_wait(full_s + 8 * st, (it / STAGES) & 1);
        const FSeg& sg = p.seg[sgi];
        const int a_rs = sg.a_mn_major ? 1 : PK, a_ks = sg.a_mn_major ? PMA : 1;
        const int b_rs = sg.b_mn_major ? 1 : PK, b_ks = sg.b_mn_major ? PMB : 1;
        const float* ap = sA + st * A_STAGE + (wm * (16 * MJ) + g) * a_rs + t * a_ks;
        const float* bp = sB + st * B_STAGE + (wn * 32 + g) * b_rs + t * b_ks;
#pragma unroll
        for (int ks = 0; ks < BK / 8; ++ks) {
            unsigned bh[4][2], bl[4][2];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float* q = bp + (j * 8) * b_rs + (ks * 8) * b_ks;
                split_tf32(q[0], bh[j][0], bl[j][0]);              // (k = t,     n = g)
                split_tf32(q[4 * b_ks], bh[j][1], bl[j][1]);       // (k = t + 4, n = g)
            }
#pragma unroll
            for (int i = 0; i < MJ; ++i) {
                const float* q = ap + (i * 16) * a_rs + (ks * 8) * a_ks;
                unsigned ah[4], al[4];
                split_tf32(q[0], ah[0], al[0]);                        // (row g,     k t)
                split_tf32(q[8 * a_rs], ah[1], al[1]);                 // (row g + 8, k t)
                split_tf32(q[4 * a_ks], ah[2], al[2]);                 // (row g,     k t + 4)
                split_tf32(q[8 * a_rs + 4 * a_ks], ah[3], al[3]);      // (row g + 8, k t + 4)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    mma_tf32(acc[i][j], al, bh[j]);     // small terms first
                    mma_tf32(acc[i][j], ah, bl[j]);
                    mma_tf32(acc[i][j], ah, bh[j]);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * st);
    }

    // ---- epilogue (two-phase per column group, see contract.cu) -----------------------------------
    const unsigned colbase = n0 + wn * 32 + 2 * t;
    if (p.splits > 1) {
        float* ws = p.ws + (size_t)blockIdx.z * p.M * p.N;
#pragma unroll
        for (int i = 0; i < MJ; ++i)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const unsigned row = m0 + wm * (16 * MJ) + i * 16 + h * 8 + g;
                if (row >= p.M) continue;
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const unsigned col = colbase + j * 8 + e;
                        if (col < p.N) ws[(size_t)row * p.N + col] = scale_acc * acc[i][j][2 * h + e];
                    }
            }
        return;
    }
    const float alpha = p.alpha * scale_acc;
    long long roff[MJ][2], droff[MJ][2];
#pragma unroll
    for (int i = 0; i < MJ; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const unsigned row = m0 + wm * (16 * MJ) + i * 16 + h * 8 + g;
            roff[i][h] = (row < p.M) ? didx_off(p.cm, row) + (long long)batch * p.c_batch : -1;
            droff[i][h] = (row < p.M && p.D) ? didx_off(p.dm, row) + (long long)batch * p.d_batch : 0;
        }
    store_tile_f32<MJ>(p, acc, alpha, roff, droff, colbase);
}

// ---- warp-specialised, persistent variant (see contract.cu): 4 producer warps stage the tiles,
// 8 consumer warps split the FP32 fragments into TF32 hi/lo pairs and issue the three MMAs ----------
constexpr int NCONS = 256, NPROD = 128, NTHREADS_WS = NCONS + NPROD;

__device__ __forceinline__ void bar_sync_producers() {
    asm volatile("bar.sync 1, %0;\n" ::"n"(NPROD) : "memory");
}

template <int VEC, int MJ>
__global__ void __launch_bounds__(NTHREADS_WS, 1) contract_f32_ws_kernel(const __grid_constant__ FParams p) {
    constexpr int BM = 32 * MJ;
    constexpr int PMA = BM + 8, PMB = BN + 8;
    constexpr int A_STAGE = (BM * PK > BK * PMA) ? BM * PK : BK * PMA;   // floats
    constexpr int B_STAGE = (BN * PK > BK * PMB) ? BN * PK : BK * PMB;
    constexpr int S = STAGES;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sA = reinterpret_cast<float*>(smem_raw);
    float* sB = sA + S * A_STAGE;
    long long* rowA = reinterpret_cast<long long*>(sB + S * B_STAGE);
    long long* rowB = rowA + BM;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(rowB + BN);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned gm = (p.M + BM - 1) / BM, gn = (p.N + BN - 1) / BN;
    const long long ntiles = (long long)gm * gn * p.nz;

    const unsigned sA_s = (unsigned)__cvta_generic_to_shared(sA);
    const unsigned sB_s = (unsigned)__cvta_generic_to_shared(sB);
    const unsigned full_s = (unsigned)__cvta_generic_to_shared(bars);
    const unsigned empty_s = full_s + 8 * S;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            mbar_init(full_s + 8 * s, NPROD);
            mbar_init(empty_s + 8 * s, NCONS / 32);
        }
    }
    __syncthreads();

    auto seg_of = [&](int kt) {
        int s = 0;
#pragma unroll
        for (int q = 1; q < PCCB_MAX_SEGMENTS; ++q)
            if (q < p.nseg && kt >= p.seg[q].kt_begin) s = q;
        return s;
    };

    if (warp >= NCONS / 32) {
        // ================================ producers ================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 56;\n");
        const int ptid = tid - NCONS;
        unsigned g = 0;
#pragma unroll 1
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const unsigned mx = (unsigned)(tile % gm);
            const long long rest = tile / gm;
            const unsigned m0 = mx * BM, n0 = (unsigned)(rest % gn) * BN;
            const int z = (int)(rest / gn);
            const int batch = z / p.splits, split = z % p.splits;
            const int kt_lo = (int)(((long long)p.total_kt * split) / p.splits);
            const int nkt = (int)(((long long)p.total_kt * (split + 1)) / p.splits) - kt_lo;
            const int mlim = (int)min((unsigned)BM, p.M - m0), nlim = (int)min((unsigned)BN, p.N - n0);
            int cur_seg = -1;
#pragma unroll 1
            for (int it = 0; it < nkt; ++it, ++g) {
                const int s = seg_of(kt_lo + it);
                const FSeg& sg = p.seg[s];
                if (s != cur_seg) {
                    bar_sync_producers();
                    for (int r = ptid; r < BM + BN; r += NPROD) {
                        if (r < BM) {
                            const unsigned m = m0 + r;
                            rowA[r] = (m < p.M) ? didx_off(sg.am, m) + (long long)batch * sg.a_batch : 0;
                        } else {
                            const unsigned n = n0 + (r - BM);
                            rowB[r - BM] = (n < p.N) ? didx_off(sg.bn, n) + (long long)batch * sg.b_batch : 0;
                        }
                    }
                    bar_sync_producers();
                    cur_seg = s;
                }
                const int st = g % S;
                if (g >= S) mbar_wait_relaxed(empty_s + 8 * st, ((g - S) / S) & 1);
                const unsigned k0 = (unsigned)(kt_lo + it - sg.kt_begin) * BK;
                const unsigned ta = sA_s + st * (A_STAGE * 4), tb = sB_s + st * (B_STAGE * 4);
                if (sg.a_mn_major) load_tile_mnmajor<VEC, BM, NPROD>(ta, sg.A, rowA, mlim, sg.ak, k0, sg.K, ptid);
                else load_tile_kmajor<VEC, BM, NPROD>(ta, sg.A, rowA, mlim, sg.ak, k0, sg.K, ptid);
                if (sg.b_mn_major) load_tile_mnmajor<VEC, BN, NPROD>(tb, sg.B, rowB, nlim, sg.bk, k0, sg.K, ptid);
                else load_tile_kmajor<VEC, BN, NPROD>(tb, sg.B, rowB, nlim, sg.bk, k0, sg.K, ptid);
                cp_async_mbar_arrive(full_s + 8 * st);
            }
        }
        asm volatile("cp.async.wait_all;\n" ::: "memory");
        return;
    }

    // ================================ consumers ================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;\n");
    const int wm = warp >> 2, wn = warp & 3;
    const int g = lane >> 2, t = lane & 3;
    unsigned g0 = 0;
#pragma unroll 1
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const unsigned mx = (unsigned)(tile % gm);
    const long long rest = tile / gm;
    const unsigned m0 = mx * BM, n0 = (unsigned)(rest % gn) * BN;
    const int z = (int)(rest / gn);
    const int batch = z / p.splits, split = z % p.splits;
    const int kt_lo = (int)(((long long)p.total_kt * split) / p.splits);
    const int nkt = (int)(((long long)p.total_kt * (split + 1)) / p.splits) - kt_lo;

    float acc[MJ][4][4];
#pragma unroll
    for (int i = 0; i < MJ; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;

    int seg_acc = -1;
    float scale_acc = 1.f;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        const int st = (g0 + it) % S;
        const int sgi = seg_of(kt_lo + it);
        if (sgi != seg_acc) {
            const float sc = p.seg[sgi].scale;
            if (seg_acc >= 0 && sc != scale_acc) {
                const float r = scale_acc / sc;
#pragma unroll
                for (int i = 0; i < MJ; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[i][j][e] *= r;
            }
            seg_acc = sgi;
            scale_acc = sc;
        }
        mbar_wait(full_s + 8 * st, ((g0 + it) / S) & 1);
        const FSeg& sg = p.seg[sgi];
        const int a_rs = sg.a_mn_major ? 1 : PK, a_ks = sg.a_mn_major ? PMA : 1;
        const int b_rs = sg.b_mn_major ? 1 : PK, b_ks = sg.b_mn_major ? PMB : 1;
        const float* ap = sA + st * A_STAGE + (wm * (16 * MJ) + g) * a_rs + t * a_ks;
        const float* bp = sB + st * B_STAGE + (wn * 32 + g) * b_rs + t * b_ks;
        // raw FP32 fragments of step ks+1 are fetched while the 12*MJ MMAs of step ks issue
        float ra[MJ][4], rb[4][2];
        auto fetch = [&](int ks) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float* q = bp + (j * 8) * b_rs + (ks * 8) * b_ks;
                rb[j][0] = q[0];
                rb[j][1] = q[4 * b_ks];
            }
#pragma unroll
            for (int i = 0; i < MJ; ++i) {
                const float* q = ap + (i * 16) * a_rs + (ks * 8) * a_ks;
                ra[i][0] = q[0];
                ra[i][1] = q[8 * a_rs];
                ra[i][2] = q[4 * a_ks];
                ra[i][3] = q[8 * a_rs + 4 * a_ks];
            }
        };
        fetch(0);
#pragma unroll
        for (int ks = 0; ks < BK / 8; ++ks) {
            unsigned bh[4][2], bl[4][2], ah[MJ][4], al[MJ][4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                split_tf32(rb[j][0], bh[j][0], bl[j][0]);
                split_tf32(rb[j][1], bh[j][1], bl[j][1]);
            }
#pragma unroll
            for (int i = 0; i < MJ; ++i)
#pragma unroll
                for (int e = 0; e < 4; ++e) split_tf32(ra[i][e], ah[i][e], al[i][e]);
            if (ks + 1 < BK / 8) fetch(ks + 1);
            // the three products of one accumulator are dependent: sweep all 4*MJ accumulators per
            // product so that consecutive MMAs are independent (small terms first)
#pragma unroll
            for (int i = 0; i < MJ; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) mma_tf32(acc[i][j], al[i], bh[j]);
#pragma unroll
            for (int i = 0; i < MJ; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) mma_tf32(acc[i][j], ah[i], bl[j]);
#pragma unroll
            for (int i = 0; i < MJ; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) mma_tf32(acc[i][j], ah[i], bh[j]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * st);
    }
    g0 += nkt;

    // ---- epilogue (two-phase per column group, see contract.cu) -----------------------------------
    const unsigned colbase = n0 + wn * 32 + 2 * t;
    if (p.splits > 1) {
        float* ws = p.ws + (size_t)z * p.M * p.N;
#pragma unroll
        for (int i = 0; i < MJ; ++i)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const unsigned row = m0 + wm * (16 * MJ) + i * 16 + h * 8 + g;
                if (row >= p.M) continue;
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const unsigned col = colbase + j * 8 + e;
                        if (col < p.N) ws[(size_t)row * p.N + col] = scale_acc * acc[i][j][2 * h + e];
                    }
            }
        continue;
    }
    const float alpha = p.alpha * scale_acc;
    long long roff[MJ][2], droff[MJ][2];
#pragma unroll
    for (int i = 0; i < MJ; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const unsigned row = m0 + wm * (16 * MJ) + i * 16 + h * 8 + g;
            roff[i][h] = (row < p.M) ? didx_off(p.cm, row) + (long long)batch * p.c_batch : -1;
            droff[i][h] = (row < p.M && p.D) ? didx_off(p.dm, row) + (long long)batch * p.d_batch : 0;
        }
    store_tile_f32<MJ>(p, acc, alpha, roff, droff, colbase);
    }   // tile loop
}

__global__ void splitk_reduce_f32_kernel(const __grid_constant__ FParams p, long long total) {
    const long long MN = (long long)p.M * p.N;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long b = idx / MN, r = idx - b * MN;
        const unsigned row = (unsigned)(r / p.N), col = (unsigned)(r - (long long)row * p.N);
        const float* w = p.ws + (size_t)b * p.splits * MN + r;
        float s = 0.f;
        for (int k = 0; k < p.splits; ++k) s += w[(size_t)k * MN];
        float v = p.alpha * s;
        float* c = p.C + didx_off(p.cm, row) + didx_off(p.cn, col) + b * p.c_batch;
        if (p.beta != 0.f) v += p.beta * (*c);
        if (p.D) v += p.gamma * p.D[didx_off(p.dm, row) + didx_off(p.dn, col) + b * p.d_batch];
        *c = v;
    }
}

// one thread per output element, plain FP32 FMA: the cross-check for the tile kernel (tests only)
__global__ void contract_f32_naive_kernel(const __grid_constant__ FParams p, long long total) {
    const long long MN = (long long)p.M * p.N;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long b = idx / MN, r = idx - b * MN;
        const unsigned row = (unsigned)(r / p.N), col = (unsigned)(r - (long long)row * p.N);
        float s = 0.f;
        for (int q = 0; q < p.nseg; ++q) {
            const FSeg& sg = p.seg[q];
            const float* a = sg.A + didx_off(sg.am, row) + b * sg.a_batch;
            const float* bb = sg.B + didx_off(sg.bn, col) + b * sg.b_batch;
            float t = 0.f;
            for (unsigned k = 0; k < sg.K; ++k) t = fmaf(a[didx_off(sg.ak, k)], bb[didx_off(sg.bk, k)], t);
            s += sg.scale * t;
        }
        float v = p.alpha * s;
        float* c = p.C + didx_off(p.cm, row) + didx_off(p.cn, col) + b * p.c_batch;
        if (p.beta != 0.f) v += p.beta * (*c);
        if (p.D) v += p.gamma * p.D[didx_off(p.dm, row) + didx_off(p.dn, col) + b * p.d_batch];
        *c = v;
    }
}

template <int MJ>
constexpr size_t smem_bytes() {
    constexpr int BM = 32 * MJ;
    constexpr int A_STAGE = (BM * PK > BK * (BM + 8)) ? BM * PK : BK * (BM + 8);
    constexpr int B_STAGE = (BN * PK > BK * (BN + 8)) ? BN * PK : BK * (BN + 8);
    return (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(float) + (BM + BN) * sizeof(long long) +
           2 * STAGES * sizeof(unsigned long long) + 16;
}

bool mult(long long x, int v) { return x % v == 0; }

// largest vector width (floats per cp.async: 4, 2 or 1) this operand allows, and its tile layout
void choose_layout(const void* base, const DIdx& mn, const DIdx& k, long long batch_stride, long long nbatch,
                   int* mn_major, int* vec) {
    const DIdx* c;      // the contiguous index
    const DIdx* s;      // the strided one
    if (k.s_lo == 1) { *mn_major = 0; c = &k; s = &mn; }
    else if (mn.s_lo == 1) { *mn_major = 1; c = &mn; s = &k; }
    else { *mn_major = 0; *vec = 1; return; }
    int best = 1;
    for (int v : {2, 4}) {
        const bool ok = (reinterpret_cast<uintptr_t>(base) % (4 * v) == 0) && (nbatch <= 1 || mult(batch_stride, v)) &&
                        (c->flat || (mult(c->div, v) && mult(c->s_hi, v))) && (s->flat || mult(s->s_hi, v)) &&
                        mult(s->s_lo, v);
        if (ok) best = v; else break;
    }
    *vec = best;
}

template <int VEC, int MJ>
int launch_tile_kernel(pccb_ctx* ctx, const FParams& p, long long batch) {
    constexpr int BM = 32 * MJ;
    constexpr size_t SMEM = smem_bytes<MJ>();
    static bool attr_set = false;
    if (!attr_set) {
        PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f32_kernel<VEC, MJ>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
        attr_set = true;
    }
    dim3 grid((unsigned)((p.M + BM - 1) / BM), (unsigned)((p.N + BN - 1) / BN), (unsigned)(batch * p.splits));
    static const bool use_ws = [] { const char* e = getenv("PCCB_WS"); return e ? atoi(e) != 0 : true; }();
    if (use_ws) {
        static bool attr_ws = false;
        if (!attr_ws) {
            PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f32_ws_kernel<VEC, MJ>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
            attr_ws = true;
        }
        FParams q = p;
        q.nz = (int)(batch * p.splits);
        const long long ntiles = (long long)grid.x * grid.y * grid.z;
        const unsigned nctas = (unsigned)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
        contract_f32_ws_kernel<VEC, MJ><<<nctas, NTHREADS_WS, SMEM, ctx->stream>>>(q);
        PCCB_CHECK_LAUNCH(ctx, "contract_f32_ws_kernel");
        return PCCB_OK;
    }
    contract_f32_kernel<VEC, MJ><<<grid, NTHREADS, SMEM, ctx->stream>>>(p);
    PCCB_CHECK_LAUNCH(ctx, "contract_f32_kernel");
    return PCCB_OK;
}

template <int VEC>
int dispatch_mj(pccb_ctx* ctx, const FParams& p, long long batch, int mj) {
    switch (mj) {
        case 1: return launch_tile_kernel<VEC, 1>(ctx, p, batch);
        case 2: return launch_tile_kernel<VEC, 2>(ctx, p, batch);
        case 3: return launch_tile_kernel<VEC, 3>(ctx, p, batch);
        default: return launch_tile_kernel<VEC, 4>(ctx, p, batch);
    }
}

int choose_mj(long long M) {
    // rows per CTA tile = 32*mj: padded rows weighted by the measured relative efficiency of the tile
    // height (a 32-row tile re-reads B four times as often and has a quarter of the MMAs per fragment load)
    static const int forced = [] { const char* e = getenv("PCCB_MJ"); return e ? atoi(e) : 0; }();
    if (forced >= 1 && forced <= 4) return forced;
    static const double eff[5] = {0.0, 0.60, 0.85, 0.95, 1.00};
    int best = 4;
    double best_cost = 1e300;
    for (int mj : {4, 3, 2, 1}) {
        const long long bm = 32 * mj;
        const double cost = (double)(((M + bm - 1) / bm) * bm) / eff[mj];
        if (cost < best_cost) { best_cost = cost; best = mj; }
    }
    return best;
}

}  // namespace

int pccb_contract_f32(pccb_ctx* ctx, const pccb_contraction* d) {
    FParams p;
    memset(&p, 0, sizeof(p));
    p.M = (unsigned)d->M;
    p.N = (unsigned)d->N;
    int kt = 0, ns = 0, vec = 4;
    for (int s = 0; s < d->nseg; ++s) {
        const pccb_segment& g = d->seg[s];
        PCCB_REQUIRE(ctx, g.K >= 0 && g.K < (1ll << 31), "pccb_contract: segment %d K out of range", s);
        if (g.K == 0 || g.scale == 0.0) continue;
        PCCB_REQUIRE(ctx, g.A_dev && g.B_dev, "pccb_contract: segment %d has a null operand", s);
        FSeg& o = p.seg[ns++];
        o.A = static_cast<const float*>(g.A_dev);
        o.B = static_cast<const float*>(g.B_dev);
        o.am = to_didx(g.am, d->M);
        o.ak = to_didx(g.ak, g.K);
        o.bn = to_didx(g.bn, d->N);
        o.bk = to_didx(g.bk, g.K);
        o.a_batch = g.a_batch;
        o.b_batch = g.b_batch;
        o.scale = (float)g.scale;
        o.K = (unsigned)g.K;
        o.kt_begin = kt;
        kt += (int)((g.K + BK - 1) / BK);
        int va, vb;
        choose_layout(o.A, o.am, o.ak, o.a_batch, d->batch, &o.a_mn_major, &va);
        choose_layout(o.B, o.bn, o.bk, o.b_batch, d->batch, &o.b_mn_major, &vb);
        vec = va < vec ? va : vec;
        vec = vb < vec ? vb : vec;
    }
    p.nseg = ns;
    p.total_kt = kt;
    p.C = static_cast<float*>(d->C_dev);
    p.cm = to_didx(d->cm, d->M);
    p.cn = to_didx(d->cn, d->N);
    p.c_batch = d->c_batch;
    p.D = static_cast<const float*>(d->D_dev);
    p.dm = to_didx(d->dm, d->M);
    p.dn = to_didx(d->dn, d->N);
    p.d_batch = d->d_batch;
    p.alpha = (float)d->alpha;
    p.beta = (float)d->beta;
    p.gamma = d->D_dev ? (float)d->gamma : 0.f;
    p.splits = 1;
    p.ws = static_cast<float*>(ctx->ws);

    const long long total = (long long)d->M * d->N * d->batch;
    if (ctx->impl == PCCB_IMPL_NAIVE || ns == 0) {
        const int threads = 256;
        long long blocks = (total + threads - 1) / threads;
        if (blocks > 148 * 64) blocks = 148 * 64;
        contract_f32_naive_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(p, total);
        PCCB_CHECK_LAUNCH(ctx, "contract_f32_naive_kernel");
        return PCCB_OK;
    }
    {   // big plain-matrix contractions with registered lo planes: tcgen05 / TMEM kernel (contract_tc05.cu)
        const int rc5 = pccb_contract_f32_tc05(ctx, d);
        if (rc5 < 0) return rc5;
        if (rc5 == 1) return PCCB_OK;
    }
    const int mj = choose_mj(d->M);
    const long long bm = 32 * mj;
    const long long gm = (d->M + bm - 1) / bm, gn = (d->N + BN - 1) / BN;
    PCCB_REQUIRE(ctx, gn <= 65535, "pccb_contract: N=%lld needs more than 65535 column tiles", (long long)d->N);
    const long long tiles = gm * gn * d->batch;
    int splits = 1;
    if (tiles < 2ll * ctx->sm_count && kt >= 16) {
        long long want = (2ll * ctx->sm_count + tiles - 1) / tiles;
        long long by_k = kt / 8;
        long long by_ws = ctx->ws ? (long long)(ctx->ws_bytes / ((size_t)total * sizeof(float))) : 1;
        long long sK = want < by_k ? want : by_k;
        if (sK > by_ws) sK = by_ws;
        if (sK > 1) splits = (int)sK;
    }
    PCCB_REQUIRE(ctx, (long long)d->batch * splits <= 65535, "pccb_contract: batch*splits=%lld exceeds grid.z",
                 (long long)d->batch * splits);
    p.splits = splits;
    int rc = vec == 4 ? dispatch_mj<4>(ctx, p, d->batch, mj)
                      : (vec == 2 ? dispatch_mj<2>(ctx, p, d->batch, mj) : dispatch_mj<1>(ctx, p, d->batch, mj));
    if (rc != PCCB_OK) return rc;
    if (splits > 1) {
        const int threads = 256;
        long long blocks = (total + threads - 1) / threads;
        if (blocks > 148 * 32) blocks = 148 * 32;
        splitk_reduce_f32_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(p, total);
        PCCB_CHECK_LAUNCH(ctx, "splitk_reduce_f32_kernel");
    }
    return PCCB_OK;
}
