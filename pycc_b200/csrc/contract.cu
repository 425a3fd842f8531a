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
//   * CTA tile (16*MI) x 128 x 16 with MI in {1,2,3,4,5,6,8} chosen per launch to minimise padded
//     rows (o^2 = 400 -> five 80-row tiles); 8 DMMA warps as 2(M) x 4(N), warp tile (8*MI) x 32:
//     4*MI DMMA per MI+4 LDS.64, so the tensor pipe, not shared memory, is the limiter.
//   * five kernels behind the one entry point (all but the last persistent over output tiles):
//       contract_f64_tma_kernel  plain 2-D operands (o^2 v^4 ladder, Wmnij, Zmbij, and the ring terms over
//                                the engine's plain-matrix copies of the o^2 v^2 operands): one producer
//                                thread issues cp.async.bulk.tensor (TMA) copies, 128-byte swizzle,
//                                fragment rows permuted so that LDS.64 stays conflict free
//                                (35.1 TFLOP/s = 0.95 of peak on the ladder);
//       contract_f64_ws_kernel   compound (gathered) operands: 4 producer warps do the index
//                                arithmetic + cp.async (LDGSTS, zero-fill at the edges), 8 consumer
//                                warps run DMMA + LDS only (30.5-32.9 TFLOP/s);
//       contract_f64_wse_kernel  the same plus a fourth warpgroup that owns the epilogue, for short
//                                contractions (8..64 k-tiles: the (T) / CC3 W build): consumers dump
//                                their accumulators to a staging tile and start the next tile at once;
//                                zero-filled k-steps of segment tails are skipped;
//       contract_f64_tmae_kernel plain operands, short K, optionally batched (3-D tensor maps): TMA producer + DMMA
//                                consumers + epilogue warps, L2-friendly tile order (m-tile groups outermost); the
//                                (T) / CC3 W build over the helper layouts of triples_engine.py: 94.5 % tensor activity
//                                on all three launches (beta = 1 as red.global.add.f64);
//       contract_f64_kernel      MI <= 2 (skinny, memory-bound launches): all warps load and
//                                compute, two CTAs per SM.
//     Stages are guarded by full/empty mbarriers; registers are re-balanced with setmaxnreg.
//   * compound (two-level) einsum indices are resolved per 16-byte chunk by the producers; that is
//     what lets a tile walk e.g. the (i,a) x (m,e) view of t2[i,m,a,e] without any transpose in HBM.
//   * shared tiles of the cp.async kernels keep the global-contiguous direction contiguous
//     ("K-major" [row][k] with pitch 20 doubles, or "MN-major" [k][row] with pitch rows+4): both
//     pitches are = 4 mod 16 eight-byte words, which makes the DMMA fragment loads conflict free.
//   * per-segment scale factors never touch the FP64 pipe in the main loop: the accumulators
//     are rescaled once when the k-loop crosses a segment boundary.
//   * small outputs (Fae, Fmi, Wmnij ...) use deterministic split-K (wave-aware split count)
//     through a caller-owned workspace, reduced in fixed order with the alpha/beta/gamma epilogue.
//   * FP32 operands (precision='SP') are routed to contract_f32.cu: plain-matrix contractions with
//     registered residual planes to the tcgen05 / TMEM 3xTF32 kernel (contract_tc05.cu), the rest to the
//     mma.sync 3xTF32 kernel.
#include <climits>
#include <cstdlib>
#include <cstring>

#include <cuda.h>

#include "common.cuh"

namespace {

constexpr int BN = 128, BK = 16, NTHREADS = 256, STAGES = 4;
constexpr int DIST = STAGES - 2;   // prefetch distance in k-tiles: one spare stage of slack between warps
constexpr int PK = BK + 4;    // pitch of a K-major tile row (doubles)

struct DSeg {
    const double* A;
    const double* B;
    DIdx am, ak, bn, bk;
    long long a_batch, b_batch;
    double scale;
    unsigned K;
    int kt_begin;          // first global k-tile of this segment
    int a_mn_major;        // A tile in shared memory: 0 = [row][k], 1 = [k][row]
    int b_mn_major;
    int bz;                // TMA kernel, batched launch: bit 0 / 1 = A / B carries the batch index (3-D tensor map)
};

struct DParams {
    unsigned M, N;
    int nseg, total_kt, splits, nz;    // nz = batch * splits (grid.z of the non-persistent kernel)
    int kchunk;                        // > 0: split-K deals chunks of kchunk k-tiles round-robin (small-tile kernel only)
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
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async8(unsigned dst, const void* src, int bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
        : "+d"(c0), "+d"(c1)
        : "d"(a), "d"(b));
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
// arrives on `bar` once all cp.async issued so far by this thread have landed
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
template <int VEC>
__device__ __forceinline__ void cp_async_vec(unsigned dst, const void* src, int bytes) {
    if (VEC == 2) cp_async16(dst, src, bytes);
    else cp_async8(dst, src, bytes);
}

// Stage an operand tile of ROWS rows x BK k into shared memory (this thread's share).
//   rows[]  : element offset of each tile row, batch offset included (rows >= rlim unused)
//   K-major : tile[row][k], pitch PK - a thread owns one k-chunk column, walks the rows
//   MN-major: tile[k][row], pitch ROWS+4 - a thread owns one row chunk, walks the k rows
template <int VEC, int ROWS, int NT = NTHREADS>
__device__ __forceinline__ void load_tile_kmajor(unsigned tile_s, const double* base, const long long* rows,
                                                 int rlim, const DIdx& kidx, unsigned k0, unsigned K, int tid) {
    constexpr int CH = BK / VEC;            // chunks per row
    constexpr int RPP = NT / CH;            // rows per pass
    constexpr int PASSES = (ROWS + RPP - 1) / RPP;
    const int kc = tid % CH, r0 = tid / CH;
    const unsigned k = k0 + kc * VEC;
    int kbytes = 0;
    const double* bk = base;
    if (k < K) {
        bk = base + didx_off(kidx, k);
        kbytes = (VEC == 2 && k + 1 < K) ? 16 : 8;
    }
    const unsigned dst0 = tile_s + (r0 * PK + kc * VEC) * 8;
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
        const int row = r0 + ps * RPP;
        if (ROWS % RPP == 0 || row < ROWS) {
            const int bytes = (row < rlim) ? kbytes : 0;
            const double* src = bytes ? bk + rows[row] : base;
            cp_async_vec<VEC>(dst0 + ps * (RPP * PK * 8), src, bytes);
        }
    }
}

template <int VEC, int ROWS, int NT = NTHREADS>
__device__ __forceinline__ void load_tile_mnmajor(unsigned tile_s, const double* base, const long long* rows,
                                                  int rlim, const DIdx& kidx, unsigned k0, unsigned K, int tid) {
    constexpr int PM = ROWS + 4;
    constexpr int CH = ROWS / VEC;                                   // chunks per k-row
    constexpr int KPP = (NT / CH) < BK ? (NT / CH) : BK; // k-rows per pass
    constexpr int PASSES = (BK + KPP - 1) / KPP;
    const int mc = tid % CH, kr0 = tid / CH;
    const int row = mc * VEC;
    if (kr0 >= KPP) return;
    int mbytes = 0;
    if (row < rlim) mbytes = (VEC == 2 && row + 1 < rlim) ? 16 : 8;
    const double* bm = base + rows[row];
    const unsigned dst0 = tile_s + (kr0 * PM + row) * 8;
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
        const int kr = kr0 + ps * KPP;
        if (BK % KPP == 0 || kr < BK) {
            const unsigned k = k0 + kr;
            const int bytes = (k < K) ? mbytes : 0;
            const double* src = bytes ? bm + didx_off(kidx, k) : base;
            cp_async_vec<VEC>(dst0 + ps * (KPP * PM * 8), src, bytes);
        }
    }
}

// Epilogue of the DMMA kernels: C = alpha * acc + beta * C + gamma * D over the thread's fragment (rows roff[i], columns
// colbase + 8 j + {c0, c1}), JG column groups at a time.  Every load of a block is issued - unconditionally, from C[0] /
// D[0] for masked elements - before the first value is used: with `if (ok) v = beta * C[..]` each element was its own
// basic block (LDG, DMUL, BRA in the SASS) and the warp paid one round trip per element.
template <int MI, int JG>
__device__ __forceinline__ void store_tile_f64(const DParams& p, double (&acc)[MI][4][2], double alpha, const long long (&roff)[MI],
                                               const long long (&droff)[MI], unsigned colbase, unsigned c0, unsigned c1) {
    const bool use_beta = p.beta != 0.0, use_d = p.D != nullptr;
#pragma unroll
    for (int j0 = 0; j0 < 4; j0 += JG) {
        long long co[JG][2];
        bool cok[JG][2];
        double add[JG][2][MI];
#pragma unroll
        for (int jj = 0; jj < JG; ++jj)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const unsigned col = colbase + (j0 + jj) * 8 + (e ? c1 : c0);
                cok[jj][e] = col < p.N;
                co[jj][e] = cok[jj][e] ? didx_off(p.cn, col) : 0;
#pragma unroll
                for (int i = 0; i < MI; ++i) add[jj][e][i] = 0.0;
            }
        if (use_beta) {
#pragma unroll
            for (int jj = 0; jj < JG; ++jj)
#pragma unroll
                for (int e = 0; e < 2; ++e)
#pragma unroll
                    for (int i = 0; i < MI; ++i) add[jj][e][i] = p.C[(cok[jj][e] && roff[i] >= 0) ? roff[i] + co[jj][e] : 0];
        }
#pragma unroll
        for (int jj = 0; jj < JG; ++jj)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int i = 0; i < MI; ++i) acc[i][j0 + jj][e] = alpha * acc[i][j0 + jj][e] + p.beta * add[jj][e][i];
        if (use_d) {
#pragma unroll
            for (int jj = 0; jj < JG; ++jj)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const unsigned col = colbase + (j0 + jj) * 8 + (e ? c1 : c0);
                    const long long dco = cok[jj][e] ? didx_off(p.dn, col) : 0;
#pragma unroll
                    for (int i = 0; i < MI; ++i) add[jj][e][i] = p.D[(cok[jj][e] && roff[i] >= 0) ? droff[i] + dco : 0];
                }
#pragma unroll
            for (int jj = 0; jj < JG; ++jj)
#pragma unroll
                for (int e = 0; e < 2; ++e)
#pragma unroll
                    for (int i = 0; i < MI; ++i) acc[i][j0 + jj][e] += p.gamma * add[jj][e][i];
        }
#pragma unroll
        for (int jj = 0; jj < JG; ++jj)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int i = 0; i < MI; ++i)
                    if (cok[jj][e] && roff[i] >= 0) p.C[roff[i] + co[jj][e]] = acc[i][j0 + jj][e];
    }
}


template <int VEC, int MI>
__global__ void __launch_bounds__(NTHREADS, (MI <= 2) ? 2 : 1) contract_f64_kernel(const __grid_constant__ DParams p) {
    constexpr int BM = 16 * MI;
    constexpr int PMA = BM + 4, PMB = BN + 4;
    constexpr int A_STAGE = (BM * PK > BK * PMA) ? BM * PK : BK * PMA;   // doubles
    constexpr int B_STAGE = (BN * PK > BK * PMB) ? BN * PK : BK * PMB;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sA = reinterpret_cast<double*>(smem_raw);
    double* sB = sA + STAGES * A_STAGE;
    long long* rowA = reinterpret_cast<long long*>(sB + STAGES * B_STAGE);
    long long* rowB = rowA + BM;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(rowB + BN);   // full[S], empty[S]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const unsigned m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int batch = blockIdx.z / p.splits, split = blockIdx.z % p.splits;
    // split-K work assignment.  Contiguous k ranges by default; with p.kchunk the k-tiles are dealt
    // round-robin in chunks, so that at any moment all CTAs read the same narrow k window of every
    // operand row (a skinny output with a huge K otherwise has #CTAs x #rows far-apart streams in
    // flight, which thrashes the TLB and the DRAM pages).
    const int kchunk = p.kchunk;
    const int kt_lo = (int)(((long long)p.total_kt * split) / p.splits);
    int nkt = (int)(((long long)p.total_kt * (split + 1)) / p.splits) - kt_lo;
    if (kchunk) {
        const int nch = (p.total_kt + kchunk - 1) / kchunk;
        const int mine = split < nch ? (nch - split + p.splits - 1) / p.splits : 0;
        nkt = mine * kchunk;
        if (mine > 0 && split + (mine - 1) * p.splits == nch - 1) nkt -= nch * kchunk - p.total_kt;
    }
    auto gkt = [&](int l) { return kchunk ? ((l / kchunk) * p.splits + split) * kchunk + (l % kchunk) : kt_lo + l; };
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

    double acc[MI][4][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    int cur_seg = -1;   // segment whose row tables are in shared memory (CTA-uniform)

    // segment of global k-tile kt
    auto seg_of = [&](int kt) {
        int s = 0;
#pragma unroll
        for (int q = 1; q < PCCB_MAX_SEGMENTS; ++q)
            if (q < p.nseg && kt >= p.seg[q].kt_begin) s = q;
        return s;
    };
    // (re)build the per-row offset tables when the load front enters a new segment (CTA-uniform)
    auto set_rows = [&](int s) {
        const DSeg& sg = p.seg[s];
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
    // issue this thread's share of the copies for global k-tile kt into stage st
    auto issue_tile = [&](int kt, int st, int s) {
        const DSeg& sg = p.seg[s];
        const unsigned k0 = (unsigned)(kt - sg.kt_begin) * BK;
        const unsigned ta = sA_s + st * (A_STAGE * 8), tb = sB_s + st * (B_STAGE * 8);
        if (sg.a_mn_major) load_tile_mnmajor<VEC, BM>(ta, sg.A, rowA, mlim, sg.ak, k0, sg.K, tid);
        else load_tile_kmajor<VEC, BM>(ta, sg.A, rowA, mlim, sg.ak, k0, sg.K, tid);
        if (sg.b_mn_major) load_tile_mnmajor<VEC, BN>(tb, sg.B, rowB, nlim, sg.bk, k0, sg.K, tid);
        else load_tile_kmajor<VEC, BN>(tb, sg.B, rowB, nlim, sg.bk, k0, sg.K, tid);
        cp_async_mbar_arrive(full_s + 8 * st);
    };

    // prologue: tiles 0 .. DIST-1   (set_rows' barriers also publish the mbarrier inits)
#pragma unroll 1
    for (int s = 0; s < DIST; ++s) {
        if (s < nkt) {
            const int sg = seg_of(gkt(s));
            if (sg != cur_seg) set_rows(sg);
            issue_tile(gkt(s), s, sg);
        }
    }
    if (cur_seg < 0) __syncthreads();   // nkt == 0 cannot happen (host), kept for safety

    const int frow = lane >> 2, fk = lane & 3;

    // Fragment addressing of the tile held by a stage (layout is a per-segment property).
    struct FragPtr {
        const double* ap;
        const double* bp;
        int a_rs, a_ks, b_rs, b_ks;
    };
    auto frag_ptr = [&](int it_) {
        const DSeg& sg = p.seg[seg_of(gkt(it_))];
        const int st = it_ % STAGES;
        FragPtr f;
        f.a_rs = sg.a_mn_major ? 1 : PK;
        f.a_ks = sg.a_mn_major ? PMA : 1;
        f.b_rs = sg.b_mn_major ? 1 : PK;
        f.b_ks = sg.b_mn_major ? PMB : 1;
        f.ap = sA + st * A_STAGE + (wm * (8 * MI) + frow) * f.a_rs + fk * f.a_ks;
        f.bp = sB + st * B_STAGE + (wn * 32 + frow) * f.b_rs + fk * f.b_ks;
        return f;
    };

    // Software pipeline over (k-tile, k-step): the fragments of step g+1 are fetched from shared
    // memory while the DMMAs of step g issue (a[i] is refilled right after its last use, only b
    // is double-buffered), so a warp never waits on LDS latency between DMMA groups.
    double a[MI], b[4], bnx[4];
    mbar_wait(full_s, 0);
    FragPtr fp = frag_ptr(0);
#pragma unroll
    for (int i = 0; i < MI; ++i) a[i] = fp.ap[i * 8 * fp.a_rs];
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = fp.bp[j * 8 * fp.b_rs];

    int seg_acc = -1;            // segment the accumulators are currently scaled for
    double scale_acc = 1.0;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        const int st = it % STAGES;
        {
            const int sgi = seg_of(gkt(it));
            if (sgi != seg_acc) {    // rare, warp-uniform: acc holds sum/scale_acc; re-express for the new scale
                const double sc = p.seg[sgi].scale;
                if (seg_acc >= 0 && sc != scale_acc) {
                    const double r = scale_acc / sc;
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) { acc[i][j][0] *= r; acc[i][j][1] *= r; }
                }
                seg_acc = sgi;
                scale_acc = sc;
            }
        }
        FragPtr fn = fp;
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            const int nks = (ks + 1) % (BK / 4);
            bool more = true;
            if (ks == BK / 4 - 1) {          // next step lives in the next k-tile
                more = it + 1 < nkt;
                if (more) {
                    mbar_wait(full_s + 8 * ((it + 1) % STAGES), ((it + 1) / STAGES) & 1);
                    fn = frag_ptr(it + 1);
                }
            }
            if (more) {
#pragma unroll
                for (int j = 0; j < 4; ++j) bnx[j] = fn.bp[j * 8 * fn.b_rs + nks * 4 * fn.b_ks];
            }
#pragma unroll
            for (int i = 0; i < MI; ++i) {
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                if (more) a[i] = fn.ap[i * 8 * fn.a_rs + nks * 4 * fn.a_ks];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bnx[j];

            if (ks == 0) {
                // refill: k-tile it + DIST goes into the stage that tile it + DIST - STAGES occupied
                const int nxt = it + DIST;
                if (nxt < nkt) {
                    const int nseg_id = seg_of(gkt(nxt));
                    if (nseg_id != cur_seg) set_rows(nseg_id);
                    if (nxt >= STAGES) mbar_wait(empty_s + 8 * (nxt % STAGES), ((nxt - STAGES) / STAGES) & 1);
                    issue_tile(gkt(nxt), nxt % STAGES, nseg_id);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * st);
        fp = fn;
    }

    // ---- epilogue ----------------------------------------------------------------------
    const unsigned colbase = n0 + wn * 32 + 2 * fk;
    if (p.splits > 1) {
        double* ws = p.ws + (size_t)blockIdx.z * p.M * p.N;
#pragma unroll
        for (int i = 0; i < MI; ++i) {
            const unsigned row = m0 + wm * (8 * MI) + i * 8 + frow;
            if (row >= p.M) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const unsigned col = colbase + j * 8 + e;
                    if (col < p.N) ws[(size_t)row * p.N + col] = scale_acc * acc[i][j][e];
                }
        }
        return;
    }
    const double alpha = p.alpha * scale_acc;
    long long roff[MI], droff[MI];
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const unsigned row = m0 + wm * (8 * MI) + i * 8 + frow;
        roff[i] = (row < p.M) ? didx_off(p.cm, row) + (long long)batch * p.c_batch : -1;
        droff[i] = (row < p.M && p.D) ? didx_off(p.dm, row) + (long long)batch * p.d_batch : 0;
    }
    store_tile_f64<MI, 1>(p, acc, alpha, roff, droff, colbase, 0u, 1u);
}

// ---- warp-specialised variant ------------------------------------------------------------------
// 8 consumer warps (2(M) x 4(N), DMMA + LDS.64 only) + 4 producer warps that own the whole
// operand staging: compound-index address arithmetic, cp.async (LDGSTS) issue and the
// full/empty mbarrier hand-shake.  The consumers' instruction stream is then DMMA-dense, and the
// producers run up to WS_STAGES k-tiles ahead.  Registers are re-balanced with setmaxnreg
// (producers 56, consumers 216 per thread).
constexpr int NCONS = 256, NPROD = 128, NTHREADS_WS = NCONS + NPROD, WS_STAGES = 4, WS_STAGES_SKINNY = 8;

__device__ __forceinline__ void bar_sync_producers() {
    asm volatile("bar.sync 1, %0;\n" ::"n"(NPROD) : "memory");
}

template <int VEC, int MI>
__global__ void __launch_bounds__(NTHREADS_WS, 1) contract_f64_ws_kernel(const __grid_constant__ DParams p) {
    constexpr int BM = 16 * MI;
    constexpr int PMA = BM + 4, PMB = BN + 4;
    constexpr int A_STAGE = (BM * PK > BK * PMA) ? BM * PK : BK * PMA;   // doubles
    constexpr int B_STAGE = (BN * PK > BK * PMB) ? BN * PK : BK * PMB;
    constexpr int S = (MI <= 2) ? WS_STAGES_SKINNY : WS_STAGES;   // skinny tiles are bandwidth-bound: deeper ring

    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sA = reinterpret_cast<double*>(smem_raw);
    double* sB = sA + S * A_STAGE;
    long long* rowA = reinterpret_cast<long long*>(sB + S * B_STAGE);
    long long* rowB = rowA + BM;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(rowB + BN);   // full[S], empty[S]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // persistent CTA: walks the output tiles (m fastest, so that the CTAs of one wave share B
    // tiles in L2); the shared-memory ring and its barrier phases run on across tiles, which lets
    // the producers stage the next tile while the consumers are still in this tile's epilogue.
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
        unsigned g = 0;     // k-tiles staged so far by this CTA (ring position)
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
            const DSeg& sg = p.seg[s];
            if (s != cur_seg) {      // per-row offset tables of the segment (producer-private)
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
            const unsigned ta = sA_s + st * (A_STAGE * 8), tb = sB_s + st * (B_STAGE * 8);
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
    const int frow = lane >> 2, fk = lane & 3;
    unsigned g0 = 0;    // ring position of this tile's first k-tile
#pragma unroll 1
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const unsigned mx = (unsigned)(tile % gm);
    const long long rest = tile / gm;
    const unsigned m0 = mx * BM, n0 = (unsigned)(rest % gn) * BN;
    const int z = (int)(rest / gn);
    const int batch = z / p.splits, split = z % p.splits;
    const int kt_lo = (int)(((long long)p.total_kt * split) / p.splits);
    const int nkt = (int)(((long long)p.total_kt * (split + 1)) / p.splits) - kt_lo;

    double acc[MI][4][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    struct FragPtr {
        const double* ap;
        const double* bp;
        int a_rs, a_ks, b_rs, b_ks;
    };
    auto frag_ptr = [&](int it_) {
        const DSeg& sg = p.seg[seg_of(kt_lo + it_)];
        const int st = (g0 + it_) % S;
        FragPtr f;
        f.a_rs = sg.a_mn_major ? 1 : PK;
        f.a_ks = sg.a_mn_major ? PMA : 1;
        f.b_rs = sg.b_mn_major ? 1 : PK;
        f.b_ks = sg.b_mn_major ? PMB : 1;
        f.ap = sA + st * A_STAGE + (wm * (8 * MI) + frow) * f.a_rs + fk * f.a_ks;
        f.bp = sB + st * B_STAGE + (wn * 32 + frow) * f.b_rs + fk * f.b_ks;
        return f;
    };

    double a[MI], b[4], bnx[4];
    mbar_wait(full_s + 8 * (g0 % S), (g0 / S) & 1);
    FragPtr fp = frag_ptr(0);
#pragma unroll
    for (int i = 0; i < MI; ++i) a[i] = fp.ap[i * 8 * fp.a_rs];
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = fp.bp[j * 8 * fp.b_rs];

    int seg_acc = -1;
    double scale_acc = 1.0;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        const int st = (g0 + it) % S;
        {
            const int sgi = seg_of(kt_lo + it);
            if (sgi != seg_acc) {
                const double sc = p.seg[sgi].scale;
                if (seg_acc >= 0 && sc != scale_acc) {
                    const double r = scale_acc / sc;
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) { acc[i][j][0] *= r; acc[i][j][1] *= r; }
                }
                seg_acc = sgi;
                scale_acc = sc;
            }
        }
        FragPtr fn = fp;
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            const int nks = (ks + 1) % (BK / 4);
            bool more = true;
            if (ks == BK / 4 - 1) {
                more = it + 1 < nkt;
                if (more) {
                    mbar_wait(full_s + 8 * ((g0 + it + 1) % S), ((g0 + it + 1) / S) & 1);
                    fn = frag_ptr(it + 1);
                }
            }
            if (more) {
#pragma unroll
                for (int j = 0; j < 4; ++j) bnx[j] = fn.bp[j * 8 * fn.b_rs + nks * 4 * fn.b_ks];
            }
#pragma unroll
            for (int i = 0; i < MI; ++i) {
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                if (more) a[i] = fn.ap[i * 8 * fn.a_rs + nks * 4 * fn.a_ks];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bnx[j];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * st);
        fp = fn;
    }
    g0 += nkt;

    // ---- epilogue ----------------------------------------------------------------------
    const unsigned colbase = n0 + wn * 32 + 2 * fk;
    if (p.splits > 1) {
        double* ws = p.ws + (size_t)z * p.M * p.N;
#pragma unroll
        for (int i = 0; i < MI; ++i) {
            const unsigned row = m0 + wm * (8 * MI) + i * 8 + frow;
            if (row >= p.M) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const unsigned col = colbase + j * 8 + e;
                    if (col < p.N) ws[(size_t)row * p.N + col] = scale_acc * acc[i][j][e];
                }
        }
        continue;
    }
    const double alpha = p.alpha * scale_acc;
    long long roff[MI], droff[MI];
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const unsigned row = m0 + wm * (8 * MI) + i * 8 + frow;
        roff[i] = (row < p.M) ? didx_off(p.cm, row) + (long long)batch * p.c_batch : -1;
        droff[i] = (row < p.M && p.D) ? didx_off(p.dm, row) + (long long)batch * p.d_batch : 0;
    }
    // (the consumers own 216 registers: up to MI = 4 the whole tile is one block of loads, MI = 5 two, MI = 6 four)
    store_tile_f64<MI, (MI <= 4) ? 4 : ((MI <= 5) ? 2 : 1)>(p, acc, alpha, roff, droff, colbase, 0u, 1u);
    }   // tile loop
}

// ---- warp-specialised variant with an asynchronous epilogue ---------------------------------------
// For short contractions (a few dozen k-tiles per output tile: the (T) / CC3 W build with
// K = 2(v + o), r2 += Wmnij tau with K = o^2) the consumers' own epilogue - 8*MI strided stores per
// thread, twice that with beta != 0 - is 15-30 % of the tile time and leaves the FP64 pipe idle
// (ncu: 81 % / 70 % busy on the (T) launches).  Here the consumers only dump their accumulators
// into a shared staging tile (STS) and go straight on to the next tile's k-loop; a fourth
// warpgroup drains the staging tile to global memory (alpha / beta / gamma, compound output
// strides) with the lanes laid along whichever output direction is contiguous, so the
// read-modify-write of C is coalesced and overlaps the next tile's DMMAs.
constexpr int NEPI = 128, NTHREADS_WSE = NCONS + NPROD + NEPI;
constexpr int SPITCH = BN + 1;      // staging pitch (doubles): odd, so column reads are conflict free
constexpr int WSE_MAX_KT = 64;      // beyond this the epilogue is < 5 % of the tile and the plain kernel has more stages
constexpr int WSE_MIN_KT = 8;       // below this the launch is a memory-bound read-modify-write of C: 8 storing warps beat 4
template <int MI> constexpr int wse_stages() { return MI <= 5 ? 4 : 3; }

template <int VEC, int MI>
__global__ void __launch_bounds__(NTHREADS_WSE, 1) contract_f64_wse_kernel(const __grid_constant__ DParams p) {
    constexpr int BM = 16 * MI;
    constexpr int PMA = BM + 4, PMB = BN + 4;
    constexpr int A_STAGE = (BM * PK > BK * PMA) ? BM * PK : BK * PMA;   // doubles
    constexpr int B_STAGE = (BN * PK > BK * PMB) ? BN * PK : BK * PMB;
    constexpr int S = (MI <= 5) ? 4 : 3;     // = wse_stages<MI>()

    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sA = reinterpret_cast<double*>(smem_raw);
    double* sB = sA + S * A_STAGE;
    double* stg = sB + S * B_STAGE;                                      // [BM][SPITCH]
    long long* rowA = reinterpret_cast<long long*>(stg + BM * SPITCH);
    long long* rowB = rowA + BM;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(rowB + BN);   // full[S], empty[S], staged, drained

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned gm = (p.M + BM - 1) / BM, gn = (p.N + BN - 1) / BN;
    const long long ntiles = (long long)gm * gn * p.nz;     // splits == 1 (host): z is the batch index

    const unsigned sA_s = (unsigned)__cvta_generic_to_shared(sA);
    const unsigned sB_s = (unsigned)__cvta_generic_to_shared(sB);
    const unsigned full_s = (unsigned)__cvta_generic_to_shared(bars);
    const unsigned empty_s = full_s + 8 * S;
    const unsigned staged_s = empty_s + 8 * S, drained_s = staged_s + 8;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            mbar_init(full_s + 8 * s, NPROD);
            mbar_init(empty_s + 8 * s, NCONS / 32);
        }
        mbar_init(staged_s, NCONS / 32);
        mbar_init(drained_s, NEPI / 32);
    }
    __syncthreads();

    auto seg_of = [&](int kt) {
        int s = 0;
#pragma unroll
        for (int q = 1; q < PCCB_MAX_SEGMENTS; ++q)
            if (q < p.nseg && kt >= p.seg[q].kt_begin) s = q;
        return s;
    };

    if (warp >= (NCONS + NPROD) / 32) {
        // ================================ epilogue warps ================================
        // (no D operand here: the host routes gamma * D contractions to the plain kernel)
        asm volatile("setmaxnreg.dec.sync.aligned.u32 88;\n");
        const int ew = (tid - (NCONS + NPROD)) >> 5;
        const bool use_beta = p.beta != 0.0;
        const bool along_m = p.cm.s_lo == 1 && p.cn.s_lo != 1;     // lanes along the contiguous output direction
        const double alpha = p.alpha, beta = p.beta;
        unsigned lt = 0;
#pragma unroll 1
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            const unsigned mx = (unsigned)(tile % gm);
            const long long rest = tile / gm;
            const unsigned m0 = mx * BM, n0 = (unsigned)(rest % gn) * BN;
            const int mlim = (int)min((unsigned)BM, p.M - m0), nlim = (int)min((unsigned)BN, p.N - n0);
            double* Cb = p.C + (rest / gn) * p.c_batch;
            mbar_wait(staged_s, lt & 1);
            if (!along_m) {
                // a warp takes rows ew, ew+4, ...; its lanes cover the 128 columns in 4 passes
                constexpr int CP = BN / 32, RU = 4;
                long long co[CP];
#pragma unroll
                for (int q = 0; q < CP; ++q) {
                    const int c = q * 32 + lane;
                    co[q] = c < nlim ? didx_off(p.cn, n0 + c) : -1;
                }
                const double* sl = stg + lane;
#pragma unroll 1
                for (int r0 = ew; r0 < mlim; r0 += 4 * RU) {
                    double v[RU][CP];
                    double* crow[RU];
#pragma unroll
                    for (int u = 0; u < RU; ++u) {
                        const int r = r0 + 4 * u;
                        crow[u] = nullptr;
                        if (r < mlim) {
                            crow[u] = Cb + didx_off(p.cm, m0 + r);
#pragma unroll
                            for (int q = 0; q < CP; ++q) {
                                double x = alpha * sl[r * SPITCH + q * 32];
                                if (use_beta && co[q] >= 0) x += beta * crow[u][co[q]];
                                v[u][q] = x;
                            }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < RU; ++u)
                        if (crow[u]) {
#pragma unroll
                            for (int q = 0; q < CP; ++q)
                                if (co[q] >= 0) crow[u][co[q]] = v[u][q];
                        }
                }
            } else {
                // a warp takes columns ew, ew+4, ...; its lanes cover the BM rows
                constexpr int RP = (BM + 31) / 32, CU = 4;
                long long ro[RP];
#pragma unroll
                for (int q = 0; q < RP; ++q) {
                    const int r = q * 32 + lane;
                    ro[q] = r < mlim ? didx_off(p.cm, m0 + r) : -1;
                }
                const double* sl = stg + lane * SPITCH;
#pragma unroll 1
                for (int c0 = ew; c0 < nlim; c0 += 4 * CU) {
                    double v[CU][RP];
                    double* ccol[CU];
#pragma unroll
                    for (int u = 0; u < CU; ++u) {
                        const int c = c0 + 4 * u;
                        ccol[u] = nullptr;
                        if (c < nlim) {
                            ccol[u] = Cb + didx_off(p.cn, n0 + c);
#pragma unroll
                            for (int q = 0; q < RP; ++q) {
                                double x = 0.0;
                                if (ro[q] >= 0) {
                                    x = alpha * sl[q * 32 * SPITCH + c];
                                    if (use_beta) x += beta * ccol[u][ro[q]];
                                }
                                v[u][q] = x;
                            }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < CU; ++u)
                        if (ccol[u]) {
#pragma unroll
                            for (int q = 0; q < RP; ++q)
                                if (ro[q] >= 0) ccol[u][ro[q]] = v[u][q];
                        }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(drained_s);
        }
        return;
    }

    if (warp >= NCONS / 32) {
        // ================================ producers ================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 56;\n");
        const int ptid = tid - NCONS;
        unsigned g = 0;     // k-tiles staged so far by this CTA (ring position)
#pragma unroll 1
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const unsigned mx = (unsigned)(tile % gm);
        const long long rest = tile / gm;
        const unsigned m0 = mx * BM, n0 = (unsigned)(rest % gn) * BN;
        const int batch = (int)(rest / gn);
        const int nkt = p.total_kt;
        const int mlim = (int)min((unsigned)BM, p.M - m0), nlim = (int)min((unsigned)BN, p.N - n0);
        int cur_seg = -1;
#pragma unroll 1
        for (int it = 0; it < nkt; ++it, ++g) {
            const int s = seg_of(it);
            const DSeg& sg = p.seg[s];
            if (s != cur_seg) {      // per-row offset tables of the segment (producer-private)
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
            const unsigned k0 = (unsigned)(it - sg.kt_begin) * BK;
            const unsigned ta = sA_s + st * (A_STAGE * 8), tb = sB_s + st * (B_STAGE * 8);
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
    asm volatile("setmaxnreg.inc.sync.aligned.u32 184;\n");
    const int wm = warp >> 2, wn = warp & 3;
    const int frow = lane >> 2, fk = lane & 3;
    unsigned g0 = 0;    // ring position of this tile's first k-tile
    unsigned lt = 0;    // tiles finished by this CTA
    const int nkt = p.total_kt;
#pragma unroll 1
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
    double acc[MI][4][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    struct FragPtr {
        const double* ap;
        const double* bp;
        int a_rs, a_ks, b_rs, b_ks;
    };
    auto frag_ptr = [&](int it_) {
        const DSeg& sg = p.seg[seg_of(it_)];
        const int st = (g0 + it_) % S;
        FragPtr f;
        f.a_rs = sg.a_mn_major ? 1 : PK;
        f.a_ks = sg.a_mn_major ? PMA : 1;
        f.b_rs = sg.b_mn_major ? 1 : PK;
        f.b_ks = sg.b_mn_major ? PMB : 1;
        f.ap = sA + st * A_STAGE + (wm * (8 * MI) + frow) * f.a_rs + fk * f.a_ks;
        f.bp = sB + st * B_STAGE + (wn * 32 + frow) * f.b_rs + fk * f.b_ks;
        return f;
    };

    double a[MI], b[4], bnx[4];
    mbar_wait(full_s + 8 * (g0 % S), (g0 / S) & 1);
    FragPtr fp = frag_ptr(0);
#pragma unroll
    for (int i = 0; i < MI; ++i) a[i] = fp.ap[i * 8 * fp.a_rs];
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = fp.bp[j * 8 * fp.b_rs];

    int seg_acc = -1;
    double scale_acc = 1.0;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        const int st = (g0 + it) % S;
        {
            const int sgi = seg_of(it);
            if (sgi != seg_acc) {
                const double sc = p.seg[sgi].scale;
                if (seg_acc >= 0 && sc != scale_acc) {
                    const double r = scale_acc / sc;
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) { acc[i][j][0] *= r; acc[i][j][1] *= r; }
                }
                seg_acc = sgi;
                scale_acc = sc;
            }
        }
        FragPtr fn = fp;
        // k-steps (of 4) of this k-tile that hold data: the tail of a segment is zero-filled by the
        // producers, and whole zero steps are skipped (K = o = 20 needs 5 steps, not 8)
        int nsteps = BK / 4;
        {
            const DSeg& sg = p.seg[seg_acc];
            const unsigned left = sg.K - (unsigned)(it - sg.kt_begin) * BK;
            if (left < (unsigned)BK) nsteps = (int)((left + 3) >> 2);
        }
        if (nsteps == BK / 4) {
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            const int nks = (ks + 1) % (BK / 4);
            // after the last k-tile the "next" fragments are re-read from the current stage (valid
            // shared memory, values unused): the fragment loads stay unconditional, so the compiler
            // has no reason to predicate / duplicate the DMMAs around them
            if (ks == BK / 4 - 1 && it + 1 < nkt) {
                mbar_wait(full_s + 8 * ((g0 + it + 1) % S), ((g0 + it + 1) / S) & 1);
                fn = frag_ptr(it + 1);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) bnx[j] = fn.bp[j * 8 * fn.b_rs + nks * 4 * fn.b_ks];
#pragma unroll
            for (int i = 0; i < MI; ++i) {
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                a[i] = fn.ap[i * 8 * fn.a_rs + nks * 4 * fn.a_ks];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bnx[j];
        }
        } else {
            // partial k-tile (rare: the last one of a segment): plain step loop, then the hand-over to
            // the next k-tile's first fragments
#pragma unroll 1
            for (int ks = 0; ks < nsteps; ++ks) {
                const bool last = ks + 1 == nsteps;
                if (last && it + 1 < nkt) {
                    mbar_wait(full_s + 8 * ((g0 + it + 1) % S), ((g0 + it + 1) / S) & 1);
                    fn = frag_ptr(it + 1);
                }
                const int nks = last ? 0 : ks + 1;
#pragma unroll
                for (int j = 0; j < 4; ++j) bnx[j] = fn.bp[j * 8 * fn.b_rs + nks * 4 * fn.b_ks];
#pragma unroll
                for (int i = 0; i < MI; ++i) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                    a[i] = fn.ap[i * 8 * fn.a_rs + nks * 4 * fn.a_ks];
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) b[j] = bnx[j];
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * st);
        fp = fn;
    }
    g0 += nkt;

    // ---- hand the tile to the epilogue warps -------------------------------------------------
    if (lt > 0) mbar_wait(drained_s, (lt - 1) & 1);      // the previous tile has left the staging buffer
    {
        double* sp = stg + (wm * (8 * MI) + frow) * SPITCH + wn * 32 + 2 * fk;
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                sp[i * 8 * SPITCH + j * 8] = scale_acc * acc[i][j][0];
                sp[i * 8 * SPITCH + j * 8 + 1] = scale_acc * acc[i][j][1];
            }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(staged_s);
    }   // tile loop
}

// ---- TMA-staged variant (plain 2-D operands: the o^2 v^4 ladder, Wmnij, Zmbij) ------------------
// When both operands of every segment are K-contiguous matrices with a uniform row stride, the
// producer is ONE thread issuing two cp.async.bulk.tensor (TMA) copies per k-tile; completion is
// counted in bytes on the stage's full barrier.  Tiles land dense ([row][16 doubles] = 128-byte
// rows) under the 128-byte swizzle: 16-byte chunk c of row r sits at chunk c ^ (r & 7).  The DMMA
// fragment loads (8 rows x 4 k per instruction, LDS.64) stay bank-conflict free because fragment
// row g of every 8-row group is mapped to tile row perm(g) = 2(g & 3) + (g >> 2): the rows a
// half-warp touches are then {0,2,4,6} or {1,3,5,7}, whose XOR patterns hit 8 distinct chunks.
// The epilogue applies the same relabeling to the output rows / columns.
constexpr int TMA_THREADS = NCONS + 128, TMA_STAGES = 6;   // producer warpgroup: one working thread (setmaxnreg needs 4 warps)

struct alignas(64) TmaMaps {
    CUtensorMap a[PCCB_MAX_SEGMENTS];
    CUtensorMap b[PCCB_MAX_SEGMENTS];
};

__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap* map, int c0, int c1, unsigned bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
        ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
}

__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* map, int c0, int c1, int c2, unsigned bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n"
        ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
        : "memory");
}

// BATCHED: the operands are 3-D tensors (k, row, batch) - the (T) / CC3 W build over plain-matrix operands, one
// launch for all k of an occupied pair; an operand that does not depend on the batch index has a batch
// dimension of extent 1 and is always read at batch coordinate 0.
template <int MI, bool BATCHED>
__global__ void __launch_bounds__(TMA_THREADS, 1)
contract_f64_tma_kernel(const __grid_constant__ DParams p, const __grid_constant__ TmaMaps maps) {
    constexpr int BM = 16 * MI;
    constexpr int A_STAGE = BM * BK, B_STAGE = BN * BK;       // doubles; rows of 128 bytes
    constexpr int S = TMA_STAGES;

    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // dynamic shared memory is only guaranteed 16-byte aligned: align the ring to 1024 by hand
    const unsigned raw_s = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned pad = (1024u - (raw_s & 1023u)) & 1023u;
    double* sA = reinterpret_cast<double*>(smem_raw + pad);
    double* sB = sA + S * A_STAGE;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sB + S * B_STAGE);   // full[S], empty[S]

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
            mbar_init(full_s + 8 * s, 1);
            mbar_init(empty_s + 8 * s, NCONS / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
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
        // ================================ TMA producer (one thread) ================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
        if (warp == NCONS / 32 && lane == 0) {
            unsigned g = 0;
#pragma unroll 1
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const unsigned mx = (unsigned)(tile % gm);
                const long long rest = tile / gm;
                const unsigned m0 = mx * BM, n0 = (unsigned)(rest % gn) * BN;
                const int z = (int)(rest / gn);
                const int split = z % p.splits, batch = z / p.splits;
                const int kt_lo = (int)(((long long)p.total_kt * split) / p.splits);
                const int nkt = (int)(((long long)p.total_kt * (split + 1)) / p.splits) - kt_lo;
#pragma unroll 1
                for (int it = 0; it < nkt; ++it, ++g) {
                    const int s = seg_of(kt_lo + it);
                    const int st = g % S;
                    if (g >= S) mbar_wait_relaxed(empty_s + 8 * st, ((g - S) / S) & 1);
                    const int k0 = (kt_lo + it - p.seg[s].kt_begin) * BK;
                    const unsigned bar = full_s + 8 * st;
                    mbar_expect_tx(bar, (unsigned)((A_STAGE + B_STAGE) * sizeof(double)));
                    if (BATCHED) {
                        const int bz = p.seg[s].bz;
                        tma_load_3d(sA_s + st * (A_STAGE * 8), &maps.a[s], k0, (int)m0, (bz & 1) ? batch : 0, bar);
                        tma_load_3d(sB_s + st * (B_STAGE * 8), &maps.b[s], k0, (int)n0, (bz & 2) ? batch : 0, bar);
                    } else {
                        tma_load_2d(sA_s + st * (A_STAGE * 8), &maps.a[s], k0, (int)m0, bar);
                        tma_load_2d(sB_s + st * (B_STAGE * 8), &maps.b[s], k0, (int)n0, bar);
                    }
                }
            }
        }
        return;
    }

    // ================================ consumers ================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;\n");
    const int wm = warp >> 2, wn = warp & 3;
    const int frow = lane >> 2, fk = lane & 3;
    const int pf = ((frow & 3) << 1) | (frow >> 2);          // tile row (within its 8-row group) of fragment row frow
    // offset (doubles) of this lane's element inside a 128-byte row, per k-step: chunk (2ks + fk/2) ^ pf
    int koff[BK / 4];
#pragma unroll
    for (int ks = 0; ks < BK / 4; ++ks) koff[ks] = ((((2 * ks) + (fk >> 1)) ^ pf) << 1) | (fk & 1);
    const int arow = (wm * (8 * MI) + pf) * BK, brow = (wn * 32 + pf) * BK;

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

    double acc[MI][4][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    double a[MI], b[4], bnx[4];
    mbar_wait(full_s + 8 * (g0 % S), (g0 / S) & 1);
    const double* ap = sA + (g0 % S) * A_STAGE + arow;
    const double* bp = sB + (g0 % S) * B_STAGE + brow;
#pragma unroll
    for (int i = 0; i < MI; ++i) a[i] = ap[i * 8 * BK + koff[0]];
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = bp[j * 8 * BK + koff[0]];

    int seg_acc = -1;
    double scale_acc = 1.0;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        const int st = (g0 + it) % S;
        {
            const int sgi = seg_of(kt_lo + it);
            if (sgi != seg_acc) {
                const double sc = p.seg[sgi].scale;
                if (seg_acc >= 0 && sc != scale_acc) {
                    const double r = scale_acc / sc;
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) { acc[i][j][0] *= r; acc[i][j][1] *= r; }
                }
                seg_acc = sgi;
                scale_acc = sc;
            }
        }
        const double* an = ap;
        const double* bn = bp;
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            const int nks = (ks + 1) % (BK / 4);
            bool more = true;
            if (ks == BK / 4 - 1) {
                more = it + 1 < nkt;
                if (more) {
                    const unsigned gn1 = g0 + it + 1;
                    mbar_wait(full_s + 8 * (gn1 % S), (gn1 / S) & 1);
                    an = sA + (gn1 % S) * A_STAGE + arow;
                    bn = sB + (gn1 % S) * B_STAGE + brow;
                }
            }
            if (more) {
#pragma unroll
                for (int j = 0; j < 4; ++j) bnx[j] = bn[j * 8 * BK + koff[nks]];
            }
#pragma unroll
            for (int i = 0; i < MI; ++i) {
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                if (more) a[i] = an[i * 8 * BK + koff[nks]];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bnx[j];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * st);
        ap = an;
        bp = bn;
    }
    g0 += nkt;

    // ---- epilogue: accumulator (frow, 2 fk + e) of an 8x8 block is tile element (perm(frow), perm(2 fk + e))
    const int pc0 = (((2 * fk) & 3) << 1) | ((2 * fk) >> 2), pc1 = (((2 * fk + 1) & 3) << 1) | ((2 * fk + 1) >> 2);
    const unsigned colbase = n0 + wn * 32;
    if (p.splits > 1) {
        double* ws = p.ws + (size_t)z * p.M * p.N;
#pragma unroll
        for (int i = 0; i < MI; ++i) {
            const unsigned row = m0 + wm * (8 * MI) + i * 8 + pf;
            if (row >= p.M) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const unsigned col = colbase + j * 8 + (e ? pc1 : pc0);
                    if (col < p.N) ws[(size_t)row * p.N + col] = scale_acc * acc[i][j][e];
                }
        }
        continue;
    }
    const double alpha = p.alpha * scale_acc;
    long long roff[MI], droff[MI];
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const unsigned row = m0 + wm * (8 * MI) + i * 8 + pf;
        roff[i] = (row < p.M) ? didx_off(p.cm, row) + (long long)batch * p.c_batch : -1;
        droff[i] = (row < p.M && p.D) ? didx_off(p.dm, row) + (long long)batch * p.d_batch : 0;
    }
    store_tile_f64<MI, 1>(p, acc, alpha, roff, droff, colbase, (unsigned)pc0, (unsigned)pc1);
    }   // tile loop
}

// ---- TMA-staged variant with an asynchronous epilogue ------------------------------------------------
// contract_f64_tma_kernel plus the fourth warpgroup of contract_f64_wse_kernel: one TMA producer thread, 8 DMMA
// consumer warps, 4 epilogue warps.  For short contractions over plain matrices with many tiles per SM - the (T) /
// CC3 W build over the helper layouts of triples_engine.py: K = 2(v + o), three launches of which two accumulate
// into W (beta = 1, strided output) - the consumers' own epilogue left the tensor pipe idle 11 % (beta = 0) to
// 22 % (beta = 1) of the launch (15.9 / 18.2 ms per launch at v = 250).  The consumers dump scale * acc into a
// staging tile and start the next tile at once; the epilogue warps apply alpha / beta with their lanes along the
// contiguous output direction.  64 / 80 / 96-row tiles only (the staging tile of 128 rows does not fit next to
// the ring); 4 stages; no split-K, no D operand.
// fire-and-forget FP64 add in L2 (each output element receives one add per launch: the result is deterministic)
__device__ __forceinline__ void red_add_f64(double* addr, double v) {
    asm volatile("red.global.add.f64 [%0], %1;\n" ::"l"(addr), "d"(v) : "memory");
}

constexpr int TMAE_THREADS = NCONS + 128 + NEPI, TMAE_STAGES = 4;

template <int MI, bool BATCHED>
__global__ void __launch_bounds__(TMAE_THREADS, 1)
contract_f64_tmae_kernel(const __grid_constant__ DParams p, const __grid_constant__ TmaMaps maps) {
    constexpr int BM = 16 * MI;
    constexpr int A_STAGE = BM * BK, B_STAGE = BN * BK;       // doubles; rows of 128 bytes
    constexpr int S = TMAE_STAGES;

    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const unsigned raw_s = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned pad = (1024u - (raw_s & 1023u)) & 1023u;
    double* sA = reinterpret_cast<double*>(smem_raw + pad);
    double* sB = sA + S * A_STAGE;
    double* stg = sB + S * B_STAGE;                                                       // [BM][SPITCH]
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(stg + BM * SPITCH);  // full[S], empty[S], staged, drained

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned gm = (p.M + BM - 1) / BM, gn = (p.N + BN - 1) / BN;
    const long long ntiles = (long long)gm * gn * p.nz;       // splits == 1 (host): z is the batch index
    const int nkt = p.total_kt;

    const unsigned sA_s = (unsigned)__cvta_generic_to_shared(sA);
    const unsigned sB_s = (unsigned)__cvta_generic_to_shared(sB);
    const unsigned full_s = (unsigned)__cvta_generic_to_shared(bars);
    const unsigned empty_s = full_s + 8 * S;
    const unsigned staged_s = empty_s + 8 * S, drained_s = staged_s + 8;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            mbar_init(full_s + 8 * s, 1);
            mbar_init(empty_s + 8 * s, NCONS / 32);
        }
        mbar_init(staged_s, NCONS / 32);
        mbar_init(drained_s, NEPI / 32);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    }
    __syncthreads();

    auto seg_of = [&](int kt) {
        int s = 0;
#pragma unroll
        for (int q = 1; q < PCCB_MAX_SEGMENTS; ++q)
            if (q < p.nseg && kt >= p.seg[q].kt_begin) s = q;
        return s;
    };

    // Tile order: groups of gridDim.x consecutive m-tiles outermost; inside a group n, then the batch index, vary
    // while every persistent CTA keeps ITS m-tile.  The A rows of a group (148 x 96 rows x K: tens of MB) stay in
    // L2 - and a CTA's own A tile mostly in its L1 - across the gn * batch tiles that share them.  With m fastest
    // over the whole M (the order of the other kernels) the (T) W build re-streamed its k-independent A operands
    // from DRAM for every (n, k): 16.9 GB read per launch for 0.25 GB of operands.
    const long long per_group = (long long)gridDim.x * gn * p.nz;
    auto decode = [&](long long tile, unsigned& mx, unsigned& nx, int& z) {
        const long long mg = tile / per_group, rem = tile - mg * per_group;
        const long long left = (long long)gm - mg * gridDim.x;
        const long long gsz = left < (long long)gridDim.x ? left : (long long)gridDim.x;
        mx = (unsigned)(mg * gridDim.x + rem % gsz);
        const long long rest = rem / gsz;
        nx = (unsigned)(rest % gn);
        z = (int)(rest / gn);
    };

    if (warp >= (NCONS + 128) / 32) {
        // ================================ epilogue warps ================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 104;\n");     // 256 x 184 + 128 x 40 + 128 x 104 = 64 K registers
        const int ew = (tid - (NCONS + 128)) >> 5;
        // The FP64 pipe belongs to the consumers' DMMAs (94 % busy): an epilogue warp that has to squeeze a DMUL /
        // DFMA per element into its gaps drains a tile more slowly than the consumers produce one (beta = 1 launches
        // ran at 79 % tensor activity).  So alpha is folded into the staged values by the consumers, beta = 0 is a plain
        // copy and beta = 1 an L2 add (red.global.add.f64); only a general beta does arithmetic here.
        const int mode = p.beta == 0.0 ? 0 : (p.beta == 1.0 ? 1 : 2);
        const bool along_m = p.cm.s_lo == 1 && p.cn.s_lo != 1;     // lanes along the contiguous output direction
        const double beta = p.beta;
        unsigned lt = 0;
#pragma unroll 1
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            unsigned mx, nx;
            int z;
            decode(tile, mx, nx, z);
            const unsigned m0 = mx * BM, n0 = nx * BN;
            const int mlim = (int)min((unsigned)BM, p.M - m0), nlim = (int)min((unsigned)BN, p.N - n0);
            double* Cb = p.C + (long long)z * p.c_batch;
            mbar_wait(staged_s, lt & 1);
            if (!along_m) {
                constexpr int CP = BN / 32, RU = 4;
                long long co[CP];
#pragma unroll
                for (int q = 0; q < CP; ++q) {
                    const int c = q * 32 + lane;
                    co[q] = c < nlim ? didx_off(p.cn, n0 + c) : -1;
                }
                const double* sl = stg + lane;
#pragma unroll 1
                for (int r0 = ew; r0 < mlim; r0 += 4 * RU) {
                    double v[RU][CP];
                    double* crow[RU];
#pragma unroll
                    for (int u = 0; u < RU; ++u) {
                        const int r = r0 + 4 * u;
                        crow[u] = nullptr;
                        if (r < mlim) {
                            crow[u] = Cb + didx_off(p.cm, m0 + r);
#pragma unroll
                            for (int q = 0; q < CP; ++q) {
                                double x = sl[r * SPITCH + q * 32];
                                if (mode == 2 && co[q] >= 0) x += beta * crow[u][co[q]];
                                v[u][q] = x;
                            }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < RU; ++u)
                        if (crow[u]) {
#pragma unroll
                            for (int q = 0; q < CP; ++q)
                                if (co[q] >= 0) {
                                    if (mode == 1) red_add_f64(crow[u] + co[q], v[u][q]);
                                    else crow[u][co[q]] = v[u][q];
                                }
                        }
                }
            } else {
                constexpr int RP = (BM + 31) / 32, CU = 4;
                long long ro[RP];
#pragma unroll
                for (int q = 0; q < RP; ++q) {
                    const int r = q * 32 + lane;
                    ro[q] = r < mlim ? didx_off(p.cm, m0 + r) : -1;
                }
                const double* sl = stg + lane * SPITCH;
#pragma unroll 1
                for (int c0 = ew; c0 < nlim; c0 += 4 * CU) {
                    double v[CU][RP];
                    double* ccol[CU];
#pragma unroll
                    for (int u = 0; u < CU; ++u) {
                        const int c = c0 + 4 * u;
                        ccol[u] = nullptr;
                        if (c < nlim) {
                            ccol[u] = Cb + didx_off(p.cn, n0 + c);
#pragma unroll
                            for (int q = 0; q < RP; ++q) {
                                double x = 0.0;
                                if (ro[q] >= 0) {
                                    x = sl[q * 32 * SPITCH + c];
                                    if (mode == 2) x += beta * ccol[u][ro[q]];
                                }
                                v[u][q] = x;
                            }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < CU; ++u)
                        if (ccol[u]) {
#pragma unroll
                            for (int q = 0; q < RP; ++q)
                                if (ro[q] >= 0) {
                                    if (mode == 1) red_add_f64(ccol[u] + ro[q], v[u][q]);
                                    else ccol[u][ro[q]] = v[u][q];
                                }
                        }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(drained_s);
        }
        return;
    }

    if (warp >= NCONS / 32) {
        // ================================ TMA producer (one thread) ================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
        if (warp == NCONS / 32 && lane == 0) {
            unsigned g = 0;
#pragma unroll 1
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                unsigned mx, nx;
                int batch;
                decode(tile, mx, nx, batch);
                const unsigned m0 = mx * BM, n0 = nx * BN;
#pragma unroll 1
                for (int it = 0; it < nkt; ++it, ++g) {
                    const int s = seg_of(it);
                    const int st = g % S;
                    if (g >= S) mbar_wait_relaxed(empty_s + 8 * st, ((g - S) / S) & 1);
                    const int k0 = (it - p.seg[s].kt_begin) * BK;
                    const unsigned bar = full_s + 8 * st;
                    mbar_expect_tx(bar, (unsigned)((A_STAGE + B_STAGE) * sizeof(double)));
                    if (BATCHED) {
                        const int bz = p.seg[s].bz;
                        tma_load_3d(sA_s + st * (A_STAGE * 8), &maps.a[s], k0, (int)m0, (bz & 1) ? batch : 0, bar);
                        tma_load_3d(sB_s + st * (B_STAGE * 8), &maps.b[s], k0, (int)n0, (bz & 2) ? batch : 0, bar);
                    } else {
                        tma_load_2d(sA_s + st * (A_STAGE * 8), &maps.a[s], k0, (int)m0, bar);
                        tma_load_2d(sB_s + st * (B_STAGE * 8), &maps.b[s], k0, (int)n0, bar);
                    }
                }
            }
        }
        return;
    }

    // ================================ consumers ================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 184;\n");
    const int wm = warp >> 2, wn = warp & 3;
    const int frow = lane >> 2, fk = lane & 3;
    const int pf = ((frow & 3) << 1) | (frow >> 2);          // tile row (within its 8-row group) of fragment row frow
    int koff[BK / 4];
#pragma unroll
    for (int ks = 0; ks < BK / 4; ++ks) koff[ks] = ((((2 * ks) + (fk >> 1)) ^ pf) << 1) | (fk & 1);
    const int arow = (wm * (8 * MI) + pf) * BK, brow = (wn * 32 + pf) * BK;
    // accumulator (frow, 2 fk + e) of an 8x8 block is tile element (perm(frow), perm(2 fk + e))
    const int pc0 = (((2 * fk) & 3) << 1) | ((2 * fk) >> 2), pc1 = (((2 * fk + 1) & 3) << 1) | ((2 * fk + 1) >> 2);

    unsigned g0 = 0, lt = 0;
#pragma unroll 1
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
    double acc[MI][4][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    double a[MI], b[4], bnx[4];
    mbar_wait(full_s + 8 * (g0 % S), (g0 / S) & 1);
    const double* ap = sA + (g0 % S) * A_STAGE + arow;
    const double* bp = sB + (g0 % S) * B_STAGE + brow;
#pragma unroll
    for (int i = 0; i < MI; ++i) a[i] = ap[i * 8 * BK + koff[0]];
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = bp[j * 8 * BK + koff[0]];

    int seg_acc = -1;
    double scale_acc = 1.0;
#pragma unroll 1
    for (int it = 0; it < nkt; ++it) {
        const int st = (g0 + it) % S;
        {
            const int sgi = seg_of(it);
            if (sgi != seg_acc) {
                const double sc = p.seg[sgi].scale;
                if (seg_acc >= 0 && sc != scale_acc) {
                    const double r = scale_acc / sc;
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) { acc[i][j][0] *= r; acc[i][j][1] *= r; }
                }
                seg_acc = sgi;
                scale_acc = sc;
            }
        }
        const double* an = ap;
        const double* bn = bp;
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            const int nks = (ks + 1) % (BK / 4);
            // after the last k-tile the "next" fragments are re-read from the current stage (valid shared memory,
            // values unused): the fragment loads stay unconditional, so ptxas has no reason to predicate /
            // duplicate the DMMAs around them (it did: 116 DMMAs per k-tile instead of 96, +50 % time)
            if (ks == BK / 4 - 1 && it + 1 < nkt) {
                const unsigned gn1 = g0 + it + 1;
                mbar_wait(full_s + 8 * (gn1 % S), (gn1 / S) & 1);
                an = sA + (gn1 % S) * A_STAGE + arow;
                bn = sB + (gn1 % S) * B_STAGE + brow;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) bnx[j] = bn[j * 8 * BK + koff[nks]];
#pragma unroll
            for (int i = 0; i < MI; ++i) {
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                a[i] = an[i * 8 * BK + koff[nks]];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bnx[j];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_s + 8 * st);
        ap = an;
        bp = bn;
    }
    g0 += nkt;

    // ---- hand the tile to the epilogue warps -------------------------------------------------
    if (lt > 0) mbar_wait(drained_s, (lt - 1) & 1);      // the previous tile has left the staging buffer
    {
        const double stage_scale = p.alpha * scale_acc;      // alpha applied here, on the consumers' own pipe slots
        double* sp = stg + (wm * (8 * MI) + pf) * SPITCH + wn * 32;
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                sp[i * 8 * SPITCH + j * 8 + pc0] = stage_scale * acc[i][j][0];
                sp[i * 8 * SPITCH + j * 8 + pc1] = stage_scale * acc[i][j][1];
            }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(staged_s);
    }   // tile loop
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

template <int MI>
constexpr size_t smem_bytes() {
    constexpr int BM = 16 * MI;
    constexpr int A_STAGE = (BM * PK > BK * (BM + 4)) ? BM * PK : BK * (BM + 4);
    constexpr int B_STAGE = (BN * PK > BK * (BN + 4)) ? BN * PK : BK * (BN + 4);
    return (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double) + (BM + BN) * sizeof(long long) +
           2 * STAGES * sizeof(unsigned long long) + 16;
}

template <int MI>
constexpr size_t smem_bytes_ws() {
    constexpr int BM = 16 * MI;
    constexpr int A_STAGE = (BM * PK > BK * (BM + 4)) ? BM * PK : BK * (BM + 4);
    constexpr int B_STAGE = (BN * PK > BK * (BN + 4)) ? BN * PK : BK * (BN + 4);
    constexpr int S = (MI <= 2) ? WS_STAGES_SKINNY : WS_STAGES;
    return (size_t)S * (A_STAGE + B_STAGE) * sizeof(double) + (BM + BN) * sizeof(long long) +
           2 * S * sizeof(unsigned long long) + 16;
}

template <int MI>
constexpr size_t smem_bytes_wse() {
    constexpr int BM = 16 * MI;
    constexpr int A_STAGE = (BM * PK > BK * (BM + 4)) ? BM * PK : BK * (BM + 4);
    constexpr int B_STAGE = (BN * PK > BK * (BN + 4)) ? BN * PK : BK * (BN + 4);
    return ((size_t)wse_stages<MI>() * (A_STAGE + B_STAGE) + (size_t)BM * SPITCH) * sizeof(double) +
           (BM + BN) * sizeof(long long) + (2 * wse_stages<MI>() + 2) * sizeof(unsigned long long) + 16;
}

bool even(long long x) { return (x & 1) == 0; }

// Shared-memory layout of one operand of one segment and whether 16-byte copies are legal.
//   mn: index walking the tile rows (M or N), k: contraction index.
void choose_layout(const void* base, const DIdx& mn, const DIdx& k, long long batch_stride, long long nbatch,
                   int* mn_major, bool* vec2_ok) {
    const bool base16 = (reinterpret_cast<uintptr_t>(base) & 15) == 0;
    const bool batch_ok = nbatch <= 1 || even(batch_stride);
    if (k.s_lo == 1) {   // K-contiguous: tile[row][k]
        *mn_major = 0;
        *vec2_ok = base16 && batch_ok && (k.flat || (even(k.div) && even(k.s_hi))) && (mn.flat || even(mn.s_hi)) &&
                   even(mn.s_lo);
        return;
    }
    if (mn.s_lo == 1) {  // row-contiguous: tile[k][row]
        *mn_major = 1;
        *vec2_ok = base16 && batch_ok && (mn.flat || (even(mn.div) && even(mn.s_hi))) && (k.flat || even(k.s_hi)) &&
                   even(k.s_lo);
        return;
    }
    *mn_major = 0;       // no contiguous direction: element-wise gather
    *vec2_ok = false;
}

template <int VEC, int MI>
int launch_tile_kernel(pccb_ctx* ctx, const DParams& p, long long batch) {
    constexpr int BM = 16 * MI;
    constexpr size_t SMEM = smem_bytes<MI>();
    static bool attr_set = false;
    if (!attr_set) {
        PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f64_kernel<VEC, MI>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
        attr_set = true;
    }
    dim3 grid((unsigned)((p.M + BM - 1) / BM), (unsigned)((p.N + BN - 1) / BN), (unsigned)(batch * p.splits));
    // tuning hook: PCCB_WS=0 selects the non-specialised kernel (all 8 warps load and compute)
    static const bool use_ws = [] { const char* e = getenv("PCCB_WS"); return e ? atoi(e) != 0 : true; }();
    // short contractions with many tiles per CTA: the variant with epilogue warps (PCCB_WSE=0 disables)
    static const bool use_wse = [] { const char* e = getenv("PCCB_WSE"); return e ? atoi(e) != 0 : true; }();
    if constexpr (MI >= 4 && MI <= 6) {
        const long long ntiles = (long long)grid.x * grid.y * grid.z;
        if (use_ws && use_wse && p.splits == 1 && !p.D && p.total_kt >= WSE_MIN_KT && p.total_kt <= WSE_MAX_KT && ntiles >= 2ll * ctx->sm_count) {
            constexpr size_t SMEM_E = smem_bytes_wse<MI>();
            static bool attr_wse = false;
            if (!attr_wse) {
                PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f64_wse_kernel<VEC, MI>,
                                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_E));
                attr_wse = true;
            }
            DParams q = p;
            q.nz = (int)batch;
            contract_f64_wse_kernel<VEC, MI><<<(unsigned)ctx->sm_count, NTHREADS_WSE, SMEM_E, ctx->stream>>>(q);
            PCCB_CHECK_LAUNCH(ctx, "contract_f64_wse_kernel");
            return PCCB_OK;
        }
    }
    // skinny tiles (MI <= 2) normally run on the 2-CTA/SM kernel; a split-K reduction over a huge K with
    // contiguous ranges (Fmi: 20 x 20 output, K = o v^2) is faster on the persistent kernel with its
    // dedicated producer warps and an 8-stage ring (0.26 -> 0.15 ms).  Routing the other skinny launches
    // (t1 x <mb|ef>: M = o padded to 32 rows, about as much DMMA time as HBM time) there was measured
    // and is not faster (0.30 vs 0.29 ms).
    const bool skinny_ws = MI <= 2 && p.kchunk == 0 && p.splits > 1;
    if (use_ws && (MI > 2 || skinny_ws)) {
        constexpr size_t SMEM_WS = smem_bytes_ws<MI>();
        static bool attr_ws = false;
        if (!attr_ws) {
            PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f64_ws_kernel<VEC, MI>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_WS));
            attr_ws = true;
        }
        DParams q = p;
        q.nz = (int)(batch * p.splits);
        const long long ntiles = (long long)grid.x * grid.y * grid.z;
        const unsigned nctas = (unsigned)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
        contract_f64_ws_kernel<VEC, MI><<<nctas, NTHREADS_WS, SMEM_WS, ctx->stream>>>(q);
        PCCB_CHECK_LAUNCH(ctx, "contract_f64_ws_kernel");
        return PCCB_OK;
    }
    contract_f64_kernel<VEC, MI><<<grid, NTHREADS, SMEM, ctx->stream>>>(p);
    PCCB_CHECK_LAUNCH(ctx, "contract_f64_kernel");
    return PCCB_OK;
}

template <int VEC>
int dispatch_mi(pccb_ctx* ctx, const DParams& p, long long batch, int mi) {
    switch (mi) {
        case 1: return launch_tile_kernel<VEC, 1>(ctx, p, batch);
        case 2: return launch_tile_kernel<VEC, 2>(ctx, p, batch);
        case 3: return launch_tile_kernel<VEC, 3>(ctx, p, batch);
        case 4: return launch_tile_kernel<VEC, 4>(ctx, p, batch);
        case 5: return launch_tile_kernel<VEC, 5>(ctx, p, batch);
        case 6: return launch_tile_kernel<VEC, 6>(ctx, p, batch);
        default: return launch_tile_kernel<VEC, 8>(ctx, p, batch);
    }
}

// ---- TMA path: eligibility, tensor maps, launch --------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// operand = K-contiguous matrix with one uniform, 16-byte-aligned row stride
// (a two-level index whose levels happen to be contiguous, s_hi == div * s_lo, is a plain dimension too:
//  the planner emits one div per index group)
bool plain_dim(const DIdx& c) { return c.flat || c.s_hi == (long long)c.div * c.s_lo; }

bool tma_operand_ok(const void* base, const DIdx& mn, const DIdx& k, long long rows) {
    if (!(plain_dim(k) && k.s_lo == 1 && plain_dim(mn))) return false;
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return false;
    if (rows > 1 && (mn.s_lo <= 0 || (mn.s_lo & 1))) return false;     // stride in doubles, must be even
    return true;
}

bool make_map(CUtensorMap* m, const void* base, long long rows, long long K, long long row_stride, int box_rows) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    if (rows <= 1) row_stride = ((K + 1) / 2) * 2;       // a single row: any legal stride
    if (row_stride < K) return false;                     // overlapping rows are not a tensor
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_stride * sizeof(double)};
    const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<void*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// batched operand: (k, row, batch); an operand without a batch stride gets a batch dimension of extent 1
bool make_map3(CUtensorMap* m, const void* base, long long rows, long long K, long long row_stride, int box_rows,
               long long nbatch, long long batch_stride) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    if (rows <= 1) row_stride = ((K + 1) / 2) * 2;
    if (row_stride < K) return false;
    if (batch_stride <= 0) { nbatch = 1; batch_stride = row_stride * (rows > 0 ? rows : 1); }
    if (batch_stride & 1) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)rows, (cuuint64_t)nbatch};
    const cuuint64_t strides[2] = {(cuuint64_t)row_stride * sizeof(double), (cuuint64_t)batch_stride * sizeof(double)};
    const cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<void*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int MI>
constexpr size_t smem_bytes_tma() {
    return (size_t)TMA_STAGES * (16 * MI + BN) * BK * sizeof(double) + 2 * TMA_STAGES * sizeof(unsigned long long) + 1024;
}

template <int MI, bool BATCHED>
int launch_tma_kernel(pccb_ctx* ctx, const DParams& p, const TmaMaps& maps, long long batch) {
    constexpr int BM = 16 * MI;
    constexpr size_t SMEM = smem_bytes_tma<MI>();
    static bool attr_set = false;
    if (!attr_set) {
        PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f64_tma_kernel<MI, BATCHED>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)SMEM));
        attr_set = true;
    }
    DParams q = p;
    q.nz = (int)(batch * p.splits);
    const long long ntiles = (long long)((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN) * q.nz;
    const unsigned nctas = (unsigned)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
    contract_f64_tma_kernel<MI, BATCHED><<<nctas, TMA_THREADS, SMEM, ctx->stream>>>(q, maps);
    PCCB_CHECK_LAUNCH(ctx, "contract_f64_tma_kernel");
    return PCCB_OK;
}

int choose_mi_epi(long long M);

template <int MI>
constexpr size_t smem_bytes_tmae() {
    return (size_t)TMAE_STAGES * (16 * MI + BN) * BK * sizeof(double) + (size_t)(16 * MI) * SPITCH * sizeof(double) +
           (2 * TMAE_STAGES + 2) * sizeof(unsigned long long) + 1024;
}

template <int MI, bool BATCHED>
int launch_tmae_kernel(pccb_ctx* ctx, const DParams& p, const TmaMaps& maps, long long batch) {
    constexpr int BM = 16 * MI;
    constexpr size_t SMEM = smem_bytes_tmae<MI>();
    static_assert(SMEM <= 232448, "staging tile + ring exceed the shared memory of an SM");
    static bool attr_set = false;
    if (!attr_set) {
        PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_f64_tmae_kernel<MI, BATCHED>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)SMEM));
        attr_set = true;
    }
    DParams q = p;
    q.nz = (int)batch;
    const long long ntiles = (long long)((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN) * q.nz;
    const unsigned nctas = (unsigned)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
    contract_f64_tmae_kernel<MI, BATCHED><<<nctas, TMAE_THREADS, SMEM, ctx->stream>>>(q, maps);
    PCCB_CHECK_LAUNCH(ctx, "contract_f64_tmae_kernel");
    return PCCB_OK;
}

// returns 1 when the TMA kernel was launched, 0 when the contraction is not eligible, < 0 on error
int try_tma(pccb_ctx* ctx, const DParams& p_in, const pccb_contraction* d, int mi) {
    static const bool use_tma = [] { const char* e = getenv("PCCB_TMA"); return e ? atoi(e) != 0 : true; }();
    static const bool use_tma_batched = [] { const char* e = getenv("PCCB_TMA_BATCHED"); return e ? atoi(e) != 0 : true; }();
    if (!use_tma || mi <= 2 || p_in.nseg < 1) return 0;
    const bool batched = d->batch != 1;
    if (batched && (!use_tma_batched || p_in.splits != 1)) return 0;
    DParams p = p_in;
    for (int s = 0; s < p.nseg; ++s) {
        DSeg& g = p.seg[s];
        if (!tma_operand_ok(g.A, g.am, g.ak, p.M) || !tma_operand_ok(g.B, g.bn, g.bk, p.N)) return 0;
        if (batched && (g.a_batch < 0 || g.b_batch < 0 || (g.a_batch & 1) || (g.b_batch & 1))) return 0;
        g.bz = (g.a_batch > 0 ? 1 : 0) | (g.b_batch > 0 ? 2 : 0);
    }
    // short contractions (8..64 k-tiles) with many tiles per SM: the variant with epilogue warps and the
    // L2-friendly tile order, on the best of the 64 / 80 / 96-row tiles (PCCB_TMAE=0 disables)
    static const bool use_tmae = [] { const char* e = getenv("PCCB_TMAE"); return e ? atoi(e) != 0 : true; }();
    bool epi = false;
    if (use_tmae && mi >= 4 && p.splits == 1 && !p.D && p.total_kt >= WSE_MIN_KT && p.total_kt <= WSE_MAX_KT) {
        const int best = choose_mi_epi(p.M);
        const long long t2 = (((long long)p.M + 16 * best - 1) / (16 * best)) * (((long long)p.N + BN - 1) / BN) * d->batch;
        if (t2 >= 2ll * ctx->sm_count) { mi = best; epi = true; }
    }
    TmaMaps maps;
    memset(&maps, 0, sizeof(maps));
    for (int s = 0; s < p.nseg; ++s) {
        const DSeg& g = p.seg[s];
        if (batched) {
            if (!make_map3(&maps.a[s], g.A, p.M, g.K, g.am.s_lo, 16 * mi, d->batch, g.a_batch)) return 0;
            if (!make_map3(&maps.b[s], g.B, p.N, g.K, g.bn.s_lo, BN, d->batch, g.b_batch)) return 0;
        } else {
            if (!make_map(&maps.a[s], g.A, p.M, g.K, g.am.s_lo, 16 * mi)) return 0;
            if (!make_map(&maps.b[s], g.B, p.N, g.K, g.bn.s_lo, BN)) return 0;
        }
    }
    int rc;
    if (epi) {
        if (batched) {
            switch (mi) {
                case 4: rc = launch_tmae_kernel<4, true>(ctx, p, maps, d->batch); break;
                case 5: rc = launch_tmae_kernel<5, true>(ctx, p, maps, d->batch); break;
                default: rc = launch_tmae_kernel<6, true>(ctx, p, maps, d->batch); break;
            }
        } else {
            switch (mi) {
                case 4: rc = launch_tmae_kernel<4, false>(ctx, p, maps, d->batch); break;
                case 5: rc = launch_tmae_kernel<5, false>(ctx, p, maps, d->batch); break;
                default: rc = launch_tmae_kernel<6, false>(ctx, p, maps, d->batch); break;
            }
        }
    } else if (batched) {
        switch (mi) {
            case 3: rc = launch_tma_kernel<3, true>(ctx, p, maps, d->batch); break;
            case 4: rc = launch_tma_kernel<4, true>(ctx, p, maps, d->batch); break;
            case 5: rc = launch_tma_kernel<5, true>(ctx, p, maps, d->batch); break;
            case 6: rc = launch_tma_kernel<6, true>(ctx, p, maps, d->batch); break;
            default: rc = launch_tma_kernel<8, true>(ctx, p, maps, d->batch); break;
        }
    } else {
        switch (mi) {
            case 3: rc = launch_tma_kernel<3, false>(ctx, p, maps, d->batch); break;
            case 4: rc = launch_tma_kernel<4, false>(ctx, p, maps, d->batch); break;
            case 5: rc = launch_tma_kernel<5, false>(ctx, p, maps, d->batch); break;
            case 6: rc = launch_tma_kernel<6, false>(ctx, p, maps, d->batch); break;
            default: rc = launch_tma_kernel<8, false>(ctx, p, maps, d->batch); break;
        }
    }
    return rc == PCCB_OK ? 1 : rc;
}

// tile height for the kernels with epilogue warps (64 / 80 / 96 rows): padded rows weighted like choose_mi
int choose_mi_epi(long long M) {
    // tuning hook: PCCB_EPI_MI=<4|5|6> forces the height
    static const int forced = [] { const char* e = getenv("PCCB_EPI_MI"); return e ? atoi(e) : 0; }();
    if (forced >= 4 && forced <= 6) return forced;
    static const int cand[] = {6, 5, 4};
    static const double eff[7] = {0, 0, 0, 0, 0.97, 0.985, 0.99};
    int best = 6;
    double best_cost = 1e300;
    for (int c : cand) {
        const long long b = 16 * c;
        const double cost = (double)(((M + b - 1) / b) * b) / eff[c];
        if (cost < best_cost) { best_cost = cost; best = c; }
    }
    return best;
}

// rows per CTA tile: minimise padded rows, prefer the taller tile on ties (more operand reuse)
int choose_mi(long long M) {
    // tuning hook for tools/bench_contract.py: PCCB_MI=<1|2|4|5|6|8> forces the tile height
    static const int forced = [] { const char* e = getenv("PCCB_MI"); return e ? atoi(e) : 0; }();
    if (forced == 1 || forced == 2 || forced == 3 || forced == 4 || forced == 5 || forced == 6 || forced == 8) return forced;
    // padded rows weighted by the measured relative rate of the tile height.  Per useful row the
    // 64...128-row tiles run at nearly the same rate (ladder, M = 1600: 13 x 128 rows 33.1 TFLOP/s,
    // 20 x 80 rows 35.2; M = 900: 10 x 96 32.5, 15 x 64 32.4, 8 x 128 29.8), so among them padding
    // decides; the 16- and 32-row tiles have a quarter / half of the DMMAs per fragment load and run on
    // the non-persistent kernel: only for outputs that are skinny anyway.
    // (48-row tiles: M = o^2 = 900 pads to 912 rows instead of 960 - the ladder at o = 30; weighted a little below
    //  the 64-row tile, so that it is taken only when it saves more than ~1.5 % of padded rows)
    static const int cand[] = {8, 6, 5, 4, 3, 2, 1};
    static const double eff[9] = {0.0, 0.35, 0.60, 0.955, 0.97, 0.985, 0.99, 0.0, 1.00};
    int best = 8;
    double best_cost = 1e300;
    for (int mi : cand) {
        const long long bm = 16 * mi;
        const double cost = (double)(((M + bm - 1) / bm) * bm) / eff[mi];
        if (cost < best_cost) { best_cost = cost; best = mi; }
    }
    return best;
}

}  // namespace

extern "C" int pccb_contract(pccb_ctx* ctx, const pccb_contraction* d) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, d != nullptr, "pccb_contract: null descriptor");
    PCCB_REQUIRE(ctx, d->dtype == PCCB_F64 || d->dtype == PCCB_F32, "pccb_contract: unknown dtype %d", d->dtype);
    PCCB_REQUIRE(ctx, d->nseg >= 0 && d->nseg <= PCCB_MAX_SEGMENTS, "pccb_contract: nseg=%d out of range", d->nseg);
    PCCB_REQUIRE(ctx, d->M > 0 && d->N > 0 && d->batch > 0, "pccb_contract: empty output (M=%lld N=%lld batch=%lld)",
                 (long long)d->M, (long long)d->N, (long long)d->batch);
    PCCB_REQUIRE(ctx, d->M < (1ll << 31) && d->N < (1ll << 31), "pccb_contract: M/N too large");
    PCCB_REQUIRE(ctx, d->C_dev != nullptr, "pccb_contract: C_dev is null");
    if (d->dtype == PCCB_F32) return pccb_contract_f32(ctx, d);   // 3xTF32 tensor-core kernel (contract_f32.cu)

    DParams p;
    memset(&p, 0, sizeof(p));
    p.M = (unsigned)d->M;
    p.N = (unsigned)d->N;
    int kt = 0, ns = 0;
    bool vec2 = true;
    for (int s = 0; s < d->nseg; ++s) {
        const pccb_segment& g = d->seg[s];
        PCCB_REQUIRE(ctx, g.K >= 0 && g.K < (1ll << 31), "pccb_contract: segment %d K out of range", s);
        if (g.K == 0 || g.scale == 0.0) continue;
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
        bool va, vb;
        choose_layout(o.A, o.am, o.ak, o.a_batch, d->batch, &o.a_mn_major, &va);
        choose_layout(o.B, o.bn, o.bk, o.b_batch, d->batch, &o.b_mn_major, &vb);
        vec2 = vec2 && va && vb;
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
        // ns == 0: pure epilogue (beta*C + gamma*D)
        const int threads = 256;
        long long blocks = (total + threads - 1) / threads;
        if (blocks > 148 * 64) blocks = 148 * 64;
        contract_f64_naive_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(p, total);
        PCCB_CHECK_LAUNCH(ctx, "contract_f64_naive_kernel");
        return PCCB_OK;
    }

    int mi = choose_mi(d->M);
    const long long bm = 16 * mi;
    const long long gm = (d->M + bm - 1) / bm, gn = (d->N + BN - 1) / BN;
    PCCB_REQUIRE(ctx, gn <= 65535, "pccb_contract: N=%lld needs more than 65535 column tiles", (long long)d->N);
    const long long tiles = gm * gn * d->batch;
    int splits = 1;
    if (tiles < 2ll * ctx->sm_count && kt >= 16) {
        // deterministic split-K for outputs with too few tiles: among the admissible split counts pick
        // the one whose tiles*splits fills whole waves of the persistent grid best (ties: fewer splits)
        const long long by_k = kt / 8;                       // at least 8 k-tiles (128 k) per split
        const long long by_ws = ctx->ws ? (long long)(ctx->ws_bytes / ((size_t)total * sizeof(double))) : 1;
        long long smax = (8ll * ctx->sm_count + tiles - 1) / tiles;
        if (smax > by_k) smax = by_k;
        if (smax > by_ws) smax = by_ws;
        double best = 0.0;
        for (long long sK = 1; sK <= smax; ++sK) {
            const long long work = tiles * sK, waves = (work + ctx->sm_count - 1) / ctx->sm_count;
            double eff = (double)work / (double)(waves * ctx->sm_count);
            if (waves < 2) eff *= 0.97;                      // a single wave has no tail overlap at all
            eff -= 0.004 * (double)sK;                       // reduction traffic grows with the split count
            if (eff > best + 1e-9) { best = eff; splits = (int)sK; }
        }
    }
    else if (tiles < 24ll * ctx->sm_count && kt >= 256 && d->batch == 1 && ctx->ws) {
        // A handful of waves of LONG tiles (the ladder of one rank out of eight: 1197 tiles of 48 x 128 x 62500 = 8.09
        // waves, the ninth wave 9 % full): split K so that the tail wave is a fraction of a tile long.  Cost in
        // full-K tile times: ceil(tiles s / SMs) / s plus the workspace traffic of the deterministic reduction.
        static const bool tail_split = [] { const char* e = getenv("PCCB_TAIL_SPLIT"); return e ? atoi(e) != 0 : true; }();
        const double t_tile = 2.0 * (double)bm * BN * (16.0 * kt) / (37.0e12 / ctx->sm_count);
        const long long by_ws = (long long)(ctx->ws_bytes / ((size_t)total * sizeof(double)));
        long long smax = kt / 64 < 8 ? kt / 64 : 8;
        if (smax > by_ws) smax = by_ws;
        double best = (double)((tiles + ctx->sm_count - 1) / ctx->sm_count);
        const double whole = best;
        int pick = 1;
        for (long long sK = 2; tail_split && sK <= smax; ++sK) {
            const long long waves = (tiles * sK + ctx->sm_count - 1) / ctx->sm_count;
            const double cost = (double)waves / (double)sK + (2.0 * sK + 1.0) * (double)total * 8.0 / 5.0e12 / t_tile;
            if (cost < best - 1e-9) { best = cost; pick = (int)sK; }
        }
        if (best < 0.98 * whole) splits = pick;
    }
    PCCB_REQUIRE(ctx, (long long)d->batch * splits <= 65535, "pccb_contract: batch*splits=%lld exceeds grid.z",
                 (long long)d->batch * splits);
    p.splits = splits;
    // round-robin chunks pay off for a few skinny tiles with many operand rows and a huge K when those
    // rows lie far apart (>= one 2 MB page each: Lambda's l2[imef] Hs[amef], rows 8 o v^2 bytes apart:
    // 2.2 -> 0.8 ms at (20,150)).  With rows close together (r_T1's t2s[mief] <ma|ef>: 8 v^2 bytes) the
    // contiguous ranges of the persistent kernel are faster (0.66 -> 0.32 ms), and so they are with few rows.
    {
        const long long rows = (d->M < bm ? d->M : bm) + (d->N < BN ? d->N : BN);
        long long kmax = 0, row_stride = 0;
        for (int s = 0; s < ns; ++s)
            if ((long long)p.seg[s].K > kmax) { kmax = p.seg[s].K; row_stride = p.seg[s].bn.s_lo; }
        static const bool no_kchunk = [] { const char* e = getenv("PCCB_NO_KCHUNK"); return e ? atoi(e) != 0 : false; }();
        p.kchunk = (!no_kchunk && splits > 1 && mi <= 2 && tiles <= 4 && rows >= 96 && row_stride * 8 >= (2ll << 20)) ? 8 : 0;
    }

    int rc = try_tma(ctx, p, d, mi);      // plain 2-D operands: TMA-staged kernel
    if (rc < 0) return rc;
    if (rc == 0) {
        // Short contractions (8..64 k-tiles) with many tiles per SM - the (T) / CC3 W build, M = v^2 - spend
        // 12-20 % of a tile in the epilogue; the kernel with epilogue warps exists for 64/80/96-row tiles only, and
        // choose_mi prefers 128 rows for a large M (v = 250: 128-row tiles ran at 0.88 / 0.79 of the DMMA rate with
        // beta = 0 / 1, profiles/r02_ncu_triples_w_o30v250.txt).  Take the best of the three heights instead.
        static const bool wse_mi = [] { const char* e = getenv("PCCB_WSE_MI"); return e ? atoi(e) != 0 : true; }();
        if (wse_mi && (mi == 8 || mi == 3) && splits == 1 && !p.D && kt >= WSE_MIN_KT && kt <= WSE_MAX_KT) {
            const int best = choose_mi_epi(d->M);
            const long long t2 = ((d->M + 16 * best - 1) / (16 * best)) * gn * d->batch;
            if (t2 >= 2ll * ctx->sm_count) mi = best;
        }
        rc = vec2 ? dispatch_mi<2>(ctx, p, d->batch, mi) : dispatch_mi<1>(ctx, p, d->batch, mi);
        if (rc != PCCB_OK) return rc;
    }
    if (splits > 1) {
        const int threads = 256;
        long long blocks = (total + threads - 1) / threads;
        if (blocks > 148 * 32) blocks = 148 * 32;
        splitk_reduce_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(p, total);
        PCCB_CHECK_LAUNCH(ctx, "splitk_reduce_kernel");
    }
    return PCCB_OK;
}
