// pccb_contract, FP32 operands, 5th-generation tensor cores (tcgen05 + TMEM) - the path of the
// big plain-matrix contractions of the mixed-precision mode (precision='SP'): the o^2 v^4 ladder
// tau[ijef] <ab|ef>, Zmbij = <mb|ef> tau[ijef], Wmnij = <mn|ef> tau[ijef]  (pycc/ccwfn.py:473-482,
// 537, 607 with pycc's precision='SP', ccwfn.py:189-196).
//
// Math: 3xTF32.  tcgen05.mma kind::tf32 reads FP32 words from shared memory and IGNORES the 13
// low mantissa bits (measured: tools/tc05_test.cu, (1 + 2^-11 + 2^-12) * 1 -> 1.0), so an FP32
// array is its own "hi" operand.  The caller registers, per operand tensor, a second array
// lo = x - trunc_tf32(x) (pccb_split_lo / pccb_register_lo_plane); a product is then
//     A B  ~=  A_hi B_hi + A_lo B_hi + A_hi B_lo
// with all three products accumulated into ONE FP32 accumulator tile in tensor memory (default:
// every stage holds A, A_lo, B, B_lo of a k-block and feeds the three products; PCCB_TC05_FUSED=0
// selects three K-concatenated passes instead).  Same accuracy model as the mma.sync kernel of contract_f32.cu (the
// lo*lo term, 2^-22 relative, is dropped), at the tcgen05 rate: a single TF32 pass runs at
// ~800 TFLOP/s on the ladder shape (tools/tc05_test.cu), the mma.sync kernel tops out at 92/3.
//
// Kernel (one 128 x 256 output tile per CTA, 192 threads):
//   warp 0, one lane : TMA producer - cp.async.bulk.tensor.2d, 128-byte swizzle, 2-stage ring of
//                      2 x (128 + 256) rows x 32 floats, completion counted in bytes on the stage's mbarrier;
//   warp 1, one lane : issues tcgen05.mma.cta_group::1.kind::tf32 (M=128, N=256, K=8; 12 per stage)
//                      from shared-memory matrix descriptors, tcgen05.commit -> the stage's "empty"
//                      barrier, and after the last k-block -> the "accumulator full" barrier;
//                      warp 1 also owns the 256-column TMEM allocation;
//   warps 2-5        : epilogue (epilogue_rows) - tcgen05.ld 32 lanes x 16 / 32 columns at a time, alpha / beta /
//                      gamma D and the compound output strides of the descriptor; thread = output row out of TMEM,
//                      and either stored that way (C contiguous along m) or through a 32 x 33 shared-memory tile
//                      per warp with lane = column (C contiguous along n); every block of C / D loads is issued
//                      before its first use.
// Edges: TMA zero-fills rows >= M / N and k >= K; the epilogue masks rows and columns.
//
// Round 2: contract_tf32x3_tc05_2sm_kernel - CTA pairs (2-CTA clusters), tcgen05.mma.cta_group::2 with M = 256 over the
// two SMs, each CTA staging its 128 rows of A and half of the B tile (64 KB stages, three of them) - runs the even part
// of the m-tiles; the kernel described above keeps an odd last tile column (and PCCB_TC05_2SM=0).
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace {

#ifndef PCCB_TC05_BKF
#define PCCB_TC05_BKF 32
#endif
constexpr int BM = 128, BN = 256, BKF = PCCB_TC05_BKF;   // 32 floats = one 128-byte swizzle row (16: 64-byte swizzle, twice the stages)
constexpr int A_BYTES = BM * BKF * 4, B_BYTES = BN * BKF * 4;
constexpr int NTHREADS = 192;
constexpr int TMEM_COLS = 256;
constexpr int MAXSEG = 2;

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap* map, int c0, int c1, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
// shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp layout): K-major tile, 128-byte
// swizzle, 8-row groups 1024 bytes apart, descriptor version 1
__device__ __forceinline__ uint64_t smem_desc(unsigned addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8 * BKF * 4) >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(BKF == 32 ? 2 : 4) << 61;      // SWIZZLE_128B / SWIZZLE_64B
    return d;
}
__device__ __forceinline__ void umma_tf32(unsigned tmem_d, uint64_t adesc, uint64_t bdesc, unsigned idesc, unsigned accumulate) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ unsigned cluster_ctarank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}

// ---- cta_group::2 forms (one MMA over the two SMs of a CTA pair; cute/arch/{copy_sm100_tma,mma_sm100_umma}.hpp) ----
// executed by both CTAs of the pair: data into the issuing CTA's shared memory, bytes counted on the LEADER's barrier
// (the shared::cluster address of the same barrier with the CTA-rank bit cleared)
__device__ __forceinline__ void tma_load_2d_2sm(unsigned dst, const CUtensorMap* map, int c0, int c1, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar & 0xFEFFFFFFu) : "memory");
}
__device__ __forceinline__ void umma_tf32_2sm(unsigned tmem_d, uint64_t adesc, uint64_t bdesc, unsigned idesc, unsigned accumulate) {
    const unsigned z = 0;
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
                 "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n}\n"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(z) : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(unsigned bar, unsigned short mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
                 ::"r"(bar), "h"(mask) : "memory");
}

struct alignas(64) TcMaps {
    CUtensorMap a[MAXSEG], alo[MAXSEG], b[MAXSEG], blo[MAXSEG];
    CUtensorMap bh[MAXSEG], blh[MAXSEG];      // B with a 128-row box: the half of the tile each CTA of a pair stages
};

struct TcParams {
    unsigned M, N;
    unsigned m_tile0;         // first m-tile of this launch (the odd last tile column runs without a cluster)
    int nseg;
    int nkb[MAXSEG];          // k-blocks (of 32) per segment
    float* C;
    DIdx cm, cn;
    const float* D;
    DIdx dm, dn;
    float alpha, beta, gamma;
    int c_tr;                 // epilogue leaves through shared memory, lane = column (set by the host: n contiguous in C)
};

// ---- epilogue (warps 2-5 of either kernel): TMEM -> alpha * acc + beta * C + gamma * D -> C -------------------------
// Warp w may touch TMEM lanes 32 (w % 4) .. +31, one output ROW per thread.  Written straight from that mapping a warp
// store touches 32 rows: fine when the m index has unit stride in C (Zr[mijB], rows (m,B)), 32 separate sectors per
// instruction when it is the n index that runs contiguously (r2[ijAb], Wr1[jBme], ...).  Those outputs go through a
// 32 x 33 shared-memory tile per warp and leave with lane = COLUMN: one 128-byte line per store.  D is added in
// whichever of the two mappings reads it along its own unit stride.
constexpr int EPI_PITCH = 33;
constexpr int EPI_BYTES = 4 * 32 * EPI_PITCH * 4;

__device__ __forceinline__ void tmem_ld16(unsigned taddr, unsigned* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr));
}

__device__ __forceinline__ void epilogue_rows(const TcParams& p, unsigned tmem, unsigned m0, unsigned n0, int warp, int lane,
                                              float* stage) {
    const unsigned lane_base = (warp & 3) * 32;
    const unsigned row = m0 + lane_base + lane;
    const bool row_ok = row < p.M;
    const bool use_beta = p.beta != 0.f, use_d = p.D != nullptr;
    const long long coff_row = row_ok ? didx_off(p.cm, row) : -1;
    const float* drow = use_d ? p.D + (row_ok ? didx_off(p.dm, row) : 0) : nullptr;
    if (!p.c_tr) {
        float* crow = p.C + (row_ok ? coff_row : 0);
        const bool c_flat = p.cn.flat && p.cn.s_lo == 1, d_flat = !use_d || (p.dn.flat && p.dn.s_lo == 1);
#pragma unroll 1
        for (unsigned c = 0; c < BN; c += 16) {
            unsigned v[16];
            tmem_ld16(tmem + (lane_base << 16) + c, v);
            asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
            if (!row_ok || n0 + c >= p.N) continue;
            // raw loads first, arithmetic after (a value consumed next to its load serialises the round trips)
            float cv[16], dv[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                const unsigned col = n0 + c + q;
                cv[q] = (use_beta && col < p.N) ? crow[c_flat ? (long long)col : didx_off(p.cn, col)] : 0.f;
            }
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                const unsigned col = n0 + c + q;
                dv[q] = (use_d && col < p.N) ? drow[d_flat ? (long long)col : didx_off(p.dn, col)] : 0.f;
            }
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                const unsigned col = n0 + c + q;
                if (col < p.N)
                    crow[c_flat ? (long long)col : didx_off(p.cn, col)] =
                        p.alpha * __uint_as_float(v[q]) + (p.beta * cv[q] + p.gamma * dv[q]);
            }
        }
        return;
    }
    // lane = column on the way out
    const bool d_tr = use_d && p.dn.s_lo == 1 && p.dm.s_lo != 1;
    const long long doff_row = (use_d && row_ok) ? didx_off(p.dm, row) : -1;
    float* tile = stage + (warp & 3) * 32 * EPI_PITCH;
#pragma unroll 1
    for (unsigned c = 0; c < BN; c += 32) {
        if (n0 + c >= p.N) break;                       // uniform over the CTA
        unsigned v[32];
        tmem_ld16(tmem + (lane_base << 16) + c, v);
        tmem_ld16(tmem + (lane_base << 16) + c + 16, v + 16);
        asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
        // (every batch of global loads is issued before its first value is used: a warp issues in order, and a load
        //  consumed right away costs one DRAM round trip per element - 4 warps per SM cannot hide that)
        float dd[32];
        if (use_d && !d_tr) {
#pragma unroll
            for (int q = 0; q < 32; ++q) {
                const unsigned col = n0 + c + q;
                dd[q] = (row_ok && col < p.N) ? drow[didx_off(p.dn, col)] : 0.f;
            }
        }
#pragma unroll
        for (int q = 0; q < 32; ++q) {
            float x = p.alpha * __uint_as_float(v[q]);
            if (use_d && !d_tr) x += p.gamma * dd[q];
            tile[lane * EPI_PITCH + q] = x;
        }
        __syncwarp();
        const unsigned col = n0 + c + lane;
        const bool col_ok = col < p.N;
        const long long coff_col = col_ok ? didx_off(p.cn, col) : 0;
        const long long doff_col = (d_tr && col_ok) ? didx_off(p.dn, col) : 0;
        float cx[32];
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const long long ro = __shfl_sync(0xffffffffu, coff_row, r);
            cx[r] = (use_beta && ro >= 0 && col_ok) ? p.C[ro + coff_col] : 0.f;
        }
        if (d_tr) {
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                const long long dro = __shfl_sync(0xffffffffu, doff_row, r);
                dd[r] = (dro >= 0 && col_ok) ? p.D[dro + doff_col] : 0.f;
            }
        }
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const long long ro = __shfl_sync(0xffffffffu, coff_row, r);
            float x = tile[r * EPI_PITCH + lane] + p.beta * cx[r];
            if (d_tr) x += p.gamma * dd[r];
            if (ro >= 0 && col_ok) p.C[ro + coff_col] = x;
        }
        __syncwarp();
    }
}

// FUSED = false: three K-concatenated passes (A,B), (A_lo,B), (A,B_lo), 4 stages of 48 KB.
// FUSED = true : each stage holds A, A_lo, B, B_lo of one k-block (96 KB, 2 stages) and feeds all three
//                products, so every operand byte crosses L2 -> shared memory once instead of 1.5 times.
template <bool FUSED>
__global__ void __launch_bounds__(NTHREADS, 1)
contract_tf32x3_tc05_kernel(const __grid_constant__ TcParams p, const __grid_constant__ TcMaps maps) {
    constexpr int STAGES = FUSED ? (BKF == 32 ? 2 : 4) : 4;
    constexpr int STAGE_BYTES = FUSED ? 2 * (A_BYTES + B_BYTES) : (A_BYTES + B_BYTES);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const unsigned raw_s = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned pad = (1024u - (raw_s & 1023u)) & 1023u;
    const unsigned ring_s = raw_s + pad;
    const unsigned full_s = ring_s + STAGES * STAGE_BYTES, empty_s = full_s + 8 * STAGES, tfull_s = empty_s + 8 * STAGES,
                   slot_s = tfull_s + 8;
    volatile unsigned* slot = reinterpret_cast<volatile unsigned*>(smem_raw + pad + STAGES * STAGE_BYTES + 16 * STAGES + 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned m0 = (p.m_tile0 + blockIdx.x) * BM, n0 = blockIdx.y * BN;
    int total_kb = 0;
    for (int s = 0; s < p.nseg; ++s) total_kb += (FUSED ? 1 : 3) * p.nkb[s];

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_s + 8 * s, 1); mbar_init(empty_s + 8 * s, 1); }
        mbar_init(tfull_s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(slot_s), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    const unsigned tmem = *slot;

    if (warp == 0) {
        if (lane == 0) {
            int g = 0;
            for (int s = 0; s < p.nseg; ++s)
                for (int pass = 0; pass < (FUSED ? 1 : 3); ++pass) {          // (A,B), (A_lo,B), (A,B_lo)
                    const CUtensorMap* ma = pass == 1 ? &maps.alo[s] : &maps.a[s];
                    const CUtensorMap* mb = pass == 2 ? &maps.blo[s] : &maps.b[s];
                    for (int kb = 0; kb < p.nkb[s]; ++kb, ++g) {
                        const int st = g % STAGES;
                        if (g >= STAGES) mbar_wait(empty_s + 8 * st, ((g / STAGES) - 1) & 1);
                        const unsigned bar = full_s + 8 * st, base = ring_s + st * STAGE_BYTES;
                        mbar_expect_tx(bar, STAGE_BYTES);
                        tma_load_2d(base, ma, kb * BKF, (int)m0, bar);
                        tma_load_2d(base + A_BYTES, mb, kb * BKF, (int)n0, bar);
                        if (FUSED) {
                            tma_load_2d(base + A_BYTES + B_BYTES, &maps.alo[s], kb * BKF, (int)m0, bar);
                            tma_load_2d(base + 2 * A_BYTES + B_BYTES, &maps.blo[s], kb * BKF, (int)n0, bar);
                        }
                    }
                }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D = F32, A = B = TF32, both K-major, N = 256, M = 128
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(BN >> 3) << 17) | ((unsigned)(BM >> 4) << 24);
            for (int g = 0; g < total_kb; ++g) {
                const int st = g % STAGES;
                mbar_wait(full_s + 8 * st, (g / STAGES) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
                const unsigned base = ring_s + st * STAGE_BYTES;
                const uint64_t ad = smem_desc(base), bd = smem_desc(base + A_BYTES);
#pragma unroll
                for (int k = 0; k < BKF / 8; ++k) {        // UMMA_K = 8 tf32 = 32 bytes: start address + 2 (x16 bytes)
                    umma_tf32(tmem, ad + 2 * k, bd + 2 * k, idesc, (g > 0 || k > 0) ? 1u : 0u);
                    if (FUSED) {
                        const uint64_t al = smem_desc(base + A_BYTES + B_BYTES), bl = smem_desc(base + 2 * A_BYTES + B_BYTES);
                        umma_tf32(tmem, al + 2 * k, bd + 2 * k, idesc, 1u);
                        umma_tf32(tmem, ad + 2 * k, bl + 2 * k, idesc, 1u);
                    }
                }
                umma_commit(empty_s + 8 * st);
            }
            umma_commit(tfull_s);
        }
    } else {
        mbar_wait(tfull_s, 0);
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
        epilogue_rows(p, tmem, m0, n0, warp, lane, reinterpret_cast<float*>(smem_raw + pad + STAGES * STAGE_BYTES + 128));
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
}

// ---- cta_group::2 kernel: one 256 x 256 output tile per CTA PAIR ------------------------------------------------
// Each CTA stages its own 128 rows of A and HALF of the B tile (128 of the 256 n-rows), both planes: 64 KB per stage,
// three stages.  The leader CTA (rank 0) issues tcgen05.mma.cta_group::2 (M = 256, N = 256, K = 8): the tensor cores of
// both SMs read A from their own shared memory and each half of B once - half the B reads per SM of the single-CTA
// kernel, whose limiter is shared-memory bandwidth.  Both producers count their bytes on the
// leader's "full" barrier; the leader's commits arrive on the "empty" / "accumulator full" barriers of both CTAs;
// each CTA's epilogue warps drain its own 128 TMEM lanes.
constexpr int STAGES_2SM = 3;
constexpr int STAGE_2SM_BYTES = 2 * A_BYTES + B_BYTES;        // A, A_lo, B half, B_lo half

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
contract_tf32x3_tc05_2sm_kernel(const __grid_constant__ TcParams p, const __grid_constant__ TcMaps maps) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const unsigned raw_s = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned pad = (1024u - (raw_s & 1023u)) & 1023u;
    const unsigned ring_s = raw_s + pad;
    const unsigned full_s = ring_s + STAGES_2SM * STAGE_2SM_BYTES, empty_s = full_s + 8 * STAGES_2SM,
                   tfull_s = empty_s + 8 * STAGES_2SM, slot_s = tfull_s + 8;
    volatile unsigned* slot = reinterpret_cast<volatile unsigned*>(smem_raw + pad + STAGES_2SM * STAGE_2SM_BYTES +
                                                                   16 * STAGES_2SM + 8);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned rank = cluster_ctarank();
    const bool leader = rank == 0;
    const unsigned m0 = (p.m_tile0 + blockIdx.x) * BM, n0 = blockIdx.y * BN;
    int total_kb = 0;
    for (int s = 0; s < p.nseg; ++s) total_kb += p.nkb[s];

    if (tid == 0) {
        for (int s = 0; s < STAGES_2SM; ++s) { mbar_init(full_s + 8 * s, 1); mbar_init(empty_s + 8 * s, 1); }
        mbar_init(tfull_s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(slot_s), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    const unsigned tmem = *slot;
    cluster_sync_all();

    if (warp == 0) {
        if (lane == 0) {
            int g = 0;
            const int nrow = (int)(n0 + rank * (BN / 2));
            for (int s = 0; s < p.nseg; ++s)
                for (int kb = 0; kb < p.nkb[s]; ++kb, ++g) {
                    const int st = g % STAGES_2SM;
                    if (g >= STAGES_2SM) mbar_wait(empty_s + 8 * st, ((g / STAGES_2SM) - 1) & 1);
                    const unsigned bar = full_s + 8 * st, base = ring_s + st * STAGE_2SM_BYTES;
                    if (leader) mbar_expect_tx(bar, 2 * STAGE_2SM_BYTES);          // both CTAs' bytes
                    tma_load_2d_2sm(base, &maps.a[s], kb * BKF, (int)m0, bar);
                    tma_load_2d_2sm(base + A_BYTES, &maps.alo[s], kb * BKF, (int)m0, bar);
                    tma_load_2d_2sm(base + 2 * A_BYTES, &maps.bh[s], kb * BKF, nrow, bar);
                    tma_load_2d_2sm(base + 2 * A_BYTES + B_BYTES / 2, &maps.blh[s], kb * BKF, nrow, bar);
                }
        }
    } else if (warp == 1) {
        if (lane == 0 && leader) {
            // instruction descriptor: D = F32, A = B = TF32, both K-major, N = 256, M = 256 (over the pair)
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(BN >> 3) << 17) | ((unsigned)((2 * BM) >> 4) << 24);
            for (int g = 0; g < total_kb; ++g) {
                const int st = g % STAGES_2SM;
                mbar_wait(full_s + 8 * st, (g / STAGES_2SM) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
                const unsigned base = ring_s + st * STAGE_2SM_BYTES;
                const uint64_t ad = smem_desc(base), al = smem_desc(base + A_BYTES), bd = smem_desc(base + 2 * A_BYTES),
                               bl = smem_desc(base + 2 * A_BYTES + B_BYTES / 2);
#pragma unroll
                for (int k = 0; k < BKF / 8; ++k) {
                    umma_tf32_2sm(tmem, ad + 2 * k, bd + 2 * k, idesc, (g > 0 || k > 0) ? 1u : 0u);
                    umma_tf32_2sm(tmem, al + 2 * k, bd + 2 * k, idesc, 1u);
                    umma_tf32_2sm(tmem, ad + 2 * k, bl + 2 * k, idesc, 1u);
                }
                umma_commit_2sm(empty_s + 8 * st, (unsigned short)3);
            }
            umma_commit_2sm(tfull_s, (unsigned short)3);
        }
    } else {
        mbar_wait(tfull_s, 0);
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
        epilogue_rows(p, tmem, m0, n0, warp, lane, reinterpret_cast<float*>(smem_raw + pad + STAGES_2SM * STAGE_2SM_BYTES + 128));
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    cluster_sync_all();                      // both epilogues are done with the pair's tensor memory
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
}


// lo[i] = x[i] - trunc_tf32(x[i])   (exact in FP32; what the tensor core drops from an FP32 word)
__global__ void split_lo_kernel(const float* __restrict__ x, float* __restrict__ lo, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = x[i];
        lo[i] = v - __uint_as_float(__float_as_uint(v) & 0xffffe000u);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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

bool make_map(CUtensorMap* m, const float* base, long long rows, long long K, long long row_stride, int box_rows) {
    EncodeTiledFn enc = encode_fn();
    if (!enc || row_stride < K) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_stride * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)BKF, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, BKF == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// operand = K-contiguous matrix with one uniform row stride, 16-byte aligned rows
// A compound index is a plain matrix dimension when it is flat (same convention as to_didx in
// common.cuh) or when its two levels happen to be contiguous (s_hi == div * s_lo: the planner emits
// one div per index group, e.g. (j,B) of Wr1[jBme] next to the strided (j,B) of <mj|eB>).
bool is_flat(const pccb_cidx& c, long long extent) { return c.div <= 1 || c.div >= extent || c.s_hi == c.div * c.s_lo; }
long long row_stride_of(const pccb_cidx& c, long long extent) { return (c.div == 1 && extent > 1) ? c.s_hi : c.s_lo; }

bool operand_ok(const void* base, const pccb_cidx& mn, const pccb_cidx& k, long long rows, long long K) {
    if (!is_flat(k, K) || !is_flat(mn, rows) || row_stride_of(k, K) != 1) return false;
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return false;
    const long long rs = row_stride_of(mn, rows);
    return rs > 0 && (rs & 3) == 0;
}

const float* lo_of(pccb_ctx* ctx, const void* ptr) {
    const char* q = static_cast<const char*>(ptr);
    for (const auto& e : ctx->lo_planes)
        if (q >= e.base && q < e.base + e.bytes) return reinterpret_cast<const float*>(e.lo + (q - e.base));
    return nullptr;
}

}  // namespace

// returns 1 when the tcgen05 kernel was launched, 0 when the contraction is not eligible, < 0 on error
int pccb_contract_f32_tc05(pccb_ctx* ctx, const pccb_contraction* d) {
    static const bool enabled = [] { const char* e = getenv("PCCB_TC05"); return e ? atoi(e) != 0 : true; }();
    // (a short M - t1[ie] <ja|be>, M = o - still pays: a quarter-filled 128-row tile at the tcgen05 rate beats the
    //  mma.sync kernel's short-K pipeline, 4.5 -> 1.x ms at (40,300))
    if (!enabled || ctx->lo_planes.empty() || d->batch != 1 || d->M < BM / 4 || d->N < BN) return 0;
    int ns = 0;
    const pccb_segment* segs[MAXSEG];
    for (int s = 0; s < d->nseg; ++s) {
        const pccb_segment& g = d->seg[s];
        if (g.K == 0 || g.scale == 0.0) continue;
        if (ns == MAXSEG) return 0;
        segs[ns++] = &g;
    }
    if (ns == 0) return 0;
    // K panels: with one segment and an A operand (both planes) larger than a slice of L2, the contraction
    // is run panel by panel over K, accumulating into C (beta = 1 after the first panel).  The A panel then
    // stays L2-resident while the big B operand streams once per panel; without this the CTAs that share
    // operand tiles drift apart and the ladder at (40,300) read 511 GB from DRAM for 65 GB of operands.
    // (the two knobs are read on every call - a getenv is nothing next to a launch - so that a test can lower them)
    const long long panel_bytes = [] { const char* e = getenv("PCCB_TC05_PANEL_MB"); return (long long)(e ? atoi(e) : 128) << 20; }();
    long long kpanel = 0;
    // (only when B is the big streaming operand, N >> M: the ladder; for square-ish products such as the ring
    //  terms the extra read-modify-write passes over C cost more than the panels save)
    const long long panel_ratio = [] { const char* e = getenv("PCCB_TC05_PANEL_RATIO"); return (long long)(e ? atoi(e) : 8); }();
    if (ns == 1 && panel_bytes > 0 && d->N >= panel_ratio * d->M) {
        const long long a_bytes = 8ll * d->M * segs[0]->K;          // two FP32 planes
        if (a_bytes > panel_bytes) {
            kpanel = panel_bytes / (8ll * d->M) / BKF * BKF;
            if (kpanel < 512) kpanel = 512;
        }
    }
    const long long min_k = [] { const char* e = getenv("PCCB_TC05_MIN_K"); return e ? atoll(e) : 256ll; }();
    for (int s = 0; s < ns; ++s) {
        const pccb_segment& g = *segs[s];
        if (g.scale != segs[0]->scale || g.K < min_k) return 0;
        if (!operand_ok(g.A_dev, g.am, g.ak, d->M, g.K) || !operand_ok(g.B_dev, g.bn, g.bk, d->N, g.K)) return 0;
        if (!lo_of(ctx, g.A_dev) || !lo_of(ctx, g.B_dev)) return 0;
    }
    const long long gm = (d->M + BM - 1) / BM, gn = (d->N + BN - 1) / BN;
    if (gn > 65535) return 0;
    constexpr size_t SMEM = (BKF == 32 ? 4 : 8) * (A_BYTES + B_BYTES) + 128 + EPI_BYTES + 1024;       // both variants: 192 KB of ring
    static const bool fused = [] { const char* e = getenv("PCCB_TC05_FUSED"); return e ? atoi(e) != 0 : true; }();
    static bool attr_set = false;
    if (!attr_set) {
        PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_tf32x3_tc05_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
        PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_tf32x3_tc05_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
        attr_set = true;
    }
    const long long Ktot = segs[0]->K;
    const long long step = kpanel > 0 ? kpanel : Ktot;
    for (long long k0 = 0; k0 < (kpanel > 0 ? Ktot : 1); k0 += step) {
        TcParams p;
        memset(&p, 0, sizeof(p));
        TcMaps maps;
        memset(&maps, 0, sizeof(maps));
        for (int s = 0; s < ns; ++s) {
            const pccb_segment& g = *segs[s];
            const long long koff = kpanel > 0 ? k0 : 0;
            const long long Kp = kpanel > 0 ? (Ktot - k0 < step ? Ktot - k0 : step) : g.K;
            const float* a = static_cast<const float*>(g.A_dev) + koff;
            const float* b = static_cast<const float*>(g.B_dev) + koff;
            const float* alo = lo_of(ctx, g.A_dev) + koff;
            const float* blo = lo_of(ctx, g.B_dev) + koff;
            const long long ars = row_stride_of(g.am, d->M), brs = row_stride_of(g.bn, d->N);
            if (!make_map(&maps.a[s], a, d->M, Kp, ars, BM) || !make_map(&maps.alo[s], alo, d->M, Kp, ars, BM) ||
                !make_map(&maps.b[s], b, d->N, Kp, brs, BN) || !make_map(&maps.blo[s], blo, d->N, Kp, brs, BN) ||
                !make_map(&maps.bh[s], b, d->N, Kp, brs, BN / 2) || !make_map(&maps.blh[s], blo, d->N, Kp, brs, BN / 2)) {
                if (k0 == 0) return 0;
                PCCB_FAIL(ctx, PCCB_ERR_ARG, "pccb_contract (tcgen05): tensor map of K panel %lld failed", (long long)k0);
            }
            p.nkb[s] = (int)((Kp + BKF - 1) / BKF);
        }
        p.M = (unsigned)d->M;
        p.N = (unsigned)d->N;
        p.nseg = ns;
        p.C = static_cast<float*>(d->C_dev);
        p.cm = to_didx(d->cm, d->M);
        p.cn = to_didx(d->cn, d->N);
        p.alpha = (float)(d->alpha * segs[0]->scale);
        {
            const char* e = getenv("PCCB_TC05_EPI");        // 0: thread = row stores for every output
            p.c_tr = (p.cn.s_lo == 1 && p.cm.s_lo != 1 && !(e && atoi(e) == 0)) ? 1 : 0;
        }
        if (k0 == 0) {
            p.D = static_cast<const float*>(d->D_dev);
            p.dm = to_didx(d->dm, d->M);
            p.dn = to_didx(d->dn, d->N);
            p.beta = (float)d->beta;
            p.gamma = d->D_dev ? (float)d->gamma : 0.f;
        } else {
            p.beta = 1.f;
        }
        // CTA pairs (cta_group::2, 256 x 256 tiles) for the even part of the m-tiles, the single-CTA kernel for an odd last
        // tile column.  Measured at (40,300) (tools/time_sp_ladder.py): ring product M = N = K = 12000 15.64 -> 14.48 ms
        // (221 -> 239 TFLOP/s), ladder (13 m-tiles = 6 pairs + 1) 137.5 -> 128.5 ms (188 -> 202).  PCCB_TC05_2SM=0 disables.
        static const int two_sm = [] { const char* e = getenv("PCCB_TC05_2SM"); return e ? atoi(e) : 1; }();
        // (an odd tile count costs a second launch: worth it only when the pair part runs many waves - one rank's ladder
        //  out of eight, 12 x 45 + 45 tiles, went 14.4 -> 15.4 ms with the split)
        const bool many_waves = (gm & ~1ll) * gn >= 8ll * ctx->sm_count;
        const long long gm2 = (fused && two_sm > 0 && gm >= 2 && ((gm & 1) == 0 || many_waves || two_sm >= 2)) ? (gm & ~1ll) : 0;
        if (gm2 > 0) {
            p.m_tile0 = 0;
            constexpr size_t SMEM2 = (size_t)STAGES_2SM * STAGE_2SM_BYTES + 128 + EPI_BYTES + 1024;
            static bool attr2 = false;
            if (!attr2) {
                PCCB_CUDA(ctx, cudaFuncSetAttribute(contract_tf32x3_tc05_2sm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM2));
                attr2 = true;
            }
            contract_tf32x3_tc05_2sm_kernel<<<dim3((unsigned)gm2, (unsigned)gn), NTHREADS, SMEM2, ctx->stream>>>(p, maps);
            PCCB_CHECK_LAUNCH(ctx, "contract_tf32x3_tc05_2sm_kernel");
        }
        if (gm2 < gm) {
            p.m_tile0 = (unsigned)gm2;
            if (fused)
                contract_tf32x3_tc05_kernel<true><<<dim3((unsigned)(gm - gm2), (unsigned)gn), NTHREADS, SMEM, ctx->stream>>>(p, maps);
            else
                contract_tf32x3_tc05_kernel<false><<<dim3((unsigned)(gm - gm2), (unsigned)gn), NTHREADS, SMEM, ctx->stream>>>(p, maps);
            PCCB_CHECK_LAUNCH(ctx, "contract_tf32x3_tc05_kernel");
        }
    }
    return 1;
}

extern "C" int pccb_split_lo(pccb_ctx* ctx, const void* x_dev, void* lo_dev, int64_t n) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, x_dev && lo_dev && n > 0, "pccb_split_lo: bad arguments");
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    split_lo_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(static_cast<const float*>(x_dev), static_cast<float*>(lo_dev), n);
    PCCB_CHECK_LAUNCH(ctx, "split_lo_kernel");
    return PCCB_OK;
}

extern "C" int pccb_register_lo_plane(pccb_ctx* ctx, const void* x_dev, const void* lo_dev, int64_t n) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, x_dev && lo_dev && n > 0, "pccb_register_lo_plane: bad arguments");
    // a new registration replaces every entry whose address range overlaps it (same base re-registered, or
    // memory recycled by the caller's allocator after the tensor that owned the old entry was freed)
    const char* b = static_cast<const char*>(x_dev);
    const size_t bytes = (size_t)n * sizeof(float);
    auto& v = ctx->lo_planes;
    for (size_t q = 0; q < v.size();) {
        if (b < v[q].base + v[q].bytes && v[q].base < b + bytes) v.erase(v.begin() + q);
        else ++q;
    }
    v.push_back({b, static_cast<const char*>(lo_dev), bytes});
    return PCCB_OK;
}

extern "C" int pccb_unregister_lo_plane(pccb_ctx* ctx, const void* x_dev) {
    if (!ctx) return PCCB_ERR_ARG;
    auto& v = ctx->lo_planes;
    for (size_t q = 0; q < v.size();) {
        if (v[q].base == static_cast<const char*>(x_dev)) v.erase(v.begin() + q);
        else ++q;
    }
    return PCCB_OK;
}

extern "C" int pccb_clear_lo_planes(pccb_ctx* ctx) {
    if (!ctx) return PCCB_ERR_ARG;
    ctx->lo_planes.clear();
    return PCCB_OK;
}
