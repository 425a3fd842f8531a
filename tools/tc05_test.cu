// Stand-alone bring-up test of a tcgen05 (5th-gen tensor core) TF32 GEMM for sm_100a:
//   C[M,N] (+)= A[M,K] * B[N,K]^T, FP32 in memory, kind::tf32, accumulators in TMEM.
// TMA (128-byte swizzle) -> shared ring -> tcgen05.mma issued by one thread -> tcgen05.ld epilogue.
// Checks (i) exact results on TF32-representable inputs against a host loop, (ii) what the tensor
// core does with the 13 low mantissa bits of an FP32 operand (truncate or round), (iii) throughput.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tc05_test tools/tc05_test.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int BM = 128, BN = 256, BKF = 32;            // tile; 32 floats = one 128-byte swizzle row
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BKF * 4, B_BYTES = BN * BKF * 4, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int NTHREADS = 192;                           // warp 0 TMA, warp 1 MMA (+TMEM alloc), warps 2-5 epilogue
constexpr int TMEM_COLS = 256;

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
// shared-memory matrix descriptor: K-major tile, 128-byte swizzle, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t smem_desc(unsigned addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFF) >> 4);            // start address
    d |= (uint64_t)1 << 16;                            // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                  // stride byte offset
    d |= (uint64_t)1 << 46;                            // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                            // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ void umma_tf32(unsigned tmem_d, uint64_t adesc, uint64_t bdesc, unsigned idesc, unsigned accumulate) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}

struct Maps { CUtensorMap a, b; };

__global__ void __launch_bounds__(NTHREADS, 1)
tf32_gemm_kernel(const __grid_constant__ Maps maps, float* __restrict__ C, int M, int N, int K, int accumulate_c) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const unsigned raw_s = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned pad = (1024u - (raw_s & 1023u)) & 1023u;
    const unsigned ring_s = raw_s + pad;
    const unsigned bars_s = ring_s + STAGES * STAGE_BYTES;          // full[S], empty[S], tmem_full, tmem slot
    const unsigned full_s = bars_s, empty_s = bars_s + 8 * STAGES, tfull_s = empty_s + 8 * STAGES, slot_s = tfull_s + 8;
    volatile unsigned* slot = reinterpret_cast<volatile unsigned*>(smem_raw + pad + STAGES * STAGE_BYTES + 16 * STAGES + 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int nkb = K / BKF;

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
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % STAGES;
                if (kb >= STAGES) mbar_wait(empty_s + 8 * s, ((kb / STAGES) - 1) & 1);
                const unsigned bar = full_s + 8 * s;
                mbar_expect_tx(bar, STAGE_BYTES);
                tma_load_2d(ring_s + s * STAGE_BYTES, &maps.a, kb * BKF, m0, bar);
                tma_load_2d(ring_s + s * STAGE_BYTES + A_BYTES, &maps.b, kb * BKF, n0, bar);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D = F32, A = B = TF32, both K-major, N = 256, M = 128
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(BN >> 3) << 17) | ((unsigned)(BM >> 4) << 24);
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % STAGES;
                mbar_wait(full_s + 8 * s, (kb / STAGES) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
                const uint64_t ad = smem_desc(ring_s + s * STAGE_BYTES), bd = smem_desc(ring_s + s * STAGE_BYTES + A_BYTES);
#pragma unroll
                for (int k = 0; k < BKF / 8; ++k)          // UMMA_K = 8 tf32 = 32 bytes: advance the start address by 2 (x16 bytes)
                    umma_tf32(tmem, ad + 2 * k, bd + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                umma_commit(empty_s + 8 * s);               // frees the stage when these MMAs have read it
            }
            umma_commit(tfull_s);                           // accumulator complete
        }
    } else {
        // epilogue: warp w may touch TMEM lanes 32 (w % 4) .. +31; thread = one output row
        mbar_wait(tfull_s, 0);
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
        const int lane_base = (warp & 3) * 32;
        const int row = m0 + lane_base + lane;
        float* crow = C + (size_t)row * N + n0;
#pragma unroll 1
        for (int c = 0; c < BN; c += 16) {
            unsigned v[16];
            const unsigned taddr = tmem + ((unsigned)lane_base << 16) + (unsigned)c;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                           "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                         : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
            if (row < M) {
#pragma unroll
                for (int q = 0; q < 16; q += 4) {
                    float4 o = make_float4(__uint_as_float(v[q]), __uint_as_float(v[q + 1]), __uint_as_float(v[q + 2]), __uint_as_float(v[q + 3]));
                    float4* dst = reinterpret_cast<float4*>(crow + c + q);
                    if (accumulate_c) { const float4 old = *dst; o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w; }
                    *dst = o;
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static bool make_map(CUtensorMap* m, const float* base, long long rows, long long K, int box_rows) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 4};
    const cuuint32_t box[2] = {(cuuint32_t)BKF, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return ((EncodeTiledFn)p)(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static int run(int M, int N, int K, int mode) {
    std::vector<float> A((size_t)M * K), B((size_t)N * K), Cref((size_t)M * N), Cg((size_t)M * N);
    srand(1);
    for (auto& x : A) x = (float)((rand() % 17) - 8) * 0.125f;      // exactly representable in TF32
    for (auto& x : B) x = (float)((rand() % 13) - 6) * 0.25f;
    if (mode == 1) {       // probe the handling of the low 13 mantissa bits
        for (auto& x : A) x = 0.f;
        for (auto& x : B) x = 0.f;
        for (int m = 0; m < M; ++m) A[(size_t)m * K] = 1.0f + 1.0f / 2048 + 1.0f / 4096;   // 1 + 2^-11 + 2^-12
        for (int n = 0; n < N; ++n) B[(size_t)n * K] = 1.0f;
    }
    float *dA, *dB, *dC;
    CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dC, Cg.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(dC, 0, Cg.size() * 4));
    Maps maps;
    if (!make_map(&maps.a, dA, M, K, BM) || !make_map(&maps.b, dB, N, K, BN)) { printf("tensor map failed\n"); return 1; }
    const size_t smem = STAGES * STAGE_BYTES + 16 * STAGES + 16 + 1024;
    CK(cudaFuncSetAttribute(tf32_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(M / BM, N / BN);
    tf32_gemm_kernel<<<grid, NTHREADS, smem>>>(maps, dC, M, N, K, 0);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(Cg.data(), dC, Cg.size() * 4, cudaMemcpyDeviceToHost));
    if (mode == 1) {
        printf("low-bit probe: (1 + 2^-11 + 2^-12) * 1 -> %.10f  (1.0 = truncation, %.10f = round to nearest)\n", Cg[0], 1.0 + 1.0 / 1024);
    } else if (mode == 0) {
        double maxd = 0;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) {
                double s = 0;
                for (int k = 0; k < K; ++k) s += (double)A[(size_t)m * K + k] * B[(size_t)n * K + k];
                const double d = fabs(s - Cg[(size_t)m * N + n]);
                if (d > maxd) maxd = d;
            }
        printf("M=%d N=%d K=%d: max |C - ref| = %g %s\n", M, N, K, maxd, maxd == 0 ? "(exact)" : "");
        if (maxd != 0) return 1;
    } else {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int i = 0; i < 2; ++i) tf32_gemm_kernel<<<grid, NTHREADS, smem>>>(maps, dC, M, N, K, 0);
        cudaEventRecord(e0);
        const int reps = 5;
        for (int i = 0; i < reps; ++i) tf32_gemm_kernel<<<grid, NTHREADS, smem>>>(maps, dC, M, N, K, 0);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= reps;
        printf("M=%d N=%d K=%d: %.3f ms, %.1f TFLOP/s (single-pass TF32)\n", M, N, K, ms, 2.0 * M * N * K / ms * 1e-9);
    }
    cudaFree(dA); cudaFree(dB); cudaFree(dC);
    return 0;
}

int main() {
    if (run(128, 256, 32, 0)) return 1;
    if (run(128, 256, 256, 0)) return 1;
    if (run(256, 512, 512, 0)) return 1;
    run(128, 256, 32, 1);
    run(1536, 22528, 22528, 2);        // ladder-like: o^2 = 1600 -> 1536, v^2 = 22500 -> 22528
    return 0;
}
