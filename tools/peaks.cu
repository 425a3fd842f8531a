// Micro-benchmarks that pin the FP64/FP32 issue-rate ceilings of one B200 (sm_100a).
// Built by tools/measure_peaks.py (nvcc -> peaks binary); prints one JSON object.
// DMMA.8x8x4 (mma.sync m8n8k4 f64) vs DFMA vs FFMA, all SMs busy, register-only operands.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

template <int NACC>
__global__ void __launch_bounds__(256) dmma_loop(double* out, int iters, double seed) {
  double a = seed + threadIdx.x * 1e-9, b = seed - threadIdx.x * 1e-9;
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = 0.0; c[i][1] = 0.0; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) dfma_loop(double* out, int iters, double seed) {
  double a = seed + threadIdx.x * 1e-9, b = 1.0 - 1e-12;
  double c[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) c[i] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i] = fma(c[i], b, a);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i];
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) ffma_loop(float* out, int iters, float seed) {
  float a = seed + threadIdx.x * 1e-6f, b = 0.999999f;
  float c[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) c[i] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i] = fmaf(c[i], b, a);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i];
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// m16n8k8 tf32 (legacy mma.sync path) for the 3xTF32 decision.
template <int NACC>
__global__ void __launch_bounds__(256) tf32_loop(float* out, int iters, float seed) {
  unsigned a0 = __float_as_uint(seed), a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 + 4, b1 = a0 + 5;
  float c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3},{%4,%5,%6,%7},{%8,%9},{%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double time_ms(F launch, int reps) {
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  launch(); CK(cudaDeviceSynchronize());
  double best = 1e30;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  void* out; CK(cudaMalloc(&out, sizeof(double) * 256 * sms * 8));
  printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
  const int iters = 20000;
  for (int wpb = 4; wpb <= 8; wpb += 4) {           // warps per CTA; 1 and 2 CTAs/SM
    for (int cps = 1; cps <= 2; ++cps) {
      int grid = sms * cps, block = wpb * 32;
      double ms = time_ms([&] { dmma_loop<8><<<grid, block>>>((double*)out, iters, 1.0); }, 5);
      double fl = 512.0 * 8 * iters * (double)wpb * grid;
      printf(", \"dmma_w%d_c%d_tflops\": %.3f", wpb, cps, fl / ms * 1e-9);
    }
  }
  {
    int grid = sms * 2, block = 256;
    double ms = time_ms([&] { dfma_loop<8><<<grid, block>>>((double*)out, iters, 1.0); }, 5);
    printf(", \"dfma_tflops\": %.3f", 2.0 * 8 * iters * (double)block * grid / ms * 1e-9);
    ms = time_ms([&] { ffma_loop<8><<<grid, block>>>((float*)out, iters, 1.0f); }, 5);
    printf(", \"ffma_tflops\": %.3f", 2.0 * 8 * iters * (double)block * grid / ms * 1e-9);
    ms = time_ms([&] { tf32_loop<8><<<grid, block>>>((float*)out, iters, 1.0f); }, 5);
    printf(", \"mma_sync_tf32_tflops\": %.3f", 2.0 * 16 * 8 * 8 * 8 * iters * (double)(block / 32) * grid / ms * 1e-9);
  }
  printf("}\n");
  return 0;
}
