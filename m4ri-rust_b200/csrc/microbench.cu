// microbench.cu — measures the roofline denominators MEASURED_PEAKS.json does not
// hold: 32-bit LOP3 issue rate, shared-memory load bandwidth (the
// real ceiling of the Four-Russians gather), legacy IMMA int8 rate, HBM copy.
// Standalone binary; prints one JSON object. Not part of libm4ri_b200.so.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e = (x);                                                               \
    if (e != cudaSuccess) {                                                            \
      fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                         \
    }                                                                                  \
  } while (0)

// ---- LOP3: 16 independent chains per thread, 3-input XOR ----------------------
__global__ void __launch_bounds__(1024, 1) lop3_kernel(uint32_t *out, int iters) {
  uint32_t r[16];
#pragma unroll
  for (int i = 0; i < 16; i++) r[i] = threadIdx.x * 2654435761u + i;
  uint32_t a = blockIdx.x | 1, b = threadIdx.x ^ 0x5555;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++)
      asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(a), "r"(b));
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s ^= r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- shared memory loads: warp-uniform row index, lane = bank (the M4RM gather) --
template <int VEC>
__global__ void __launch_bounds__(1024, 1) lds_kernel(uint32_t *out, int iters) {
  extern __shared__ __align__(16) uint32_t sm[];
  const int words = 32768;  // 128 KiB
  for (int i = threadIdx.x; i < words; i += blockDim.x) sm[i] = i * 2654435761u;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  uint32_t acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
  uint32_t idx = threadIdx.x >> 5;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      idx = (idx * 5 + 1) & 255;  // warp-uniform pseudo-random row
      if (VEC == 1) {
        acc0 ^= sm[idx * 32 + lane];
      } else if (VEC == 2) {
        uint2 v = *(const uint2 *)&sm[(idx * 32 + lane) * 2 & (words - 1)];
        acc0 ^= v.x; acc1 ^= v.y;
      } else {
        uint4 v = *(const uint4 *)&sm[(idx * 32 + lane) * 4 & (words - 1)];
        acc0 ^= v.x; acc1 ^= v.y; acc2 ^= v.z; acc3 ^= v.w;
      }
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 ^ acc1 ^ acc2 ^ acc3;
}

// ---- legacy IMMA: mma.sync m16n8k32 s8 ------------------------------------------
__global__ void __launch_bounds__(256, 4) imma_kernel(int *out, int iters) {
  int c[8][4];
#pragma unroll
  for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
  uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      asm volatile(
          "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
          "{%0,%1,%2,%3};"
          : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
          : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void copy_kernel(uint4 *dst, const uint4 *src, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}

template <typename F>
static float time_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; i++) launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best;
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  uint32_t *out;
  CK(cudaMalloc(&out, (size_t)sms * 8 * 1024 * 4));
  printf("{\"device\": \"%s\", \"sms\": %d", p.name, sms);

  {  // LOP3
    int iters = 4096;
    float ms = time_ms([&] { lop3_kernel<<<sms, 1024>>>(out, iters); }, 10);
    double ops = (double)sms * 1024 * iters * 16;
    printf(", \"lop3_per_s\": %.4e, \"lop3_ms\": %.3f", ops / (ms * 1e-3), ms);
  }
  {  // LDS
    int iters = 4096;
    CK(cudaFuncSetAttribute(lds_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
    CK(cudaFuncSetAttribute(lds_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
    CK(cudaFuncSetAttribute(lds_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
    float ms1 = time_ms([&] { lds_kernel<1><<<sms, 1024, 131072>>>(out, iters); }, 10);
    float ms2 = time_ms([&] { lds_kernel<2><<<sms, 1024, 131072>>>(out, iters); }, 10);
    float ms4 = time_ms([&] { lds_kernel<4><<<sms, 1024, 131072>>>(out, iters); }, 10);
    double loads = (double)sms * 1024 * iters * 8;
    printf(", \"lds32_bytes_per_s\": %.4e, \"lds64_bytes_per_s\": %.4e, \"lds128_bytes_per_s\": %.4e",
           loads * 4 / (ms1 * 1e-3), loads * 8 / (ms2 * 1e-3), loads * 16 / (ms4 * 1e-3));
  }
  {  // IMMA
    int iters = 2048;
    float ms = time_ms([&] { imma_kernel<<<sms * 4, 256>>>((int *)out, iters); }, 10);
    double ops = (double)sms * 4 * 8 * iters * 8 * (2.0 * 16 * 8 * 32);
    printf(", \"imma_sync_ops_per_s\": %.4e", ops / (ms * 1e-3));
  }
  {  // HBM copy, 1 GiB each way
    size_t bytes = (size_t)1 << 30;
    uint4 *a, *b;
    CK(cudaMalloc(&a, bytes));
    CK(cudaMalloc(&b, bytes));
    CK(cudaMemset(a, 1, bytes));
    float ms = time_ms([&] { copy_kernel<<<sms * 16, 512>>>(b, a, bytes / 16); }, 10);
    printf(", \"hbm_copy_bytes_per_s\": %.4e", 2.0 * bytes / (ms * 1e-3));
    CK(cudaFree(a));
    CK(cudaFree(b));
  }
  int clk = 0;
  CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
  printf(", \"clock_rate_khz\": %d}\n", clk);
  return 0;
}
