// microbench.cu -- measures the roofline denominators SURVEY.md §6/§8d asks for and MEASURED_PEAKS.json lacks:
//   L2 random-gather bandwidth (8 B loads over a 2 MB table, all SMs), shared-memory random-gather bandwidth
//   (8 B loads, conflict-free and random), FP64 add/mul/fma rate, pinned H2D / D2H bandwidth.
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
// Prints one JSON object.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s; }

// every thread issues ITER dependent-free random 8 B loads from a table that lives in L2
template <int ITER>
__global__ void l2_gather(const uint2* __restrict__ table, uint32_t mask, uint64_t* sink) {
  uint32_t s = blockIdx.x * blockDim.x + threadIdx.x + 12345u;
  uint64_t acc = 0;
#pragma unroll 8
  for (int i = 0; i < ITER; i++) {
    uint32_t idx = lcg(s) >> 8 & mask;
    uint2 v;
    asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(table + idx));  // bypass L1
    acc += v.x + v.y;
  }
  if (acc == 0x123456789ULL) *sink = acc;
}
template <int ITER>
__global__ void l1_gather(const uint2* __restrict__ table, uint32_t mask, uint64_t* sink) {
  uint32_t s = blockIdx.x * blockDim.x + threadIdx.x + 12345u;
  uint64_t acc = 0;
#pragma unroll 8
  for (int i = 0; i < ITER; i++) { uint32_t idx = lcg(s) >> 8 & mask; uint2 v = __ldg(table + idx); acc += v.x + v.y; }
  if (acc == 0x123456789ULL) *sink = acc;
}

// shared-memory gather: RANDOM = 1 random 8 B words (bank conflicts as they fall), 0 = lane-linear (conflict free)
template <int ITER, int RANDOM>
__global__ void smem_gather(uint64_t* sink) {
  extern __shared__ uint2 sm[];
  const int words = 16384;  // 128 KB
  for (int i = threadIdx.x; i < words; i += blockDim.x) sm[i] = make_uint2(i, i * 7);
  __syncthreads();
  uint32_t s = blockIdx.x * blockDim.x + threadIdx.x + 999u;
  uint64_t acc = 0; uint32_t base = threadIdx.x & 31;
#pragma unroll 8
  for (int i = 0; i < ITER; i++) {
    uint32_t r = lcg(s) >> 8;
    uint32_t idx = RANDOM ? (r & (words - 1)) : (((r & (words / 32 - 1)) * 32 + base));
    uint2 v = sm[idx]; acc += v.x + v.y;
  }
  if (acc == 0x123456789ULL) *sink = acc;
}

template <int ITER, int MODE>  // MODE 0: fma, 1: separate mul+add
__global__ void fp64_rate(double* sink, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3;
#pragma unroll 4
  for (int i = 0; i < ITER; i++) {
    if (MODE == 0) { x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b); }
    else { x0 = __dadd_rn(__dmul_rn(x0, a), b); x1 = __dadd_rn(__dmul_rn(x1, a), b); x2 = __dadd_rn(__dmul_rn(x2, a), b); x3 = __dadd_rn(__dmul_rn(x3, a), b); }
  }
  if (x0 + x1 + x2 + x3 == 1.2345) *sink = x0;
}

template <class F> static float time_ms(F&& f, int reps = 5) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) { CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms; }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int sms = p.multiProcessorCount;
  uint64_t* sink; CK(cudaMalloc(&sink, 8));
  printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
  {  // L2 gather, 2 MB table
    const uint32_t words = (2u << 20) / 8; uint2* t; CK(cudaMalloc(&t, words * 8)); CK(cudaMemset(t, 1, words * 8));
    constexpr int ITER = 4096; int blocks = sms * 8, threads = 256;
    float ms = time_ms([&] { l2_gather<ITER><<<blocks, threads>>>(t, words - 1, sink); });
    double bytes = (double)blocks * threads * ITER * 8;
    printf(", \"l2_gather_8B_2MB_GBs\": %.1f", bytes / ms / 1e6);
    ms = time_ms([&] { l1_gather<ITER><<<blocks, threads>>>(t, words - 1, sink); });
    printf(", \"l1l2_gather_8B_2MB_GBs\": %.1f", bytes / ms / 1e6);
    CK(cudaFree(t));
  }
  {  // shared memory gather
    constexpr int ITER = 8192; int blocks = sms, threads = 1024; size_t smem = 128 << 10;
    CK(cudaFuncSetAttribute(smem_gather<ITER, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(smem_gather<ITER, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    double bytes = (double)blocks * threads * ITER * 8;
    float ms = time_ms([&] { smem_gather<ITER, 0><<<blocks, threads, smem>>>(sink); });
    printf(", \"smem_gather_8B_linear_GBs\": %.1f", bytes / ms / 1e6);
    ms = time_ms([&] { smem_gather<ITER, 1><<<blocks, threads, smem>>>(sink); });
    printf(", \"smem_gather_8B_random_GBs\": %.1f", bytes / ms / 1e6);
  }
  {  // FP64
    constexpr int ITER = 8192; int blocks = sms * 4, threads = 512; double* ds; CK(cudaMalloc(&ds, 8));
    float ms = time_ms([&] { fp64_rate<ITER, 0><<<blocks, threads>>>(ds, 1.0000001, 1e-9); });
    printf(", \"fp64_fma_TFLOPs\": %.2f", (double)blocks * threads * ITER * 4 * 2 / ms / 1e9);
    ms = time_ms([&] { fp64_rate<ITER, 1><<<blocks, threads>>>(ds, 1.0000001, 1e-9); });
    printf(", \"fp64_mul_add_unfused_TFLOPs\": %.2f", (double)blocks * threads * ITER * 4 * 2 / ms / 1e9);
  }
  {  // PCIe, pinned
    size_t bytes = size_t(1) << 30; void *h, *d; CK(cudaHostAlloc(&h, bytes, cudaHostAllocDefault)); CK(cudaMalloc(&d, bytes));
    float ms = time_ms([&] { CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice)); }, 3);
    printf(", \"h2d_pinned_GBs\": %.1f", bytes / ms / 1e6);
    ms = time_ms([&] { CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost)); }, 3);
    printf(", \"d2h_pinned_GBs\": %.1f", bytes / ms / 1e6);
    // HBM copy for context (device to device)
    void* d2; CK(cudaMalloc(&d2, bytes));
    ms = time_ms([&] { CK(cudaMemcpyAsync(d2, d, bytes, cudaMemcpyDeviceToDevice)); }, 5);
    printf(", \"hbm_copy_GBs\": %.1f", 2.0 * bytes / ms / 1e6);
  }
  printf("}\n");
  return 0;
}
