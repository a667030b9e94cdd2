// umma_probe.cu -- checks, against plain integer arithmetic on the host, how tcgen05.mma kind::i8 reads the K-major
// no-swizzle shared-memory layout the scoring kernels use (core matrix = 8 rows x 16 bytes, contiguous 128 B):
//     byte (row r, k) of an operand with R rows sits at  (k / 16) * (R * 16) + r * 16 + (k % 16)
// i.e. stride between 8-row groups (SBO) = 128 B, stride between the two 16-byte K halves of one MMA (LBO) = R * 16 B.
// One CTA: D[128 x N] (int32, TMEM) = A[128 x K] * B[N x K]^T over K / 32 MMAs, read back with tcgen05.ld.32x32b.
// Variants tried: (0) LBO / SBO as above, (1) the two descriptor fields swapped.  Prints which one is exact.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_probe tools/umma_probe.cu && tools/umma_probe
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;      // descriptor version for sm_100
  return d;                    // base offset 0, layout type 0 = no swizzle
}

template <int N>
__global__ void __launch_bounds__(128, 1) probe_kernel(const int8_t* __restrict__ A, const int8_t* __restrict__ B, int K, int variant, int32_t* __restrict__ D) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                       // [K/16][128][16]
  uint8_t* sB = smem + 128 * K;             // [K/16][N][16]
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 128 * K; i += 128) { int r = i / K, k = i % K; sA[(k / 16) * (128 * 16) + r * 16 + (k % 16)] = (uint8_t)A[i]; }
  for (int i = tid; i < N * K; i += 128) { int r = i / K, k = i % K; sB[(k / 16) * (N * 16) + r * 16 + (k % 16)] = (uint8_t)B[i]; }
  if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(64) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the operand bytes were written by ordinary stores
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tmem_base;
  if (tid == 0) {
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);   // S32 accum, S8 x S8, K-major both
    for (int ks = 0; ks < K / 32; ks++) {
      uint32_t a_addr = smem_u32(sA) + ks * 2 * (128 * 16), b_addr = smem_u32(sB) + ks * 2 * (N * 16);
      uint64_t da = variant == 0 ? make_desc(a_addr, 128 * 16, 128) : make_desc(a_addr, 128, 128 * 16);
      uint64_t db = variant == 0 ? make_desc(b_addr, N * 16, 128) : make_desc(b_addr, 128, N * 16);
      uint32_t acc = ks > 0 ? 1u : 0u, z = 0;
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                   "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
                   ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(z), "r"(z), "r"(z), "r"(z) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  // everybody waits for the MMAs (phase 0 of the barrier)
  asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < N; c0 += 8) {
    uint32_t v[8];
    const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16) + c0;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 8; j++) D[tid * N + c0 + j] = (int32_t)v[j];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(64) : "memory");
}

template <int N> static int run(int K) {
  std::vector<int8_t> A(128 * K), B(N * K); std::vector<int32_t> ref(128 * N), out(128 * N);
  srand(7 + N + K);
  for (auto& x : A) x = (int8_t)(rand() % 131 - 65);
  for (auto& x : B) x = (int8_t)(rand() % 131 - 65);
  for (int i = 0; i < 128; i++) for (int j = 0; j < N; j++) { int s = 0; for (int k = 0; k < K; k++) s += (int)A[i * K + k] * (int)B[j * K + k]; ref[i * N + j] = s; }
  int8_t *dA, *dB; int32_t* dD;
  cudaMalloc(&dA, A.size()); cudaMalloc(&dB, B.size()); cudaMalloc(&dD, out.size() * 4);
  cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
  int good = -1;
  for (int variant = 0; variant < 1; variant++) {   // variant 1 (descriptor fields swapped) faults with an illegal memory access on the B200: profiles/umma_probe_r02.txt
    cudaMemset(dD, 0xFF, out.size() * 4);
    size_t sm = (size_t)(128 + N) * K + 1024;
    cudaFuncSetAttribute(probe_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    probe_kernel<N><<<1, 128, sm>>>(dA, dB, K, variant, dD);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N=%d K=%d variant %d: CUDA error %s\n", N, K, variant, cudaGetErrorString(e)); return -2; }
    cudaMemcpy(out.data(), dD, out.size() * 4, cudaMemcpyDeviceToHost);
    long bad = 0; for (size_t i = 0; i < out.size(); i++) bad += out[i] != ref[i];
    printf("N=%d K=%d variant %d (%s): %ld of %zu entries differ; D[0][0]=%d ref=%d D[5][3]=%d ref=%d\n", N, K, variant,
           variant == 0 ? "LBO=K-halves stride, SBO=128" : "fields swapped", bad, out.size(), out[0], ref[0], out[5 * N + 3], ref[5 * N + 3]);
    if (bad == 0 && good < 0) good = variant;
  }
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
  return good;
}

int main() {
  int g1 = run<32>(64), g2 = run<32>(256), g3 = run<16>(64), g4 = run<64>(128);
  printf("exact variant: N=32,K=64 -> %d; N=32,K=256 -> %d; N=16,K=64 -> %d; N=64,K=128 -> %d\n", g1, g2, g3, g4);
  return (g1 == 0 && g2 == 0 && g3 == 0 && g4 == 0) ? 0 : 1;
}
