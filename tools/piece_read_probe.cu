// Dev aid: DRAM read rate of the key pre-pass's access pattern -- every CTA column (blockIdx.y) reads one piece of PIECE bytes of
// every 800-byte row of its tiles -- as a function of the piece length.  nvcc -O3 -arch=sm_100a -o piece_read_probe piece_read_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int PIECE_DOUBLES>
__global__ void __launch_bounds__(512, 2) probe(const double* __restrict__ X, int F, long n, int E, long ntiles, double* sink) {
  const int j = threadIdx.x % PIECE_DOUBLES, ev0 = threadIdx.x / PIECE_DOUBLES, evstep = 512 / PIECE_DOUBLES;
  const int f = blockIdx.y * PIECE_DOUBLES + j;
  double acc = 0;
  for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const double* xt = X + tile * (long)E * F + f;
    for (int e = ev0; e < E; e += evstep * 5) {
      double v[5];
#pragma unroll
      for (int k = 0; k < 5; k++) { int ee = e + k * evstep; v[k] = ee < E ? __ldcs(xt + (long)ee * F) : 0.0; }
#pragma unroll
      for (int k = 0; k < 5; k++) acc += v[k];
    }
  }
  if (acc == 1.2345e-300) sink[0] = acc;
}
template <int PD> static void run(const double* X, int F, long n, double* sink) {
  const int E = 640; const long ntiles = n / E; const int nby = F / PD; int gx = 2 * 148 / nby; if (gx < 1) gx = 1;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int it = 0; it < 3; it++) probe<PD><<<dim3(gx, nby), 512>>>(X, F, n, E, ntiles, sink);
  cudaEventRecord(a);
  for (int it = 0; it < 10; it++) probe<PD><<<dim3(gx, nby), 512>>>(X, F, n, E, ntiles, sink);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  printf("{\"piece_bytes\": %d, \"F\": %d, \"ctas\": %d, \"GBs\": %.1f}\n", PD * 8, F, gx * nby, (double)ntiles * E * nby * PD * 8 * 10 / (ms * 1e-3) / 1e9);
}
int main() {
  const int F = 96; const long n = 640L * 444 * 4;   // rows of 768 bytes (a multiple of every piece length tried), 873 MB: larger than L2
  double* X; double* sink; cudaMalloc(&X, n * F * 8); cudaMalloc(&sink, 8); cudaMemset(X, 0, n * F * 8);
  run<8>(X, F, n, sink); run<16>(X, F, n, sink); run<32>(X, F, n, sink); run<96>(X, F, n, sink);
  return 0;
}
