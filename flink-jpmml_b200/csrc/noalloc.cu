// noalloc.cu -- fjp_score_sparse_noalloc: sparse (CSR) scoring for callers that must not allocate device memory.
//
// Who needs this.  fjp_score_sparse / fjp_score_batch stage their rows in per-stream arenas that the engine grows with
// cudaMalloc (and shrinks with cudaFree) whenever a bigger batch shows up.  Both calls can block until the whole GPU is
// idle, and both can simply fail when another tenant of the device holds the rest of HBM.  A Flink subtask that shares
// its GPU with other jobs under a fixed memory budget, or that sits on a latency-bound path where a device-wide stall in
// the middle of a micro-batch is not acceptable, therefore wants a scoring call that touches no allocator at all: it
// brings its own (pre-allocated) input and output buffers, and the one piece of working memory the path needs -- the
// dense rows + present bits that the CSR rows are expanded into before the model kernel reads them -- is reserved inside
// the library image itself.
//
// That reservation is the file-scope array below: FJP_NOALLOC_SCRATCH_BYTES (8 MiB) of device memory on EVERY GPU that
// loads libfjpmml.so, held until the process exits.  A call fits when  n * (vec_size*8 + ceil(vec_size/32)*4) <= 8 MiB
// (about 10 000 rows of 100 fields); anything larger is refused with an error, never truncated.
//
// The array is one shared work area, not a per-call buffer: it still holds whatever the previous call expanded, and two
// calls whose kernels overlap on the GPU (different streams, or different host threads) overwrite each other's rows.
// Serialise calls on one device -- one stream, or wait for the previous call's work before issuing the next.  Nothing
// has to be cleared between calls: the expansion kernel rewrites every value and every present word of the rows it uses.
//
// Set-up: the first call on a device looks up the device address of the array (once per device, cached).  In host code
// the array's name only stands for the symbol; it is not an address a kernel can use, so the kernels always get the
// looked-up pointer.  Make that first call during start-up, before any stream capture begins on that device.
//
// This entry point has no counterpart in the reference.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "engine_internal.h"
#include "kernels.h"

using namespace fjp;

static __device__ __align__(256) unsigned char g_noalloc_scratch[FJP_NOALLOC_SCRATCH_BYTES];

static constexpr int kMaxDevices = 64;
static unsigned char* g_scratch_addr[kMaxDevices];   // device address of g_noalloc_scratch, per device ordinal
static std::mutex g_scratch_mu;

static int noalloc_fail(const std::string& msg) { internal_set_error(msg.c_str()); return -1; }

// Looks the array's device address up once per device; `device` must be the current device.
static cudaError_t scratch_for_device(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_addr[device]) {
    void* p = nullptr;
    cudaError_t ce = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (ce != cudaSuccess) return ce;
    g_scratch_addr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_addr[device];
  return cudaSuccess;
}

extern "C" int fjp_score_sparse_noalloc(fjp_engine* e, fjp_model* m, const int64_t* d_row_ptr, const int32_t* d_indices,
                                        const double* d_values, long n, int vec_size, double* d_score, uint8_t* d_status,
                                        void* cuda_stream) {
  if (n <= 0) return 0;
  if (!e || !m) return noalloc_fail("fjp_score_sparse_noalloc: engine and model are required");
  const DeviceModel* dm = internal_device_model(m);
  if (!dm) return noalloc_fail("fjp_score_sparse_noalloc: the model is not resident on a device");
  const int device = fjp_engine_device(e);
  if (internal_model_device(m) != device) return noalloc_fail("fjp_score_sparse_noalloc: the model lives on another device");
  if (device < 0 || device >= kMaxDevices) return noalloc_fail("fjp_score_sparse_noalloc: device ordinal out of range");
  if (vec_size != internal_model_fields(m))
    return noalloc_fail("fjp_score_sparse_noalloc: vec_size differs from the model's number of active fields");
  if (internal_model_derived(m) > 0)
    return noalloc_fail("fjp_score_sparse_noalloc: models with derived fields need a larger work area than the built-in one; use fjp_score_sparse");
  const int W = (vec_size + 31) / 32;
  const size_t x_bytes = (size_t)n * vec_size * 8, p_bytes = (size_t)n * W * 4;   // rows are 8-byte, present words 4-byte aligned: no padding, the documented bound is exact
  if (x_bytes + p_bytes > (size_t)FJP_NOALLOC_SCRATCH_BYTES)
    return noalloc_fail("fjp_score_sparse_noalloc: batch needs " + std::to_string(x_bytes + p_bytes) + " bytes of scratch, the built-in area holds " +
                        std::to_string((size_t)FJP_NOALLOC_SCRATCH_BYTES) + "; split the batch (rows * (vec_size*8 + ceil(vec_size/32)*4) must fit)");
  cudaError_t ce = cudaSetDevice(device);
  if (ce != cudaSuccess) return noalloc_fail(std::string("cudaSetDevice: ") + cudaGetErrorString(ce));
  unsigned char* scratch = nullptr;
  ce = scratch_for_device(device, &scratch);
  if (ce != cudaSuccess) return noalloc_fail(std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(ce));
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  double* X = reinterpret_cast<double*>(scratch);
  uint32_t* present = reinterpret_cast<uint32_t*>(scratch + x_bytes);
  int launches = 0;
  ce = launch_expand_sparse(reinterpret_cast<const long long*>(d_row_ptr), d_indices, d_values, n, vec_size, X, present, st, &launches);
  if (ce != cudaSuccess) return noalloc_fail(std::string("expand kernel: ") + cudaGetErrorString(ce));
  BatchArgs a{};
  a.X = X; a.present = present; a.present_stride = W; a.n = n; a.score = d_score; a.status = d_status;
  ce = launch_score(*dm, a, st, &launches);
  internal_count_launches(e, launches);
  if (ce != cudaSuccess) return noalloc_fail(std::string("scoring kernel: ") + cudaGetErrorString(ce));
  return 0;
}
