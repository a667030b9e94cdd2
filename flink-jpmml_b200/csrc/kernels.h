// kernels.h -- launch interface between the engine (host) and the CUDA kernels.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "model.h"

namespace fjp {

// One micro-batch.  All pointers are DEVICE pointers.
struct BatchArgs {
  const double* X;           // n x F row-major (the Flink Vector payloads, api/converter/VectorConverter.scala:68-86)
  const uint32_t* present;   // n x present_stride words, bit i of row = slot i present (sparse-absent = 0); NULL = dense
  const double* replace;     // n, NaN = none (replaceNan of PmmlModel.predict, api/PmmlModel.scala:109); NULL = none
  long n;
  int present_stride;
  double* score;             // n   Score.value (NaN when status != 0)
  uint8_t* status;           // n   0 = Score, else EmptyScore with the reason (model.h ST_*)
  int32_t* aux;              // n   optional: 1-based cluster position / node id (single tree) / winning label index
  int32_t* node_ids;         // n x node_stride optional: per-segment node ids of an ensemble
  int node_stride;
  double* probs;             // n x prob_stride optional: class probabilities (regression, NN) / winning distance
  int prob_stride;
  // re-scoring of selected rows (the exact FP64 kernels behind the tensor-core path): when set, work item i is row row_map[i]
  // and the number of work items is *n_dev (both in device memory, written by an earlier kernel on the same stream)
  const int32_t* row_map;
  const int32_t* n_dev;
  unsigned long long* stats; // optional device counters: [0] events that took the tensor-core path, [1] of those re-scored in FP64
};

struct ForestLaunch {  // tile geometry of forest_complete_kernel
  int events_per_cta, trees_per_stage, row_stride, trees_interleaved;
  uint32_t off_ring, off_rows, off_votes, off_bars, off_fail, smem_bytes;
  int slices; double* leaf_out;   // small-batch latency path: tree groups split over gridDim.y, leaf values to scratch
};

bool plan_forest_complete(const DeviceModel& m, ForestLaunch* out);
cudaError_t launch_score(const DeviceModel& m, const BatchArgs& a, cudaStream_t s, int* launches, void* slice_scratch = nullptr);
size_t forest_slice_scratch_bytes(const DeviceModel& m, long n);     // work area of either forest side path (0: none wanted)
size_t forest_compact_scratch_bytes(const DeviceModel& m, long n);   // throughput path only
cudaError_t launch_expand_sparse(const long long* row_ptr, const int32_t* idx, const double* val, long n, int F,
                                 double* X, uint32_t* present, cudaStream_t s, int* launches);
cudaError_t launch_widen_f32(const float* src, double* dst, long count, cudaStream_t s, int* launches);   // src, dst 16-byte aligned
size_t derived_scratch_bytes(const DeviceModel& m, long n);
size_t sliced_scratch_bytes(const DeviceModel& m, const BatchArgs& a);
size_t dense_fma_scratch_bytes(const DeviceModel& m, const BatchArgs& a);  // FMA paths of the dense tile kernel (0: not taken for this batch)   // tensor-core path of dense layers (0: not taken for this batch)
cudaError_t launch_score_any(const DeviceModel& m, const BatchArgs& a, void* scratch, cudaStream_t s, int* launches);
cudaError_t launch_fill_status(const BatchArgs& a, int status, cudaStream_t s, int* launches);

}  // namespace fjp
