// kernels_internal.h -- what the kernel translation units (kernels.cu, sliced_gemm.cu) share on the host side.
#pragma once
#include <cuda_runtime.h>

#include "kernels.h"

namespace fjp {
void device_facts(int* num_sms, int* smem_optin);    // of the current device (cached per device ordinal)
int env_int(const char* name, int dflt);
cudaError_t launch_neural_pair(const DeviceModel& m, const BatchArgs& a, cudaStream_t s);   // FP64 MLP tiles; honours a.row_map / a.n_dev
// tensor-core path of dense layers: false = not taken for this model / batch (nothing launched); true = launched, *err says how it went
bool launch_sliced(const DeviceModel& m, const BatchArgs& a, void* scratch, cudaStream_t s, int* launches, cudaError_t* err);
// fused tensor-core layer for softmax RegressionModels over plain dense rows: takes(): is this batch for it; launch: events it cannot
// decide go on the list (count + rows)
bool sliced_dense_takes(const DeviceModel& m, const BatchArgs& a);
cudaError_t launch_sliced_dense(const DeviceModel& m, const BatchArgs& a, int32_t* fb_count, int32_t* fb_rows, cudaStream_t s);
}  // namespace fjp
