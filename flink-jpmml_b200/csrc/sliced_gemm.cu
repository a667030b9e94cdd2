// sliced_gemm.cu -- translation unit of the tensor-core path of dense layers (kernels in sliced_gemm.cuh) and its launch plan.
#include <cuda_runtime.h>

#include <cstdint>

#include "device_common.cuh"
#include "kernels_internal.h"
#include "sliced_gemm.cuh"

namespace fjp {

static thread_local int g_num_sms = 0, g_smem_optin = 0;

// ---- tensor-core path of dense layers (sliced_gemm.cuh): geometry, work area, launch ---------------------------------------
// FJP_SLICED=0 switches the path off, =2 takes it for every batch size (tests); by default batches of at least one tile per two
// SMs take it, smaller ones stay on the FP64 tiles (a handful of tiles cannot hide the four launches of a pass).
struct SlicedPlan { long pass_rows, pass_tiles; size_t off_rows[2], off_a[MAX_LAYERS], off_fb_rows, bytes; };
static bool plan_sliced(const DeviceModel& m, const BatchArgs& a, SlicedPlan* out) {
  if (m.kind != MK_NEURAL || !m.nn.sg.enabled || a.probs || a.node_ids || a.row_map) return false;
  const int mode = env_int("FJP_SLICED", 1);
  if (!mode) return false;
  device_facts(&g_num_sms, &g_smem_optin);
  if (mode != 2 && a.n < (long)SG_M * g_num_sms / 2) return false;
  const SgDesc& sg = m.nn.sg;
  for (int l = 0; l < sg.n_layers; l++) if (sg::smem_plan(sg.layers[l].Kp / SG_KC, sg.layers[l].Np).bytes > (uint32_t)g_smem_optin) return false;
  SlicedPlan p{}; const long tiles = (a.n + SG_M - 1) / SG_M, cap = (long)env_int("FJP_SG_PASS_TILES", (int)sg::PASS_TILES_PER_SM) * g_num_sms;
  p.pass_tiles = tiles < cap ? tiles : cap; p.pass_rows = p.pass_tiles * SG_M;
  size_t off = 256;                                            // [0]: the re-scoring counter
  for (int k = 0; k < 2; k++) { p.off_rows[k] = off; off += ((size_t)p.pass_rows * sizeof(SgRow) + 255) & ~size_t(255); }
  for (int l = 0; l < sg.n_layers; l++) { p.off_a[l] = off; off += (size_t)p.pass_tiles * (sg.layers[l].Kp / SG_KC) * sg::A_CHUNK; }
  p.off_fb_rows = off; off += ((size_t)a.n * 4 + 255) & ~size_t(255);
  p.bytes = off; *out = p; return true;
}
size_t sliced_scratch_bytes(const DeviceModel& m, const BatchArgs& a) { SlicedPlan p; return plan_sliced(m, a, &p) ? p.bytes : 0; }

static cudaError_t launch_sliced_plan(const DeviceModel& m, const BatchArgs& a, const SlicedPlan& p, void* scratch, cudaStream_t s, int* launches) {
  uint8_t* base = static_cast<uint8_t*>(scratch); const SgDesc& sg = m.nn.sg;
  int32_t* fb_count = reinterpret_cast<int32_t*>(base); int32_t* fb_rows = reinterpret_cast<int32_t*>(base + p.off_fb_rows);
  cudaError_t e = cudaMemsetAsync(fb_count, 0, 4, s); if (e != cudaSuccess) return e;
#define FJP_SG_ATTR(MM) cudaFuncSetAttribute(sg::sg_layer_kernel<sg::MM>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin)
  FJP_SG_ATTR(M_RELU); FJP_SG_ATTR(M_IDENT); FJP_SG_ATTR(M_TANH); FJP_SG_ATTR(M_LOGISTIC); FJP_SG_ATTR(M_LAST);
#undef FJP_SG_ATTR
  int nl = 0;
  for (long r0 = 0; r0 < a.n; r0 += p.pass_rows) {
    sg::Pass ps{}; ps.row_begin = r0; ps.n_rows = a.n - r0 < p.pass_rows ? a.n - r0 : p.pass_rows; ps.fb_count = fb_count; ps.fb_rows = fb_rows;
    const long tiles = (ps.n_rows + SG_M - 1) / SG_M;
    ps.a_out = reinterpret_cast<int8_t*>(base + p.off_a[0]); ps.rows_out = reinterpret_cast<SgRow*>(base + p.off_rows[0]);
    const bool plain = sg.inputs_identity && m.prep_identity && a.present == nullptr && a.replace == nullptr;
    if (plain) sg::sg_slice_inputs_kernel<true><<<(unsigned)((ps.n_rows * 4 + 255) / 256), 256, 0, s>>>(m, a, ps);
    else sg::sg_slice_inputs_kernel<false><<<(unsigned)((ps.n_rows * 4 + 255) / 256), 256, 0, s>>>(m, a, ps);
    e = cudaGetLastError(); if (e != cudaSuccess) return e;
    nl++;
    for (int l = 0; l < sg.n_layers; l++) {
      const bool last = l == sg.n_layers - 1;
      ps.layer = l; ps.a_in = reinterpret_cast<const int8_t*>(base + p.off_a[l]); ps.a_out = last ? nullptr : reinterpret_cast<int8_t*>(base + p.off_a[l + 1]);
      ps.rows_in = reinterpret_cast<const SgRow*>(base + p.off_rows[l & 1]); ps.rows_out = reinterpret_cast<SgRow*>(base + p.off_rows[(l & 1) ^ 1]);
      const uint32_t smem = sg::smem_plan(sg.layers[l].Kp / SG_KC, sg.layers[l].Np).bytes;
      const int grid = (int)(tiles < g_num_sms ? tiles : g_num_sms);
      const int act = sg.layers[l].act;
      if (last) sg::sg_layer_kernel<sg::M_LAST><<<grid, sg::THREADS, smem, s>>>(m, a, ps);
      else if (act == ACT_RECTIFIER) sg::sg_layer_kernel<sg::M_RELU><<<grid, sg::THREADS, smem, s>>>(m, a, ps);
      else if (act == ACT_IDENTITY) sg::sg_layer_kernel<sg::M_IDENT><<<grid, sg::THREADS, smem, s>>>(m, a, ps);
      else if (act == ACT_TANH) sg::sg_layer_kernel<sg::M_TANH><<<grid, sg::THREADS, smem, s>>>(m, a, ps);
      else sg::sg_layer_kernel<sg::M_LOGISTIC><<<grid, sg::THREADS, smem, s>>>(m, a, ps);
      e = cudaGetLastError(); if (e != cudaSuccess) return e;
      nl++;
    }
  }
  // events the margin test (or an irregular input) handed back: the FP64 kernel in JPMML's order, row by row from the list
  BatchArgs fb = a; fb.row_map = fb_rows; fb.n_dev = fb_count;
  e = launch_neural_pair(m, fb, s); if (e != cudaSuccess) return e;
  if (launches) *launches += nl;   // (launch_score already counted one)
  return cudaSuccess;
}

// fused single dense layer (softmax RegressionModel over plain dense rows, sliced_gemm.cuh namespace dense)
bool sliced_dense_takes(const DeviceModel& m, const BatchArgs& a) {
  if (m.kind != MK_REGRESSION || !m.reg.sg.enabled || a.probs || a.node_ids || a.row_map || a.present || a.replace || !m.prep_identity || (m.F & 1)) return false;
  const int mode = env_int("FJP_SLICED", 1);
  if (!mode) return false;
  device_facts(&g_num_sms, &g_smem_optin);
  if (mode != 2 && a.n < (long)SG_M * g_num_sms / 2) return false;
  return sg::dense::smem_plan().bytes <= (uint32_t)g_smem_optin;
}
cudaError_t launch_sliced_dense(const DeviceModel& m, const BatchArgs& a, int32_t* fb_count, int32_t* fb_rows, cudaStream_t s) {
  device_facts(&g_num_sms, &g_smem_optin);
  const uint32_t smem = sg::dense::smem_plan().bytes;
  cudaFuncSetAttribute(sg::dense::sg_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const long tiles = (a.n + SG_M - 1) / SG_M; const int grid = (int)(tiles < g_num_sms ? tiles : g_num_sms);
  sg::dense::sg_dense_kernel<<<grid, sg::dense::THREADS, smem, s>>>(m, a, fb_count, fb_rows);
  return cudaGetLastError();
}

bool launch_sliced(const DeviceModel& m, const BatchArgs& a, void* scratch, cudaStream_t s, int* launches, cudaError_t* err) {
  SlicedPlan p;
  if (!plan_sliced(m, a, &p)) return false;
  *err = launch_sliced_plan(m, a, p, scratch, s, launches);
  return true;
}

}  // namespace fjp
