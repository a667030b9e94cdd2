// device_common.cuh -- device helpers shared by the kernel translation units (kernels.cu, sliced_gemm.cu): Java comparison
// semantics, input preparation (JPMML FieldValueUtil.prepareInputValue), result writers, link functions, key encodings,
// mbarrier / bulk-copy wrappers, NeuralNetwork activations.  Everything is inline: each unit compiles its own copy.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.h"

namespace fjp {

// ================================================================================================ helpers
__device__ __forceinline__ bool java_eq(double a, double b) {  // Double.equals
  return __double_as_longlong(a) == __double_as_longlong(b) || (a != a && b != b);
}
__device__ __forceinline__ int java_cmp(double a, double b) {  // Double.compare
  if (a < b) return -1;
  if (a > b) return 1;
  if (a == b) {
    if (a == 0.0) { long long x = __double_as_longlong(a), y = __double_as_longlong(b); return x == y ? 0 : (x < 0 ? -1 : 1); }
    return 0;
  }
  bool an = a != a, bn = b != b;
  return an == bn ? 0 : (an ? 1 : -1);
}
__device__ __forceinline__ bool total_less(double a, double b) { return java_cmp(a, b) < 0; }

// JPMML FieldValueUtil.prepareInputValue for a Double input (SURVEY.md App. A "Input preparation").
// returns 0: value in `out`; 1: missing; 2: preparation failure (InputPreparationException)
__device__ inline int prep_value(const PrepField& pf, const double* __restrict__ pool, bool present, double raw, double& out) {
  if (present) {
    bool castable = true; double v = raw;
    switch (pf.dtype) {
      case DT_DOUBLE: break;
      case DT_FLOAT: v = (double)(float)raw; break;
      case DT_INTEGER: castable = isfinite(raw) && raw == floor(raw) && fabs(raw) <= 2147483647.0; break;
      case DT_BOOLEAN: castable = raw == 0.0 || raw == 1.0; break;
      default: castable = false;  // string & friends: a Double cannot be cast (api/PmmlModelSpec.scala:131-136)
    }
    int st = 0;  // 0 valid, 1 invalid, 2 missing
    if (!castable) st = 1;
    else {
      const double* p = pool + pf.pool_off + 3 * pf.n_intervals;
      bool matched = false;
      for (int i = 0; i < pf.n_missing; i++) if (java_eq(v, p[pf.n_valid + pf.n_invalid + i])) { st = 2; matched = true; break; }
      if (!matched) for (int i = 0; i < pf.n_invalid; i++) if (java_eq(v, p[pf.n_valid + i])) { st = 1; matched = true; break; }
      if (!matched && pf.has_valid_space) {
        bool ok = false;
        for (int i = 0; i < pf.n_valid; i++) if (java_eq(v, p[i])) { ok = true; break; }
        st = ok ? 0 : 1; matched = true;
      }
      if (!matched && pf.continuous && pf.n_intervals) {
        const double* iv = pool + pf.pool_off; bool in = false;
        for (int i = 0; i < pf.n_intervals; i++) {
          int fl = (int)iv[3 * i + 2];
          bool lo = !(fl & 1) || ((fl & 4) ? v > iv[3 * i] : v >= iv[3 * i]);
          bool hi = !(fl & 2) || ((fl & 8) ? v < iv[3 * i + 1] : v <= iv[3 * i + 1]);
          if (lo && hi) { in = true; break; }
        }
        st = in ? 0 : 1;
      }
    }
    if (st == 0) {
      if (pf.outlier != OT_AS_IS) {
        if (pf.has_low && v < pf.low) { if (pf.outlier == OT_AS_MISSING) goto missing; v = pf.low; }
        else if (pf.has_high && v > pf.high) { if (pf.outlier == OT_AS_MISSING) goto missing; v = pf.high; }
      }
      out = v; return 0;
    }
    if (st == 1) {
      if (pf.ivt == IVT_RETURN_INVALID) return 2;
      if (pf.ivt == IVT_AS_IS) { if (!castable) return 2; out = v; return 0; }
    }
  }
missing:
  if (pf.has_repl == 1) { out = pf.repl; return 0; }
  return pf.has_repl == 2 ? 2 : 1;
}

// one event's slot f: applies replaceNan (api/PmmlModel.scala:148) then the preparation above
__device__ __forceinline__ int load_field(const DeviceModel& m, const BatchArgs& a, long row, int f, double& out) {
  double raw = a.X[row * (long)m.F + f];
  bool present = a.present ? ((a.present[row * (long)a.present_stride + (f >> 5)] >> (f & 31)) & 1u) : true;
  if (!present && a.replace) { double r = a.replace[row]; if (r == r) { raw = r; present = true; } }
  if (m.prep_simple) {
    if (!present) return 1;
    out = m.prep[f].dtype == DT_FLOAT ? (double)(float)raw : raw;
    return 0;
  }
  return prep_value(m.prep[f], m.prep_pool, present, raw, out);
}

// The reference prepares EVERY active field before it evaluates the model (S/api/PmmlModel.scala:143-152: activeFields.map
// { EvaluatorUtil.prepare }), so an invalid value fails the event even in a field the model never looks at.  Kernels that
// only load the fields their model references ask this first.  Models whose fields cannot fail (prep_simple) skip it.
__device__ __noinline__ inline bool row_fails_preparation(const DeviceModel& m, const BatchArgs& a, long row) {
  if (m.prep_simple) return false;
  for (int f = 0; f < m.F; f++) { double v; if (load_field(m, a, row, f, v) == 2) return true; }
  return false;
}

__device__ __forceinline__ double apply_target(const TargetXform& t, double v) {
  if (t.has_min && v < t.mn) v = t.mn;
  if (t.has_max && v > t.mx) v = t.mx;
  v = v * t.factor + t.constant;
  if (t.cast == 1) v = floor(v + 0.5); /* Math.round: half goes up, also below zero */ else if (t.cast == 2) v = ceil(v); else if (t.cast == 3) v = floor(v);
  return v;
}

__device__ __forceinline__ void write_out(const BatchArgs& a, long row, double score, int status, int aux) {
  a.score[row] = status == ST_OK ? score : __longlong_as_double(0x7FF8000000000000LL);
  a.status[row] = (uint8_t)status;
  if (a.aux) a.aux[row] = aux;
}
// classification / clustering winner -> (score, status) through the label table (Pipeline.scala:94-98)
__device__ __forceinline__ void write_label(const DeviceModel& m, const BatchArgs& a, long row, int label, int aux) {
  int st = m.target_status ? m.target_status : m.label_status[label];
  write_out(a, row, m.label_value[label], st, aux);
}

// ---- link functions and class normalisation (RegressionModel; also the modelChain epilogue) ------------------
__device__ __forceinline__ double link_fn(int norm, double y) {
  switch (norm) {
    case NORM_NONE: return y;
    case NORM_SOFTMAX: case NORM_LOGIT: return 1.0 / (1.0 + exp(-y));
    case NORM_EXP: return exp(y);
    case NORM_PROBIT: return 0.5 * erfc(-y / sqrt(2.0));
    case NORM_CLOGLOG: return 1.0 - exp(-exp(y));
    case NORM_LOGLOG: return exp(-exp(-y));
    case NORM_CAUCHIT: return 0.5 + atan(y) / 3.14159265358979323846;
    default: return y;
  }
}
__device__ __forceinline__ double ipow(double x, int e) {  // Math.pow(x, integer) for the small exponents PMML uses
  if (e == 1) return x; if (e == 0) return 1.0; if (e == 2) return x * x;
  return pow(x, (double)e);
}

// classification epilogue shared by regression / NN: y[0..C) in (strided) memory -> probabilities -> first maximal label
template <class Get, class Set>
__device__ void normalize_classes(int norm, int C, Get y, Set p) {
  if (norm == NORM_SOFTMAX) { double s = 0; for (int i = 0; i < C; i++) { double e = exp(y(i)); p(i, e); s += e; } for (int i = 0; i < C; i++) p(i, y(i) / s); }
  else if (norm == NORM_SIMPLEMAX) { double s = 0; for (int i = 0; i < C; i++) s += y(i); for (int i = 0; i < C; i++) p(i, y(i) / s); }
  else if (norm == NORM_NONE) { double s = 0; for (int i = 0; i + 1 < C; i++) s += y(i); p(C - 1, 1.0 - s); }
  else { double s = 0; for (int i = 0; i + 1 < C; i++) { double v = link_fn(norm, y(i)); p(i, v); s += v; } p(C - 1, 1.0 - s); }
}


// modelChain epilogue: RegressionTables over the channel values of the tree segments (RegTerm.slot = channel)
template <class GetCh>
__device__ void chain_finish(const DeviceModel& m, const BatchArgs& a, long row, GetCh ch) {
  const RegressionDesc& rd = m.reg; double y[64];
  for (int t = 0; t < rd.n_tables; t++) {
    double acc = rd.intercept[t];
    for (int i = rd.term_off[t]; i < rd.term_off[t + 1]; i++) { RegTerm tm = rd.terms[i]; acc += tm.coef * ipow(ch(tm.slot), tm.exponent); }
    y[t] = acc;
  }
  if (!rd.classify) { write_out(a, row, apply_target(m.tx, link_fn(rd.norm, y[0])), m.target_status, -1); return; }
  const int C = rd.n_tables;
  normalize_classes(rd.norm, C, [&](int i) { return y[i]; }, [&](int i, double v) { y[i] = v; });
  int best = 0; for (int i = 1; i < C; i++) if (java_cmp(y[i], y[best]) > 0) best = i;
  write_label(m, a, row, best, best);
  if (a.probs) for (int i = 0; i < C && i < a.prob_stride; i++) a.probs[row * (long)a.prob_stride + i] = y[i];
}
__device__ __forceinline__ double channel_value(const ChannelSpec& cs, double sum) {
  return apply_target(cs.tx, cs.average ? sum / (double)cs.n_trees : sum);
}

// ---- key encodings for tree tests (model.h ForestDesc) -------------------------------------------------
__device__ __forceinline__ uint32_t ordkey(float f) {
  uint32_t u = __float_as_uint(f);
  return (f != f) ? 0xFFC00000u : ((u & 0x80000000u) ? ~u : (u | 0x80000000u));
}
__device__ __forceinline__ uint32_t enc_float(double v, int mode) {
  float f = mode == 1 ? __double2float_ru(v) : (mode == 2 ? __double2float_rd(v) : (float)v);
  return ordkey(f);
}
__device__ __forceinline__ uint32_t enc_rank(double v, const double* __restrict__ t, int n) {
  int lo = 0, hi = n;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (total_less(t[mid], v)) lo = mid + 1; else hi = mid; }
  bool eq = lo < n && !total_less(v, t[lo]);
  return eq ? 2u * lo + 2u : 2u * lo + 1u;
}
__device__ __forceinline__ uint32_t encode_key(const ForestDesc& fd, int f, double v) {
  if (fd.enc == ENC_FLOAT) return enc_float(v, fd.feat_mode[f]);
  return enc_rank(v, fd.rank_thr + fd.rank_off[f], fd.rank_off[f + 1] - fd.rank_off[f]);
}

// ---- mbarrier + bulk async copy (TMA unit, 1-D) ----------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst_smem, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- NeuralNetwork activations and NormContinuous (J6) -------------------------------------------------------------
__device__ __forceinline__ double activate(int act, double z, double thr) {
  switch (act) {
    case ACT_THRESHOLD: return z > thr ? 1.0 : 0.0;
    case ACT_LOGISTIC: return 1.0 / (1.0 + exp(-z));
    case ACT_TANH: return tanh(z);
    case ACT_IDENTITY: return z;
    case ACT_EXPONENTIAL: return exp(z);
    case ACT_RECIPROCAL: return 1.0 / z;
    case ACT_SQUARE: return z * z;
    case ACT_GAUSS: return exp(-(z * z));
    case ACT_SINE: return sin(z);
    case ACT_COSINE: return cos(z);
    case ACT_ELLIOTT: return z / (1.0 + fabs(z));
    case ACT_ARCTAN: return 2.0 * atan(z) / 3.14159265358979323846;
    default: return fmax(0.0, z);  // rectifier
  }
}
__device__ inline double norm_cont(const double* __restrict__ pts, int n, int outliers, double x, bool& missing) {
  auto interp = [&](int i) { double a0 = pts[2 * i], b0 = pts[2 * i + 1], a1 = pts[2 * i + 2], b1 = pts[2 * i + 3]; return b0 + (x - a0) / (a1 - a0) * (b1 - b0); };
  if (x < pts[0]) { if (outliers == 1) { missing = true; return 0; } if (outliers == 2) return pts[1]; return interp(0); }
  if (x > pts[2 * (n - 1)]) { if (outliers == 1) { missing = true; return 0; } if (outliers == 2) return pts[2 * (n - 1) + 1]; return interp(n - 2); }
  int i = 0; while (i + 2 < n && x > pts[2 * (i + 1)]) i++;
  return interp(i);
}

}  // namespace fjp
