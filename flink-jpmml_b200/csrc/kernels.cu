// kernels.cu -- the scoring kernels (sm_100a).  Together they replace, for a whole micro-batch at once, what the
// reference does per event in PmmlModel.predict
//   flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/PmmlModel.scala:109-174
// i.e. prepareInput (:143-152, JPMML EvaluatorUtil.prepare), evaluateInput (:159-160, JPMML ModelEvaluator.evaluate
// for ClusteringModel / TreeModel / MiningModel / RegressionModel / NeuralNetwork) and extractTarget (:167-174,
// Pipeline.scala:61-98).  Input preparation and target extraction are fused into every kernel.
//
// Arithmetic follows the JVM: double accumulation in document order, no fused multiply-add (built with -fmad=false),
// comparisons in java.lang.Double.compareTo's total order where JPMML compares through Comparable.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <mutex>

#include "device_common.cuh"
#include "kernels.h"
#include "kernels_internal.h"

namespace fjp {

// ================================================================================================ forest, complete trees
// One CTA scores a tile of E = blockDim.x events (one event per thread) against ALL trees:
//   phase A  rows are read coalesced from HBM (software-pipelined through registers), prepared, turned into 32-bit
//            order keys and written TRANSPOSED into shared memory (keys[f][e]) so that the per-lane gather in phase B
//            is bank-conflict free;
//   phase B  tree blobs (nodes+leaves, 2^(D+1) x 8 B each) stream through a 2-stage shared-memory ring filled by
//            bulk async copies (cp.async.bulk + mbarrier) from the L2-resident model; every warp walks the same tree
//            for 32 different events, TI trees interleaved per thread for ILP; leaf values accumulate in double in
//            segment order (bit-identical to a sequential sum).
// Shared memory:  keys F*E*4 | ring 2*G*blob (phase A's row stage aliases it when it fits) | votes | barriers
// MODE: 0 sum/average, 1 weighted sum/average, 2 majority vote, 3 modelChain (per-channel sums -> RegressionTables).
// DBG: also write per-tree node ids.
constexpr int FC_RMAX = 10;  // row-stage elements a thread keeps in flight
template <int D, int TI, int MODE, bool DBG, bool SLICE = false>
__global__ void __launch_bounds__(512, 1)
forest_complete_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const ForestLaunch fl) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ForestDesc& fd = m.forest;
  const int E = blockDim.x, F = m.F, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = E >> 5;
  const int depth = D ? D : fd.depth;
  const uint32_t blob_words = 2u << depth, blob_bytes = blob_words * 8u;
  const int G = fl.trees_per_stage, S = fl.row_stride;
  uint32_t* keys = reinterpret_cast<uint32_t*>(smem);
  uint64_t* ring = reinterpret_cast<uint64_t*>(smem + fl.off_ring);
  double* xs = reinterpret_cast<double*>(smem + fl.off_rows);
  uint16_t* votes = reinterpret_cast<uint16_t*>(smem + fl.off_votes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + fl.off_bars);
  uint8_t* row_fail = smem + fl.off_fail;
  const int T = fd.n_trees, NG = (T + G - 1) / G;
  const long ntiles = (a.n + E - 1) / E;
  // latency path for small batches: gridDim.y CTAs share one tile, each walks a slice of the tree groups and writes the
  // leaf values to fl.leaf_out[row][tree]; forest_ordered_sum_kernel then adds them in segment order (same result bits)
  const int g_lo = SLICE ? (int)((long)blockIdx.y * NG / gridDim.y) : 0, g_hi = SLICE ? (int)((long)(blockIdx.y + 1) * NG / gridDim.y) : NG;
  double* const leaf_out = SLICE ? fl.leaf_out : nullptr;
  const bool rows_alias_ring = fl.off_rows == fl.off_ring;
  const uint32_t inv_f = 0xFFFFFFFFu / (uint32_t)F + 1u;  // i / F == umulhi(i, inv_f) for i < 2^16

  // bars: [0],[1] ring slot filled (bulk-copy bytes), [2],[3] ring slot drained (one arrival per warp)
  if (tid == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], nwarps); mbar_init(&bars[3], nwarps); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();

  uint32_t q_issue = 0, q_wait = 0;  // running stage counters (producer: thread 0 only)
  auto issue = [&](int g) {  // thread 0: fill ring slot q_issue&1 with tree group g
    uint32_t slot = q_issue & 1u; int t0 = g * G; int nt = min(G, T - t0); uint32_t bytes = (uint32_t)nt * blob_bytes;
    mbar_expect_tx(&bars[slot], bytes);
    bulk_g2s(ring + (size_t)slot * G * blob_words, fd.blobs + (size_t)t0 * blob_words, bytes, &bars[slot]);
    q_issue++;
  };

  for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long row0 = tile * E;
    if (!rows_alias_ring && tid == 0) { fence_proxy_async(); issue(g_lo); if (g_lo + 1 < g_hi) issue(g_lo + 1); }
    row_fail[tid] = 0;
    // ---------------- phase A: load, prepare, encode, transpose (32 rows per chunk; lane <-> row when encoding)
    {
      double v[FC_RMAX];
      uint32_t pw_next[4] = {~0u, ~0u, ~0u, ~0u}, pw_cur[4];   // this lane's row: present words (rows with <= 128 fields)
      const bool pw_regs = a.present != nullptr && a.present_stride <= 4;
      auto chunk_cnt = [&](int rb) -> int { long left = a.n - (row0 + rb); int rows = left >= 32 ? 32 : (left > 0 ? (int)left : 0); return rows * F; };
      auto load_chunk = [&](int rb) {
        const double* src = a.X + (row0 + rb) * (long)F; const int cnt = chunk_cnt(rb);
#pragma unroll
        for (int k = 0; k < FC_RMAX; k++) { int i = tid + k * E; if (i < cnt) v[k] = __ldcs(src + i); }
        if (pw_regs && row0 + rb + lane < a.n) {
          const uint32_t* pp = a.present + (row0 + rb + lane) * (long)a.present_stride;
#pragma unroll
          for (int w = 0; w < 4; w++) if (w < a.present_stride) pw_next[w] = __ldg(pp + w);
        }
      };
      load_chunk(0);
      for (int rb = 0; rb < E; rb += 32) {
        const int cnt = chunk_cnt(rb);
#pragma unroll
        for (int k = 0; k < FC_RMAX; k++) {
          int i = tid + k * E;
          if (i < cnt) { int r = (int)__umulhi((uint32_t)i, inv_f); xs[r * S + (i - r * F)] = v[k]; }
        }
        {  // wide rows: the part that does not fit the register batch
          const double* src = a.X + (row0 + rb) * (long)F;
          for (int i = tid + FC_RMAX * E; i < cnt; i += E) { int r = (int)__umulhi((uint32_t)i, inv_f); xs[r * S + (i - r * F)] = __ldcs(src + i); }
        }
        __syncthreads();
#pragma unroll
        for (int w = 0; w < 4; w++) pw_cur[w] = pw_next[w];
        if (rb + 32 < E) load_chunk(rb + 32);  // in flight while this chunk is encoded
        long row = row0 + rb + lane;
        if (row < a.n) {
          double rep = 0; bool has_rep = false;
          if (a.replace) { rep = a.replace[row]; has_rep = rep == rep; }
          uint32_t pw = 0xFFFFFFFFu; int pw_idx = -1;
          for (int f = warp; f < F; f += nwarps) {
            double raw = xs[lane * S + f]; bool present = true;
            if (pw_regs) { int wi = f >> 5; uint32_t wv = wi == 0 ? pw_cur[0] : (wi == 1 ? pw_cur[1] : (wi == 2 ? pw_cur[2] : pw_cur[3])); present = (wv >> (f & 31)) & 1u; }
            else if (a.present) { int wi = f >> 5; if (wi != pw_idx) { pw = a.present[row * (long)a.present_stride + wi]; pw_idx = wi; } present = (pw >> (f & 31)) & 1u; }
            if (!present && has_rep) { raw = rep; present = true; }
            double val; int st;
            if (m.prep_simple) { st = present ? 0 : 1; val = raw; if (fd.enc != ENC_FLOAT && m.prep[f].dtype == DT_FLOAT) val = (double)(float)raw; }
            else st = prep_value(m.prep[f], m.prep_pool, present, raw, val);
            uint32_t key = 0;
            if (st == 0) key = encode_key(fd, f, val); else if (st == 2) row_fail[rb + lane] = 1;
            keys[f * E + rb + lane] = key;
          }
        }
        __syncthreads();
      }
    }
    if (rows_alias_ring && tid == 0) { fence_proxy_async(); issue(g_lo); if (g_lo + 1 < g_hi) issue(g_lo + 1); }
    // ---------------- phase B: walk all trees
    const uint32_t kbase = smem_u32(keys) + 4u * tid;   // byte address of this event's key column
    const uint32_t E4 = 4u * E;
    double sum = 0.0, wsum = 0.0;
    uint32_t seen = 0;  // classification: number of distinct labels seen so far (insertion order)
    uint32_t cur_ch = 0;   // modelChain: channel of the running sum
    double* chv = reinterpret_cast<double*>(votes);   // modelChain: finished channel values [channel][event]
    if (MODE == 2) for (int l = 0; l < m.n_labels; l++) { votes[l * E + tid] = 0; votes[(m.n_labels + l) * E + tid] = 0xFFFF; }
    for (int g = g_lo; g < g_hi; g++) {
      uint32_t slot = q_wait & 1u;
      mbar_wait(&bars[slot], (q_wait >> 1) & 1u); q_wait++;
      const uint32_t base = smem_u32(ring + (size_t)slot * G * blob_words);
      const int nt = min(G, T - g * G);
      for (int t = 0; t < nt; t += TI) {
        uint32_t idx[TI], bl[TI];
#pragma unroll
        for (int k = 0; k < TI; k++) { idx[k] = 1; bl[k] = base + (uint32_t)min(t + k, nt - 1) * blob_bytes; }
#pragma unroll
        for (int d = 0; d < (D ? D : 16); d++) {
          if (!D && d >= depth) break;
#pragma unroll
          for (int k = 0; k < TI; k++) {
            uint32_t nx, ny, key;
            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(nx), "=r"(ny) : "r"(bl[k] + idx[k] * 8u));
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(key) : "r"(kbase + (nx & 0x7FFFFFFFu) * E4));
            uint32_t vv = key - (nx >> 31);
            // ny holds ~ka: vv + ~ka carries out exactly when vv > ka ("go right"), and the carry is added
            // into idx + idx:  idx = 2*idx + (vv > ka)  in two integer instructions
            asm("{\n\t.reg .u32 t;\n\tadd.cc.u32 t, %2, %3;\n\taddc.u32 %0, %1, %1;\n\t}" : "=r"(idx[k]) : "r"(idx[k]), "r"(ny), "r"(vv));
          }
        }
#pragma unroll
        for (int k = 0; k < TI; k++) {
          if (t + k < nt) {
            if (MODE != 2) {
              double lv, w;
              asm volatile("ld.shared.f64 %0, [%1];" : "=d"(lv) : "r"(bl[k] + idx[k] * 8u));
              if (MODE == 1) { asm volatile("ld.shared.f64 %0, [%1];" : "=d"(w) : "r"(bl[k])); sum += w * lv; wsum += w; }
              else if (MODE == 3) {
                uint32_t ch; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(ch) : "r"(bl[k]));
                if (ch != cur_ch) { chv[cur_ch * E + tid] = channel_value(fd.channels[cur_ch], sum); sum = 0.0; cur_ch = ch; }   // trees arrive channel by channel
                sum += lv;
              } else if (SLICE) { if (row0 + tid < a.n) leaf_out[(long)(g * G + t + k) * a.n + row0 + tid] = lv; }   // [tree][event]: coalesced
              else sum += lv;
            } else {
              uint32_t lab;
              asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lab) : "r"(bl[k] + idx[k] * 8u));
              uint16_t c = votes[lab * E + tid];
              if (c == 0) votes[(m.n_labels + lab) * E + tid] = (uint16_t)seen++;
              votes[lab * E + tid] = c + 1;
            }
            if (DBG) if (a.node_ids && row0 + tid < a.n) a.node_ids[(row0 + tid) * (long)a.node_stride + g * G + t + k] =
                fd.leaf_node_id[((size_t)(g * G + t + k) << depth) + (idx[k] - (1u << depth))];
          }
        }
      }
      __syncwarp();     // this warp is done with the slot; only the producer waits for the others before refilling it
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bars[2 + slot])) : "memory");
      if (tid == 0 && g + 2 < g_hi) { mbar_wait(&bars[2 + slot], ((q_wait - 1) >> 1) & 1u); fence_proxy_async(); issue(g + 2); }
    }
    // ---------------- epilogue: combine, Targets, extraction
    long row = row0 + tid;
    if (SLICE) { if (row < a.n && blockIdx.y == 0) a.status[row] = row_fail[tid] ? ST_PREP : ST_OK; }   // the sum kernel finishes the row
    else if (row < a.n) {
      if (row_fail[tid]) write_out(a, row, 0, ST_PREP, -1);
      else if (MODE == 3) {
        chv[cur_ch * E + tid] = channel_value(fd.channels[cur_ch], sum);
        chain_finish(m, a, row, [&](int c) { return chv[c * E + tid]; });
      } else if (MODE != 2) {
        double v = sum;
        if (fd.method == FM_AVERAGE) v = sum / (double)T; else if (fd.method == FM_WAVERAGE) v = sum / wsum;
        write_out(a, row, apply_target(m.tx, v), m.target_status, -1);
      } else {
        int best = -1; uint32_t bc = 0, bo = 0;
        for (int l = 0; l < m.n_labels; l++) {
          uint32_t c = votes[l * E + tid], o = votes[(m.n_labels + l) * E + tid];
          if (c && (best < 0 || c > bc || (c == bc && o < bo))) { best = l; bc = c; bo = o; }  // first maximal in insertion order
        }
        write_label(m, a, row, best, best);
      }
    }
    __syncthreads();  // keys / votes / row_fail / ring (row stage) are reused by the next tile
  }
}

// Second half of the small-batch latency path: one thread per event adds the T leaf values in segment order (the same
// sequence of double additions as the one-CTA walk, so the bits are the same), then average / Targets / extraction.
__global__ void __launch_bounds__(128) forest_ordered_sum_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const double* __restrict__ leaf) {
  long row = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (row >= a.n) return;
  if (a.status[row] != ST_OK) { write_out(a, row, 0, ST_PREP, -1); return; }
  const int T = m.forest.n_trees; const double* lv = leaf + row;   // [tree][event]
  double sum = 0.0; int t = 0;
  for (; t + 8 <= T; t += 8) {   // eight loads in flight, added strictly in order
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = __ldcs(lv + (long)(t + k) * a.n);
#pragma unroll
    for (int k = 0; k < 8; k++) sum += v[k];
  }
  for (; t < T; t++) sum += __ldcs(lv + (long)t * a.n);
  if (m.forest.method == FM_AVERAGE) sum = sum / (double)T;
  write_out(a, row, apply_target(m.tx, sum), m.target_status, -1);
}

// ================================================================================================ forest, general form
// Thread per event, nodes and features straight from global memory (L1/L2).  Handles every compiled tree shape:
// explicit successors, null predictions, last-prediction leaves; also the fallback when a key tile does not fit on chip.
__global__ void __launch_bounds__(256) forest_general_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a) {
  const ForestDesc& fd = m.forest;
  long row = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (row >= a.n) return;
  const int T = fd.n_trees; const bool classify = fd.classify;
  double sum = 0, wsum = 0; bool fail = row_fails_preparation(m, a, row), null_result = false;
  constexpr int MAXL = 64;
  static_assert(MAX_MEDIAN_SEGMENTS <= MAXL, "median values share the vote array");
  double votes[MAXL]; int order[MAXL]; int seen = 0;   // regression median / max: votes[] holds the active segments' values
  if (classify) for (int l = 0; l < m.n_labels && l < MAXL; l++) { votes[l] = 0; order[l] = -1; }
  int last_node_id = -1, n_active = 0, first_label = -1;
  double chv[MAX_CHANNELS]; int cur_ch = 0;   // modelChain: finished channel values
  for (int t = 0; t < T && !fail; t++) {
    int32_t s = fd.troot[t];
    while (s >= 0) {
      GNode nd = fd.gnodes[s];
      double v; int st = load_field(m, a, row, (int)nd.feat, v);
      if (st == 2) { fail = true; break; }
      if (st == 1) s = nd.on_missing;
      else s = encode_key(fd, (int)nd.feat, v) <= nd.k ? nd.on_true : nd.on_false;
    }
    if (fail) break;
    if (s == LEAF_SKIP) { if (a.node_ids) a.node_ids[row * (long)a.node_stride + t] = -1; continue; }   // Segment predicate not TRUE
    n_active++;
    if (a.node_ids) a.node_ids[row * (long)a.node_stride + t] = s == LEAF_NULL ? 0 : fd.gleaf_node_id[~s];
    if (s == LEAF_NULL) {
      null_result = true;
      if (fd.method == FM_SELECT_FIRST) { if (a.node_ids) for (int u = t + 1; u < T; u++) a.node_ids[row * (long)a.node_stride + u] = -1; break; }
      continue;
    }
    int leaf = ~s; last_node_id = fd.gleaf_node_id[leaf];
    double w = fd.tweight[t];
    if (fd.method == FM_SELECT_FIRST) {   // the first segment whose predicate holds decides alone; later segments are not evaluated
      if (classify) first_label = fd.gleaf_label[leaf]; else sum = fd.gleaf_value[leaf];
      if (a.node_ids) for (int u = t + 1; u < T; u++) a.node_ids[row * (long)a.node_stride + u] = -1;
      break;
    }
    if (!classify) {
      double v = fd.gleaf_value[leaf];
      if (fd.method == FM_CHAIN) { int ch = fd.tree_channel[t]; if (ch != cur_ch) { chv[cur_ch] = channel_value(fd.channels[cur_ch], sum); sum = 0.0; cur_ch = ch; } sum += v; }
      else if (fd.method == FM_WSUM || fd.method == FM_WAVERAGE) { sum += w * v; wsum += w; }
      else if (fd.method == FM_MEDIAN || fd.method == FM_MAX) {   // insertion into the sorted prefix
        int k = n_active - 1; while (k > 0 && votes[k - 1] > v) { votes[k] = votes[k - 1]; k--; } votes[k] = v;
      }
      else sum += v;
    } else {
      if (fd.method == FM_AVG_PROB || fd.method == FM_WAVG_PROB) {   // sum of (weighted) leaf distributions; /sum(w) does not move the winner
        const double ww = fd.method == FM_WAVG_PROB ? w : 1.0;
        for (int i = fd.gdist_off[leaf]; i < fd.gdist_off[leaf + 1]; i++) {
          int lab = fd.gdist_label[i];
          if (order[lab] < 0) order[lab] = seen++;
          votes[lab] += ww * fd.gdist_prob[i];
        }
        wsum += ww;
      } else {
        int lab = fd.gleaf_label[leaf];
        if (order[lab] < 0) order[lab] = seen++;
        votes[lab] += (fd.method == FM_WVOTE) ? w : 1.0; wsum += (fd.method == FM_WVOTE) ? w : 1.0;
      }
    }
  }
  if (fail) { write_out(a, row, 0, ST_PREP, -1); return; }
  const bool alone = fd.method == FM_SINGLE || fd.method == FM_SELECT_FIRST;   // the result is one tree's own
  // a null prediction of the deciding tree (or no active segment at all) is "no target value"; a null prediction inside an
  // aggregation fails the evaluation itself, before target extraction is looked at
  if (null_result) { write_out(a, row, 0, (alone || fd.method == FM_CHAIN) && m.target_status ? m.target_status : ST_EVAL, -1); return; }
  if (n_active == 0) { write_out(a, row, 0, m.target_status ? m.target_status : ST_EVAL, -1); return; }
  if (fd.method == FM_CHAIN) { chv[cur_ch] = channel_value(fd.channels[cur_ch], sum); chain_finish(m, a, row, [&](int c) { return chv[c]; }); return; }
  if (!classify) {
    double v = sum;
    if (fd.method == FM_AVERAGE) v = sum / (double)n_active; else if (fd.method == FM_WAVERAGE) v = sum / wsum;
    else if (fd.method == FM_MAX) v = votes[n_active - 1];
    else if (fd.method == FM_MEDIAN) {   // 50th percentile as commons-math's Percentile (LEGACY estimate) computes it: position (n + 1) / 2
      const int h = n_active / 2;
      if (n_active & 1) v = votes[h]; else { double lo = votes[h - 1], hi = votes[h]; v = lo + 0.5 * (hi - lo); }
    }
    write_out(a, row, apply_target(m.tx, v), m.target_status, alone ? last_node_id : -1);
  } else if (fd.method == FM_SELECT_FIRST) {
    write_label(m, a, row, first_label, last_node_id);
  } else {
    int best = -1;
    for (int l = 0; l < m.n_labels; l++) {
      if (order[l] < 0) continue;
      if (best < 0) { best = l; continue; }
      int c = java_cmp(votes[l], votes[best]);
      if (c > 0 || (c == 0 && order[l] < order[best])) best = l;
    }
    write_label(m, a, row, best, fd.method == FM_SINGLE ? last_node_id : best);
  }
}

// ================================================================================================ clustering (J1)
// CTA = 256 threads = 64 events x 4 cluster groups.  The tile's prepared field values sit transposed in shared
// memory (xs[f][e]); centres are broadcast-loaded through L1.  Distances accumulate in double in ClusteringField order.
constexpr int CL_E = 64, CL_G = 4, CL_XS = CL_E + 1, CL_SS = CL_E + 4;
template <bool FAST>
__global__ void __launch_bounds__(CL_E* CL_G) cluster_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ClusterDesc& cd = m.cluster; const int nf = cd.nf, K = cd.k;
  double* xs = reinterpret_cast<double*>(smem);                    // nf x CL_XS  (padded: conflict-free both ways)
  uint8_t* st_s = smem + (size_t)nf * CL_XS * 8;                   // nf x CL_SS  (0 value, 1 missing, 2 fail)
  double* bd = reinterpret_cast<double*>(st_s + (((size_t)nf * CL_SS + 15) & ~size_t(15)));  // CL_G x CL_E
  int* bc = reinterpret_cast<int*>(bd + CL_G * CL_E);
  const int tid = threadIdx.x, e = tid & (CL_E - 1), g = tid / CL_E;
  // re-scoring mode (a.row_map): work item i is row row_map[i], *n_dev of them (the events the FMA path handed back)
  const long n_items = a.n_dev ? min((long)*a.n_dev, a.n) : a.n;
  auto row_of = [&](long item) { return a.row_map ? (long)a.row_map[item] : item; };
  for (long row0 = blockIdx.x * (long)CL_E; row0 < n_items; row0 += gridDim.x * (long)CL_E) {
  __syncthreads();   // the previous tile's reductions are done with the buffers
  // load + prepare: consecutive threads take consecutive clustering fields of one row
  for (int i = tid; i < nf * CL_E; i += blockDim.x) {
    int r = i / nf, f = i - r * nf; long item = row0 + r;
    double v = 0; int st = 1;
    if (item < n_items) { int slot = cd.slot[f]; if (slot >= 0) st = load_field(m, a, row_of(item), slot, v); }
    xs[f * CL_XS + r] = v; st_s[f * CL_SS + r] = (uint8_t)st;
  }
  __syncthreads();
  const bool live = row0 + e < n_items; const long row = live ? row_of(row0 + e) : 0;
  bool fail = live && !cd.slots_identity && row_fails_preparation(m, a, row); double adj = 1.0;   // active fields that are not clustering fields
  {
    double sum_q = 0, sum_nm = 0;
    for (int f = 0; f < nf; f++) { int st = st_s[f * CL_SS + e]; if (st == 2) fail = true; double q = cd.q[f]; sum_q += q; sum_nm += st == 0 ? q : 0.0; }
    adj = sum_q / sum_nm;
  }
  int best = -1; double bestd = 0;
  for (int c = g; c < K; c += CL_G) {
    const double* ctr = cd.centers + (size_t)c * nf;
    double acc = 0;
    for (int f = 0; f < nf; f++) {
      if (st_s[f * CL_SS + e] != 0) continue;  // missing fields are skipped
      double x = xs[f * CL_XS + e], y = __ldg(ctr + f), term;
      if (FAST) { double z = fabs(x - y); term = z * z; }
      else {
        double z; int cf = cd.compare[f];
        if (cf == CF_ABSDIFF) z = fabs(x - y);
        else if (cf == CF_GAUSSSIM) { double s = cd.sim_scale[f], dd = x - y; z = exp(-log(2.0) * dd * dd / (s * s)); }
        else if (cf == CF_DELTA) z = java_eq(x, y) ? 0.0 : 1.0;
        else z = java_eq(x, y) ? 1.0 : 0.0;
        double p = (cd.measure == CM_EUCLIDEAN || cd.measure == CM_SQEUCLIDEAN) ? z * z
                 : (cd.measure == CM_MINKOWSKI ? (cd.p == 2.0 ? z * z : (cd.p == 1.0 ? z : pow(z, cd.p))) : z);
        term = cd.weight[f] * p;
      }
      if (!FAST && cd.measure == CM_CHEBYCHEV) acc = fmax(acc, term); else acc += term;
    }
    double t = acc * adj, d;
    if (cd.measure == CM_EUCLIDEAN) d = sqrt(t);
    else if (cd.measure == CM_MINKOWSKI) d = cd.p == 1.0 ? t : (cd.p == 2.0 ? sqrt(t) : pow(t, 1.0 / cd.p));
    else d = t;
    if (best < 0 || java_cmp(d, bestd) < 0) { best = c; bestd = d; }  // strictly smaller wins: first cluster kept on ties
  }
  bd[g * CL_E + e] = bestd; bc[g * CL_E + e] = best;
  __syncthreads();
  if (g == 0 && live) {
    for (int gg = 1; gg < CL_G; gg++) {
      int c2 = bc[gg * CL_E + e]; if (c2 < 0) continue; double d2 = bd[gg * CL_E + e];
      int cmp = best < 0 ? -1 : java_cmp(d2, bestd);
      if (cmp < 0 || (cmp == 0 && c2 < best)) { best = c2; bestd = d2; }
    }
    if (fail) write_out(a, row, 0, ST_PREP, -1);
    else write_label(m, a, row, best, best + 1);
    if (a.probs) a.probs[row * (long)a.prob_stride] = bestd;
  }
  }
}

// ---- fast centre-based path (squaredEuclidean / euclidean, absDiff, unit weights): config C1 / C2 ----------
// Register tile: every thread owns ONE event and CK_TC = 8 clusters (8 running sums); a warp = 32 events x 8 clusters.
// CTA = 16 warps = (16 / groups) event blocks x `groups` cluster groups, persistent over tiles.  Centres sit in shared
// memory transposed [field][cluster] so the 8 centres a thread needs for one field are 4 broadcast LDS.128; the
// tile's prepared inputs sit transposed [field][event] (one conflict-free LDS.64 per field).  Per field and thread:
// 1 + 4 shared loads feed 24 double-precision instructions (sub, mul, add x 8), so the FP64 pipe is the limiter.
// Sums run over the fields in ClusteringField order in double, unfused: bit-identical to the per-event interpreter.
constexpr int CK_TC = 8, CK_WARPS = 16;
struct ClusterPlan { int groups, event_blocks, events, chunk, kp, pipelined; uint32_t off_x, off_st, off_best, smem_bytes; };
template <bool HASMISS>
__global__ void __launch_bounds__(CK_WARPS * 32, 1) cluster_fast_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const ClusterPlan pl) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ClusterDesc& cd = m.cluster; const int nf = cd.nf, K = cd.k, KP = pl.kp, E = pl.events, EP = E + 1;
  double* cent = reinterpret_cast<double*>(smem);                       // nf x KP   [f][c]
  double* xs = reinterpret_cast<double*>(smem + pl.off_x);               // chunk x EP
  uint8_t* sts = smem + pl.off_st;                                       // chunk x (E + 4)
  double* bestd = reinterpret_cast<double*>(smem + pl.off_best);         // groups x E
  int* bestc = reinterpret_cast<int*>(bestd + pl.groups * E);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int eb = warp % pl.event_blocks, cg0 = warp / pl.event_blocks;   // event block, first cluster group of this warp
  const int warps_per_eb = CK_WARPS / pl.event_blocks;
  const int e = eb * 32 + lane;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  for (int i = tid; i < nf * KP; i += blockDim.x) { int f = i / KP, c = i - f * KP; cent[i] = c < K ? cd.centers[(size_t)c * nf + f] : 0.0; }
  const long ntiles = (a.n + E - 1) / E;
  if (pl.pipelined) {
    // Plain double rows (optionally with present bits: HASMISS), all fields in one chunk, one cluster group per warp: the
    // next tile's rows (and present words) are copied into the second buffer (cp.async, straight into the transposed
    // position) while this tile's distances are computed.
    const int W = a.present_stride;                                          // <= 4 words in this path (HASMISS)
    uint32_t* pbuf = reinterpret_cast<uint32_t*>(sts);                       // 2 x E x 4 present words
    auto fill_async = [&](long tile, int buf) {
      const long row0 = tile * E; double* xb = xs + (size_t)buf * pl.chunk * EP;
      const int sh = 31 - __clz(nf); const bool pow2 = (nf & (nf - 1)) == 0;
      for (int i = tid; i < E * nf; i += blockDim.x) {
        int r = pow2 ? (i >> sh) : (i / nf), f = i - r * nf; long rr = row0 + r;
        if (rr < a.n) cp_async8(xb + f * EP + r, a.X + rr * (long)m.F + cd.slot[f]); else xb[f * EP + r] = 0.0;
      }
      if (HASMISS) for (int i = tid; i < E * W; i += blockDim.x) {
        int r = i / W, w = i - r * W; long rr = row0 + r; uint32_t* dst = pbuf + (buf * E + r) * 4 + w;
        if (rr < a.n) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(a.present + rr * (long)W + w) : "memory");
        else *dst = 0u;
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int buf = 0;
    if ((long)blockIdx.x < ntiles) fill_async(blockIdx.x, 0);
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, buf ^= 1) {
      const long row = tile * E + e;
      if (tile + gridDim.x < ntiles) { fill_async(tile + gridDim.x, buf ^ 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
      else asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();   // this tile's rows (and, the first time, the centres) are in place
      double acc[CK_TC];
#pragma unroll
      for (int k = 0; k < CK_TC; k++) acc[k] = 0.0;
      const double* xb = xs + (size_t)buf * pl.chunk * EP; const double* cp = cent + cg0 * CK_TC;
      double adj = 1.0;
      if (HASMISS) {   // the fields a row supplies, as bits; a missing field is skipped and the sum rescaled (all weights are 1)
        uint32_t pw[4]; int npresent = 0;
#pragma unroll
        for (int w = 0; w < 4; w++) {
          pw[w] = w < W ? pbuf[(buf * E + e) * 4 + w] : 0u;
          const int lim = nf - 32 * w; if (lim < 32) pw[w] &= lim > 0 ? (1u << lim) - 1u : 0u;
          npresent += __popc(pw[w]);
        }
        adj = (double)nf / (double)npresent;
#pragma unroll
        for (int w = 0; w < 4; w++) {
          uint32_t bits = pw[w]; const int fend = min(nf, 32 * (w + 1));
          for (int f = 32 * w; f < fend; f++, cp += KP, bits >>= 1) {
            const double x = xb[f * EP + e]; const bool present = bits & 1u;
            const double2* c2 = reinterpret_cast<const double2*>(cp);
#pragma unroll
            for (int k = 0; k < CK_TC / 2; k++) {
              double2 z = c2[k];
              double d0 = fabs(x - z.x), d1 = fabs(x - z.y);
              double t0 = d0 * d0, t1 = d1 * d1;
              if (present) { acc[2 * k] += t0; acc[2 * k + 1] += t1; }
            }
          }
        }
      } else {
        for (int f = 0; f < nf; f++, cp += KP) {
          const double x = xb[f * EP + e];
          const double2* c2 = reinterpret_cast<const double2*>(cp);
#pragma unroll
          for (int k = 0; k < CK_TC / 2; k++) {
            double2 z = c2[k];
            double d0 = fabs(x - z.x), d1 = fabs(x - z.y);
            acc[2 * k] += d0 * d0; acc[2 * k + 1] += d1 * d1;
          }
        }
      }
      int bc = -1; double bd = 0;
#pragma unroll
      for (int k = 0; k < CK_TC; k++) {
        int c = cg0 * CK_TC + k;
        if (c < K) { double t = HASMISS ? acc[k] * adj : acc[k], d = cd.measure == CM_EUCLIDEAN ? sqrt(t) : t; if (bc < 0 || java_cmp(d, bd) < 0) { bc = c; bd = d; } }
      }
      bestd[cg0 * E + e] = bc < 0 ? INF : bd; bestc[cg0 * E + e] = bc;
      __syncthreads();
      if (cg0 == 0 && row < a.n) {
        for (int g = 1; g < warps_per_eb; g++) {
          int c2 = bestc[g * E + e]; if (c2 < 0) continue; double d2 = bestd[g * E + e];
          int cmp = bc < 0 ? -1 : java_cmp(d2, bd);
          if (cmp < 0 || (cmp == 0 && c2 < bc)) { bc = c2; bd = d2; }
        }
        write_label(m, a, row, bc, bc + 1);
        if (a.probs) a.probs[row * (long)a.prob_stride] = bd;
      }
      __syncthreads();   // bestd / bestc and this buffer are free for the next rounds
    }
    return;
  }
  for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long row0 = tile * E, row = row0 + e;
    int bc = -1; double bd = 0; bool fail = false;
    const int passes = (pl.groups + warps_per_eb - 1) / warps_per_eb;   // uniform trip count: every warp joins the barriers
    for (int pass = 0; pass < passes; pass++) {
      const int cg = cg0 + pass * warps_per_eb; const bool active = cg < pl.groups;
      double acc[CK_TC];
#pragma unroll
      for (int k = 0; k < CK_TC; k++) acc[k] = 0.0;
      double nmiss_q = 0, all_q = 0;
      for (int f0 = 0; f0 < nf; f0 += pl.chunk) {
        const int nfc = min(pl.chunk, nf - f0);
        __syncthreads();   // previous chunk fully consumed (and the centres written) before xs is overwritten
        const int sh = 31 - __clz(nfc); const bool pow2 = (nfc & (nfc - 1)) == 0;
        for (int i = tid; i < E * nfc; i += blockDim.x) {
          int r = pow2 ? (i >> sh) : (i / nfc), f = i - r * nfc; long rr = row0 + r; double v = 0; int st = 1;
          if (HASMISS) { if (rr < a.n) st = load_field(m, a, rr, cd.slot[f0 + f], v); sts[f * (E + 4) + r] = (uint8_t)st; }
          else if (rr < a.n) v = __ldcs(a.X + rr * (long)m.F + cd.slot[f0 + f]);   // plain doubles, all present
          xs[f * EP + r] = v;
        }
        __syncthreads();
        const double* cp = cent + (size_t)f0 * KP + (active ? cg : 0) * CK_TC;
        if (active) for (int f = 0; f < nfc; f++, cp += KP) {
          double x = xs[f * EP + e];
          bool present = true;
          if (HASMISS) { int st = sts[f * (E + 4) + e]; present = st == 0; fail = fail || st == 2; all_q += 1.0; nmiss_q += present ? 1.0 : 0.0; }
          const double2* c2 = reinterpret_cast<const double2*>(cp);
#pragma unroll
          for (int k = 0; k < CK_TC / 2; k++) {
            double2 z = c2[k];
            double d0 = fabs(x - z.x), d1 = fabs(x - z.y);
            double t0 = d0 * d0, t1 = d1 * d1;
            if (HASMISS) { t0 = present ? t0 : 0.0; t1 = present ? t1 : 0.0; }   // a skipped field adds nothing (sums are >= +0)
            acc[2 * k] += t0; acc[2 * k + 1] += t1;
          }
        }
      }
      const double adj = HASMISS ? all_q / nmiss_q : 1.0;
#pragma unroll
      for (int k = 0; k < CK_TC; k++) {
        int c = cg * CK_TC + k;
        if (active && c < K) {
          double t = acc[k] * adj, d = cd.measure == CM_EUCLIDEAN ? sqrt(t) : t;
          if (bc < 0 || java_cmp(d, bd) < 0) { bc = c; bd = d; }
        }
      }
    }
    bestd[cg0 * E + e] = bc < 0 ? INF : bd; bestc[cg0 * E + e] = bc;
    __syncthreads();
    if (cg0 == 0 && row < a.n) {
      for (int g = 1; g < warps_per_eb; g++) {
        int c2 = bestc[g * E + e]; if (c2 < 0) continue; double d2 = bestd[g * E + e];
        int cmp = bc < 0 ? -1 : java_cmp(d2, bd);
        if (cmp < 0 || (cmp == 0 && c2 < bc)) { bc = c2; bd = d2; }
      }
      if (fail) write_out(a, row, 0, ST_PREP, -1); else write_label(m, a, row, bc, bc + 1);
      if (a.probs) a.probs[row * (long)a.prob_stride] = bd;
    }
  }
}

// ---- the same register tile with TWO events per thread (plain double rows, with or without present bits) --------------
// A warp-wide broadcast LDS.128 of centres costs two shared-memory wavefronts; with one event per thread the 4 centre loads
// + 1 input load of a field (11 wavefronts) nearly fill the 12 cycles its 24 FP64 instructions take.  Two events per thread
// (lane l owns events l and l + 32 of its warp's 64) halve the centre traffic per distance term.
// CTA = 16 warps = event blocks x cluster groups (groups <= 16); a tile of 64 x event_blocks events is processed in chunks
// of fields, and the chunks of this and the next tile are double-buffered with cp.async (8 B each, straight into the
// transposed position).
struct ClusterPairPlan { int groups, event_blocks, events, kp, chunk, nchunks; uint32_t off_x, off_p, off_best, smem_bytes; };
template <bool HASMISS>
__global__ void __launch_bounds__(512, 1) cluster_pair_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const ClusterPairPlan pl) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ClusterDesc& cd = m.cluster; const int nf = cd.nf, K = cd.k, KP = pl.kp, E = pl.events, EP = E + 1;
  double* cent = reinterpret_cast<double*>(smem);                       // nf x KP   [f][c]
  double* xs = reinterpret_cast<double*>(smem + pl.off_x);               // 2 x chunk x EP
  uint32_t* pbuf = reinterpret_cast<uint32_t*>(smem + pl.off_p);         // 2 x E x 4 present words (by tile parity)
  double* bestd = reinterpret_cast<double*>(smem + pl.off_best);         // groups x E
  int* bestc = reinterpret_cast<int*>(bestd + pl.groups * E);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int eb = warp % pl.event_blocks, cg = warp / pl.event_blocks;    // warps = event_blocks x groups exactly
  const int e0 = eb * 64 + lane;                                          // this thread's events: e0 and e0 + 32
  const int W = a.present_stride, NC = pl.nchunks;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  for (int i = tid; i < nf * KP; i += blockDim.x) { int f = i / KP, c = i - f * KP; cent[i] = c < K ? cd.centers[(size_t)c * nf + f] : 0.0; }
  const long ntiles = (a.n + E - 1) / E;
  const long my_tiles = (long)blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long items = my_tiles * NC;                                       // (tile, chunk) pairs of this CTA, in order
  auto fill_async = [&](long item) {
    const long ti = item / NC; const int c = (int)(item - ti * NC), buf = (int)(item & 1);
    const long row0 = (blockIdx.x + ti * gridDim.x) * E; double* xb = xs + (size_t)buf * pl.chunk * EP;
    const int f0 = c * pl.chunk, nfc = min(pl.chunk, nf - f0);
    const int sh = 31 - __clz(nfc); const bool pow2 = (nfc & (nfc - 1)) == 0;
    for (int i = tid; i < E * nfc; i += blockDim.x) {
      int r = pow2 ? (i >> sh) : (i / nfc), f = i - r * nfc; long rr = row0 + r;
      if (rr < a.n) cp_async8(xb + f * EP + r, a.X + rr * (long)m.F + cd.slot[f0 + f]); else xb[f * EP + r] = 0.0;
    }
    if (HASMISS && c == 0) for (int i = tid; i < E * W; i += blockDim.x) {
      int r = i / W, w = i - r * W; long rr = row0 + r; uint32_t* dst = pbuf + ((int)(ti & 1) * E + r) * 4 + w;
      if (rr < a.n) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(a.present + rr * (long)W + w) : "memory");
      else *dst = 0u;
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (items > 0) fill_async(0);
  double acc[2][CK_TC]; double adj[2] = {1.0, 1.0}; uint32_t pw[2][4] = {{0u, 0u, 0u, 0u}, {0u, 0u, 0u, 0u}};
  for (long item = 0; item < items; item++) {
    const long ti = item / NC; const int c = (int)(item - ti * NC), buf = (int)(item & 1);
    const long tile = blockIdx.x + ti * gridDim.x;
    if (item + 1 < items) { fill_async(item + 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();   // this chunk's values (and, the first time, the centres) are in place
    if (c == 0) {
#pragma unroll
      for (int k = 0; k < CK_TC; k++) { acc[0][k] = 0.0; acc[1][k] = 0.0; }
      if (HASMISS) {   // the fields a row supplies, as bits; a missing field is skipped and the sum rescaled (all weights are 1)
#pragma unroll
        for (int v = 0; v < 2; v++) {
          int npresent = 0;
#pragma unroll
          for (int w = 0; w < 4; w++) {
            pw[v][w] = w < W ? pbuf[((int)(ti & 1) * E + e0 + 32 * v) * 4 + w] : 0u;
            const int lim = nf - 32 * w; if (lim < 32) pw[v][w] &= lim > 0 ? (1u << lim) - 1u : 0u;
            npresent += __popc(pw[v][w]);
          }
          adj[v] = (double)nf / (double)npresent;
        }
      }
    }
    const int f0 = c * pl.chunk, f1 = min(nf, f0 + pl.chunk);
    const double* xb = xs + (size_t)buf * pl.chunk * EP + e0; const double* cp = cent + (size_t)f0 * KP + cg * CK_TC;
    if (HASMISS) {
#pragma unroll
      for (int w = 0; w < 4; w++) {
        const int fa = max(f0, 32 * w), fb = min(f1, 32 * (w + 1));   // the part of present word w inside this chunk (chunks are multiples of 32)
        uint32_t b0 = pw[0][w] >> (fa & 31), b1 = pw[1][w] >> (fa & 31);
        for (int f = fa; f < fb; f++, cp += KP, xb += EP, b0 >>= 1, b1 >>= 1) {
          const double x0 = xb[0], x1 = xb[32]; const bool p0 = b0 & 1u, p1 = b1 & 1u;
          const double2* c2 = reinterpret_cast<const double2*>(cp);
#pragma unroll
          for (int k = 0; k < CK_TC / 2; k++) {
            double2 z = c2[k];
            double d00 = fabs(x0 - z.x), d01 = fabs(x0 - z.y), d10 = fabs(x1 - z.x), d11 = fabs(x1 - z.y);
            double t00 = d00 * d00, t01 = d01 * d01, t10 = d10 * d10, t11 = d11 * d11;
            if (p0) { acc[0][2 * k] += t00; acc[0][2 * k + 1] += t01; }
            if (p1) { acc[1][2 * k] += t10; acc[1][2 * k + 1] += t11; }
          }
        }
      }
    } else {
      for (int f = f0; f < f1; f++, cp += KP, xb += EP) {
        const double x0 = xb[0], x1 = xb[32];
        const double2* c2 = reinterpret_cast<const double2*>(cp);
#pragma unroll
        for (int k = 0; k < CK_TC / 2; k++) {
          double2 z = c2[k];
          double d00 = fabs(x0 - z.x), d01 = fabs(x0 - z.y), d10 = fabs(x1 - z.x), d11 = fabs(x1 - z.y);
          acc[0][2 * k] += d00 * d00; acc[0][2 * k + 1] += d01 * d01;
          acc[1][2 * k] += d10 * d10; acc[1][2 * k + 1] += d11 * d11;
        }
      }
    }
    if (c == NC - 1) {
#pragma unroll
      for (int v = 0; v < 2; v++) {
        int bc = -1; double bd = 0;
#pragma unroll
        for (int k = 0; k < CK_TC; k++) {
          int cc = cg * CK_TC + k;
          if (cc < K) { double t = HASMISS ? acc[v][k] * adj[v] : acc[v][k], d = cd.measure == CM_EUCLIDEAN ? sqrt(t) : t; if (bc < 0 || java_cmp(d, bd) < 0) { bc = cc; bd = d; } }
        }
        bestd[cg * E + e0 + 32 * v] = bc < 0 ? INF : bd; bestc[cg * E + e0 + 32 * v] = bc;
      }
      __syncthreads();
      if (cg == 0) {
#pragma unroll
        for (int v = 0; v < 2; v++) {
          const int e = e0 + 32 * v; const long row = tile * E + e;
          if (row >= a.n) continue;
          int bc = bestc[e]; double bd = bestd[e];
          for (int g = 1; g < pl.groups; g++) {
            int c2 = bestc[g * E + e]; if (c2 < 0) continue; double d2 = bestd[g * E + e];
            int cmp = bc < 0 ? -1 : java_cmp(d2, bd);
            if (cmp < 0 || (cmp == 0 && c2 < bc)) { bc = c2; bd = d2; }
          }
          write_label(m, a, row, bc, bc + 1);
          if (a.probs) a.probs[row * (long)a.prob_stride] = bd;
        }
      }
    }
    __syncthreads();   // this buffer (and bestd / bestc) are free for the next rounds
  }
}

// ================================================================================================ regression (J4)
// Generic: thread per event, predictor lists in document order (CSR), values prepared on the fly.
constexpr int REG_MAXC = 64;
__global__ void __launch_bounds__(128) regression_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a) {
  const RegressionDesc& rd = m.reg;
  long row = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (row >= a.n) return;
  double y[REG_MAXC]; bool fail = row_fails_preparation(m, a, row), null_result = false;
  for (int t = 0; t < rd.n_tables && !fail; t++) {
    double acc = rd.intercept[t];
    for (int i = rd.term_off[t]; i < rd.term_off[t + 1]; i++) {
      RegTerm tm = rd.terms[i]; double v; int st = tm.slot < 0 ? 1 : load_field(m, a, row, tm.slot, v);
      if (st == 2) { fail = true; break; }
      if (st == 1) { null_result = true; continue; }  // missing numeric predictor: the table has no result
      acc += tm.coef * ipow(v, tm.exponent);
    }
    for (int i = rd.cat_off[t]; i < rd.cat_off[t + 1] && !fail; i++) {
      RegCat c = rd.cats[i]; double v; int st = c.slot < 0 ? 1 : load_field(m, a, row, c.slot, v);
      if (st == 2) { fail = true; break; }
      if (st == 0 && java_eq(v, c.value)) acc += c.coef;
    }
    y[t] = acc;
  }
  if (fail) { write_out(a, row, 0, ST_PREP, -1); return; }
  if (null_result) { write_out(a, row, 0, m.target_status ? m.target_status : ST_EVAL, -1); return; }
  if (!rd.classify) { write_out(a, row, apply_target(m.tx, link_fn(rd.norm, y[0])), m.target_status, -1); return; }
  const int C = rd.n_tables;
  normalize_classes(rd.norm, C, [&](int i) { return y[i]; }, [&](int i, double v) { y[i] = v; });
  int best = 0; for (int i = 1; i < C; i++) if (java_cmp(y[i], y[best]) > 0) best = i;
  write_label(m, a, row, best, best);
  if (a.probs) for (int i = 0; i < C && i < a.prob_stride; i++) a.probs[row * (long)a.prob_stride + i] = y[i];
}

// ---- dense tables, register-tiled (multinomial logistic, config C3) -------------------------------------------
// Same tiling as cluster_fast_kernel: a thread owns one event and 8 tables; a warp = 32 events x 8 tables; the CTA's
// warps = event blocks x table groups.  Inputs and the transposed coefficient block [field][table] stream through
// shared memory in chunks of fields; y[c] accumulates in double, unfused, in field order (= NumericPredictor order).
struct DensePlan { int groups, event_blocks, events, chunk, kp, warps; uint32_t off_x, off_st, off_y, off_flags, smem_bytes; };
// What the tile kernel multiplies: y_t = icpt_t + sum_f wt[f][t] x_f for C tables.  RegressionModel: the RegressionTables.
// ClusteringModel (FAST = 2): wt = -2 c_t, icpt = |c_t|^2, so that y_t = |x - c_t|^2 - |x|^2 and the nearest centre is the smallest y.
struct DenseView {
  const double* wt; const double* icpt; int C;
  const double* csq;      // FAST = 3: ClusterDesc::fast_csq
  double amax, bmax;      // max_t sum_f |wt[f][t]|, max_t |icpt_t|   (FAST = 1)  /  max |c|, unused (FAST = 2)
  int32_t* fb_count; int32_t* fb_rows;   // events the margin test hands to the FAST = 0 kernel
};
// EPT = events per thread (lane l owns events l, l + 32 of its warp's 32 * EPT): a warp-wide broadcast LDS.128 of coefficients
// costs two shared-memory wavefronts, so with one event per thread the coefficient loads alone (11 wavefronts per field for 10
// tables) outlast the 10 cycles the field's 20 FP64 instructions take; two events share them.
// FAST = 0: the reference's arithmetic -- unfused multiply and add in predictor order (JPMML on the JVM).
// FAST = 1 / 2: fused multiply-add (twice the FP64 rate) for softmax classification / nearest centre.  y then differs from the
//   reference's by at most eps = (2 F + 8) 2^-53 (|icpt| + sum |w| |x|), bounded per event from the largest |x|; a label is written
//   only when the winner's margin exceeds 2 eps, every other event goes on the list the FAST = 0 kernel re-scores (a.row_map).
// FAST = 3: nearest centre over rows with missing values (present bits, at most 128 fields): a missing field counts as x = 0 in the
//   dot products and its c^2 (= (wt / 2)^2) is taken back out of |c|^2 afterwards, field by field from the event's missing bits;
//   JPMML's adjustment factor sum(q) / sum(present q) scales all of an event's distances alike and does not move the winner.
template <bool HASMISS, int TC, int EPT, int FAST>
__global__ void __launch_bounds__(256, EPT == 1 ? 3 : 2) regression_tile_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const DensePlan pl, const DenseView dv) {
  extern __shared__ __align__(128) uint8_t smem[];
  const RegressionDesc& rd = m.reg; const int F = m.F, C = dv.C, KP = pl.kp, E = pl.events, EP = E + 1;
  double* wt = reinterpret_cast<double*>(smem);                          // chunk x KP  [f][c]
  double* xs = reinterpret_cast<double*>(smem + pl.off_x);               // chunk x EP
  uint8_t* sts = smem + pl.off_st;                                       // chunk x (E + 4)
  double* ys = reinterpret_cast<double*>(smem + pl.off_y);               // KP x EP
  int* eflags = reinterpret_cast<int*>(smem + pl.off_flags);             // E: bit0 preparation failure, bit1 missing predictor
  uint32_t* pws = reinterpret_cast<uint32_t*>(smem + pl.off_flags) + E;  // FAST = 3: E x 4 present words
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int eb = warp % pl.event_blocks, cg = warp / pl.event_blocks, e = eb * 32 * EPT + lane;
  for (int i = tid; i < E; i += blockDim.x) eflags[i] = 0;
  // re-scoring mode (a.row_map): work item i is row row_map[i], *n_dev of them
  const long n_items = a.n_dev ? min((long)*a.n_dev, a.n) : a.n;
  auto row_of = [&](long item) { return a.row_map ? (long)a.row_map[item] : item; };
  const long ntiles = (n_items + E - 1) / E;
  for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long row0 = tile * E;
    double acc[EPT][TC]; int flags[EPT];  // bit0 preparation failure, bit1 missing predictor
    int xm[EPT];                           // FAST: high word of the largest |x| of the event
#pragma unroll
    for (int v = 0; v < EPT; v++) {
      flags[v] = 0; xm[v] = 0;
#pragma unroll
      for (int k = 0; k < TC; k++) { int c = cg * TC + k; acc[v][k] = c < C ? dv.icpt[c] : 0.0; }
    }
    // double-buffered chunks: the copies of chunk c+1 (cp.async, 8 B each, written straight into the transposed
    // position) are in flight while chunk c is multiplied
    const int nchunks = (F + pl.chunk - 1) / pl.chunk;
    auto fill = [&](int buf, int f0) {
      const int nfc = min(pl.chunk, F - f0);
      double* wb = wt + (size_t)buf * pl.chunk * KP; double* xb = xs + (size_t)buf * pl.chunk * EP; uint8_t* sb = sts + (size_t)buf * pl.chunk * (E + 4);
      for (int i = tid; i < nfc * KP; i += blockDim.x) cp_async8(wb + i, dv.wt + (size_t)f0 * KP + i);
      const int sh = 31 - __clz(nfc); const bool pow2 = (nfc & (nfc - 1)) == 0;
      if (!HASMISS && pow2 && nfc <= (int)blockDim.x && (blockDim.x & (nfc - 1)) == 0) {
        // plain doubles, power-of-two chunk: a thread keeps its field and walks down the rows with pointer increments
        const int f = tid & (nfc - 1), rstep = blockDim.x >> sh, live = (int)min((long)E, n_items - row0);
        int r = tid >> sh;
        const double* src = a.X + (row0 + r) * (long)F + f0 + f; double* dst = xb + f * EP + r; const long sstep = (long)rstep * F;
        for (; r < live; r += rstep, src += sstep, dst += rstep) cp_async8(dst, src);
        for (; r < E; r += rstep, dst += rstep) *dst = 0.0;
      } else
      for (int i = tid; i < E * nfc; i += blockDim.x) {
        int r = pow2 ? (i >> sh) : (i / nfc), f = i - r * nfc; long it = row0 + r;
        if (HASMISS) { double v = 0; int st = 0; if (it < n_items) st = load_field(m, a, row_of(it), f0 + f, v); sb[f * (E + 4) + r] = (uint8_t)st; xb[f * EP + r] = v; }
        else if (it < n_items) cp_async8(xb + f * EP + r, a.X + it * (long)F + f0 + f);   // plain doubles, all present
        else xb[f * EP + r] = 0.0;
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    __syncthreads();   // the previous tile's epilogue is done with ys / the buffers
    if (FAST == 3) {
      const int W = (F + 31) >> 5;
      for (int i = tid; i < E * 4; i += blockDim.x) {
        const int r = i >> 2, w = i & 3; const long it = row0 + r; uint32_t bits = 0;
        if (it < n_items && w < W) { bits = a.present[it * (long)a.present_stride + w]; if (w == W - 1 && (F & 31)) bits &= (1u << (F & 31)) - 1u; }
        pws[i] = bits;
      }
      __syncthreads();
    }
    fill(0, 0);
    if (FAST == 3) {   // |c|^2 (the accumulators' start) is summed over every field: take the missing ones back out, two fields per step, while
                       // the first chunk's copies are in flight
      const double* csq = dv.csq + cg * TC;
#pragma unroll
      for (int v = 0; v < EPT; v++) {
        for (int w = 0; w < ((F + 31) >> 5); w++) {
          uint32_t miss = ~pws[(e + 32 * v) * 4 + w]; if (w == ((F + 31) >> 5) - 1 && (F & 31)) miss &= (1u << (F & 31)) - 1u;
          while (miss) {
            const int fa = 32 * w + __ffs(miss) - 1; miss &= miss - 1u;
            const int fb = miss ? 32 * w + __ffs(miss) - 1 : F; miss &= miss - 1u;     // row F is all zero; 0 & 0xFFFFFFFF stays 0
            const double2* qa = reinterpret_cast<const double2*>(csq + (size_t)fa * KP); const double2* qb = reinterpret_cast<const double2*>(csq + (size_t)fb * KP);
            double2 va[TC / 2], vb[TC / 2];
#pragma unroll
            for (int k = 0; k < TC / 2; k++) { va[k] = __ldg(qa + k); vb[k] = __ldg(qb + k); }
#pragma unroll
            for (int k = 0; k < TC / 2; k++) { acc[v][2 * k] -= va[k].x; acc[v][2 * k + 1] -= va[k].y; acc[v][2 * k] -= vb[k].x; acc[v][2 * k + 1] -= vb[k].y; }
          }
        }
      }
    }

    for (int c = 0; c < nchunks; c++) {
      const int f0 = c * pl.chunk, nfc = min(pl.chunk, F - f0), buf = c & 1;
      if (c + 1 < nchunks) { fill(buf ^ 1, f0 + pl.chunk); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
      else asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();
      const double* wp = wt + (size_t)buf * pl.chunk * KP + cg * TC; const double* xb = xs + (size_t)buf * pl.chunk * EP;
      const uint8_t* sb = sts + (size_t)buf * pl.chunk * (E + 4);
      if (FAST == 3) {   // whatever sits in a missing slot becomes 0 before the chunk is multiplied (chunks are powers of two: whole present words, or an aligned part of one)
        double* xw = xs + (size_t)buf * pl.chunk * EP; const int nw = (nfc + 31) >> 5;
        for (int i = tid; i < E * nw; i += blockDim.x) {
          const int r = i % E, wq = i / E;
          uint32_t miss = ~pws[r * 4 + (f0 >> 5) + wq] >> (f0 & 31); if (nfc - 32 * wq < 32) miss &= (1u << (nfc - 32 * wq)) - 1u;
          while (miss) { const int bq = __ffs(miss) - 1; miss &= miss - 1u; xw[(32 * wq + bq) * EP + r] = 0.0; }
        }
        __syncthreads();
      }
      for (int f = 0; f < nfc; f++, wp += KP) {
        double x[EPT];
#pragma unroll
        for (int v = 0; v < EPT; v++) {
          x[v] = xb[f * EP + e + 32 * v];
          if (HASMISS) { int st = sb[f * (E + 4) + e + 32 * v]; flags[v] |= st == 2 ? 1 : (st == 1 ? 2 : 0); }
          if (FAST) xm[v] = max(xm[v], __double2hiint(x[v]) & 0x7fffffff);
        }
        const double2* w2 = reinterpret_cast<const double2*>(wp);
#pragma unroll
        for (int k = 0; k < TC / 2; k++) {
          double2 w = w2[k];
#pragma unroll
          for (int v = 0; v < EPT; v++) {
            if (FAST != 0) { acc[v][2 * k] = fma(w.x, x[v], acc[v][2 * k]); acc[v][2 * k + 1] = fma(w.y, x[v], acc[v][2 * k + 1]); }
            else { acc[v][2 * k] += w.x * x[v]; acc[v][2 * k + 1] += w.y * x[v]; }
          }
        }
      }
      __syncthreads();   // buffer `buf` may be refilled by the next iteration's fill (and, after the last chunk, reused for ys)
    }
#pragma unroll
    for (int v = 0; v < EPT; v++) {
#pragma unroll
      for (int k = 0; k < TC; k++) ys[(cg * TC + k) * EP + e + 32 * v] = acc[v][k];
      if (HASMISS) { if (flags[v]) atomicOr(eflags + e + 32 * v, flags[v]); }
    }
    __syncthreads();
#pragma unroll
    for (int v = 0; v < EPT; v++) {
      const int e = eb * 32 * EPT + lane + 32 * v; const long item = row0 + e;
      bool rescore = false; long row = 0;
      if (cg == 0 && item < n_items) {
        row = row_of(item);
        int fl = HASMISS ? eflags[e] : 0;
        if (fl & 1) write_out(a, row, 0, ST_PREP, -1);
        else if (fl & 2) write_out(a, row, 0, m.target_status ? m.target_status : ST_EVAL, -1);
        else if (FAST == 2 || FAST == 3) {
          // nearest centre: smallest y, the first one on ties -- decided here only when the runner-up is farther than the bound
          const double xmax = __hiloint2double(xm[v] + 1, 0), s = xmax + dv.amax;
          const double eps = (double)(2 * F + 8) * 2.220446049250313e-16 * (double)F * s * s;
          int best = 0; double yb = ys[e], y2 = __longlong_as_double(0x7FF0000000000000LL);
          for (int i = 1; i < C; i++) { const double y = ys[i * EP + e]; if (y < yb) { y2 = yb; yb = y; best = i; } else y2 = fmin(y2, y); }
          bool any_present = true;
          if (FAST == 3) { const uint32_t* pw = pws + e * 4; any_present = (pw[0] | pw[1] | pw[2] | pw[3]) != 0u; }   // no field at all: the exact kernel's business
          if (any_present && (C == 1 || (y2 - yb > 2.0 * eps && s < 1e150))) write_label(m, a, row, best, best + 1); else rescore = true;
        }
        else if (!rd.classify) write_out(a, row, apply_target(m.tx, link_fn(rd.norm, ys[e])), m.target_status, -1);
        else {
          bool decided = true;
          if (FAST == 1) {
            const double xmax = __hiloint2double(xm[v] + 1, 0);
            const double eps = (double)(2 * F + 8) * 1.1102230246251565e-16 * (dv.bmax + dv.amax * xmax);
            double yb = ys[e], y2 = -__longlong_as_double(0x7FF0000000000000LL), ya = fabs(yb);
            for (int i = 1; i < C; i++) { const double y = ys[i * EP + e]; ya = fmax(ya, fabs(y)); if (y > yb) { y2 = yb; yb = y; } else y2 = fmax(y2, y); }
            decided = C == 1 || (yb - y2 > 2.0 * eps && ya < 600.0);      // |y| < 600: exp() stays finite and non-zero, so the order of p is the order of y
          }
          if (!decided) rescore = true;
          else {
            normalize_classes(rd.norm, C, [&](int i) { return ys[i * EP + e]; }, [&](int i, double v) { ys[i * EP + e] = v; });
            int best = 0; double bp = ys[e];
            for (int i = 1; i < C; i++) { double p = ys[i * EP + e]; if (java_cmp(p, bp) > 0) { best = i; bp = p; } }
            write_label(m, a, row, best, best);
            if (a.probs) for (int i = 0; i < C && i < a.prob_stride; i++) a.probs[row * (long)a.prob_stride + i] = ys[i * EP + e];
          }
        }
      }
      if (FAST) {   // (cg == 0 warps only have candidates; the ballot is over the whole warp)
        const unsigned need = __ballot_sync(0xffffffffu, rescore), alive = __ballot_sync(0xffffffffu, cg == 0 && item < n_items);
        if (a.stats && lane == 0 && alive) { atomicAdd(a.stats + 2, (unsigned long long)__popc(alive)); if (need) atomicAdd(a.stats + 3, (unsigned long long)__popc(need)); }
        if (need) {
          int base = 0;
          if (lane == 0) base = atomicAdd(dv.fb_count, __popc(need));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (rescore) dv.fb_rows[base + __popc(need & ((1u << lane) - 1u))] = (int32_t)row;
        }
      }
    }
    if (HASMISS) { __syncthreads(); for (int i = tid; i < E; i += blockDim.x) eflags[i] = 0; }
  }
}

// ================================================================================================ general regression (J5)
// Thread per event: design vector from the PPMatrix (factor indicators, covariate powers, products), one dot product
// per ParamMatrix row in PCell order, link / multinomial softmax, first maximal category.
__device__ __forceinline__ double inv_link(int link, double p, double eta) {
  switch (link) {
    case GL_IDENTITY: return eta;
    case GL_LOG: return exp(eta);
    case GL_LOGIT: return 1.0 / (1.0 + exp(-eta));
    case GL_PROBIT: return 0.5 * erfc(-eta / sqrt(2.0));
    case GL_CLOGLOG: return 1.0 - exp(-exp(eta));
    case GL_LOGLOG: return exp(-exp(-eta));
    case GL_LOGC: return 1.0 - exp(eta);
    case GL_CAUCHIT: return 0.5 + atan(eta) / 3.14159265358979323846;
    case GL_POWER: return p == 0.0 ? exp(eta) : pow(eta, 1.0 / p);
    default: return 1.0 / (p * (exp(-eta) - 1.0));  // negbin
  }
}
__global__ void __launch_bounds__(128) general_regression_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a) {
  const GeneralRegDesc& gd = m.greg;
  long row = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (row >= a.n) return;
  double x[GR_MAX_PARAMS]; bool fail = row_fails_preparation(m, a, row), null_result = false;
  for (int p = 0; p < gd.n_params; p++) {
    double v = 1.0;
    for (int i = gd.cell_off[p]; i < gd.cell_off[p + 1]; i++) {
      GCell c = gd.cells[i]; double f; int st = c.slot < 0 ? 1 : load_field(m, a, row, c.slot, f);
      if (st == 2) fail = true; else if (st == 1) null_result = true;
      else v *= c.is_factor ? (java_eq(f, c.value) ? 1.0 : 0.0) : (c.value == 1.0 ? f : pow(f, c.value));
    }
    x[p] = v;
  }
  if (fail) { write_out(a, row, 0, ST_PREP, -1); return; }
  if (null_result) { write_out(a, row, 0, m.target_status ? m.target_status : ST_EVAL, -1); return; }
  auto dot = [&](int r) { double s = 0; for (int i = gd.beta_off[r]; i < gd.beta_off[r + 1]; i++) s += gd.beta[i] * x[gd.beta_param[i]]; return s; };
  if (!gd.classify) {
    double eta = dot(0) + gd.offset;
    double v = gd.model_type == GR_GENERAL_LINEAR ? inv_link(gd.link, gd.link_param, eta) : eta;
    write_out(a, row, apply_target(m.tx, v), m.target_status, -1);
    return;
  }
  double p[GR_MAX_CATS]; const int C = gd.n_cats;
  if (gd.model_type == GR_MULTINOMIAL) {
    double s = 0;
    for (int c = 0; c < C; c++) { double eta = 0; for (int r = 0; r < gd.n_rows; r++) if (gd.row_cat[r] == c) eta = dot(r); p[c] = exp(eta); s += p[c]; }
    for (int c = 0; c < C; c++) p[c] /= s;
  } else {
    double q = inv_link(gd.link, gd.link_param, dot(0) + gd.offset);
    for (int c = 0; c < C; c++) p[c] = c == gd.row_cat[0] ? q : (C == 2 ? 1.0 - q : 0.0);
  }
  int best = 0; for (int c = 1; c < C; c++) if (java_cmp(p[c], p[best]) > 0) best = c;
  write_label(m, a, row, best, best);
  if (a.probs) for (int c = 0; c < C && c < a.prob_stride; c++) a.probs[row * (long)a.prob_stride + c] = p[c];
}

// ================================================================================================ neural network (J6)
// CTA = 256 threads (8 warps) x NN_E = 32 events (lane <-> event).  Activations ping-pong between two transposed
// shared-memory buffers act[neuron][event]; a warp produces 4 neurons at a time, weights are warp-broadcast loads.
// Accumulation runs in connection order in double, as JPMML's NeuralNetworkEvaluator does.
constexpr int NN_E = 32, NN_WARPS = 8, NN_J = 4;
__global__ void __launch_bounds__(NN_E* NN_WARPS) neural_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const NeuralDesc& nd = m.nn; const int W = nd.max_width;
  double* act0 = reinterpret_cast<double*>(smem); double* act1 = act0 + (size_t)W * NN_E;
  __shared__ int flags[NN_E];  // bit0: preparation failure, bit1: missing input (no result)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long row = blockIdx.x * (long)NN_E + lane;
  if (warp == 0) flags[lane] = 0;
  __syncthreads();
  for (int i = warp; i < nd.n_inputs; i += NN_WARPS) {
    NeuralInput in = nd.inputs[i]; double out = 0; int fl = 0;
    if (row < a.n) {
      if (in.kind == NI_CONSTANT) out = in.value;
      else {
        double v; int st = in.slot < 0 ? 1 : load_field(m, a, row, in.slot, v);
        if (st == 2) fl = 1;
        else if (st == 1) { if (in.has_map_missing) out = in.map_missing; else fl = 2; }
        else if (in.kind == NI_FIELDREF) out = v;
        else if (in.kind == NI_NORMDISC) out = java_eq(v, in.value) ? 1.0 : 0.0;
        else { bool miss = false; out = norm_cont(nd.pts + 2 * in.pts_off, in.n_pts, in.outliers, v, miss); if (miss) fl = 2; }
      }
    }
    act0[i * NN_E + lane] = out;
    if (fl) atomicOr(&flags[lane], fl);
  }
  __syncthreads();
  double* cur = act0; double* nxt = act1;
  for (int l = 0; l < nd.n_layers; l++) {
    const NeuralLayer L = nd.layers[l];
    const double* Wm = nd.weights + L.w_off; const double* B = nd.biases + L.b_off;
    if (L.act == ACT_RADIAL) {   // radial basis layer: squared distance to the weight vector instead of a weighted sum; no bias
      for (int j = warp; j < L.n_out; j += NN_WARPS) {
        double z = 0;
        for (int i = 0; i < L.n_in; i++) { const double dlt = cur[i * NN_E + lane] - __ldg(Wm + (size_t)j * L.n_in + i); z += dlt * dlt; }
        nxt[j * NN_E + lane] = exp(B[2 * L.n_out + j] - z / B[L.n_out + j]);
      }
    } else
    for (int j0 = warp * NN_J; j0 < L.n_out; j0 += NN_WARPS * NN_J) {
      double z[NN_J];
#pragma unroll
      for (int k = 0; k < NN_J; k++) z[k] = B[min(j0 + k, L.n_out - 1)];
      for (int i = 0; i < L.n_in; i++) {
        double x = cur[i * NN_E + lane];
#pragma unroll
        for (int k = 0; k < NN_J; k++) z[k] += __ldg(Wm + (size_t)min(j0 + k, L.n_out - 1) * L.n_in + i) * x;
      }
#pragma unroll
      for (int k = 0; k < NN_J; k++) if (j0 + k < L.n_out) nxt[(j0 + k) * NN_E + lane] = activate(L.act, z[k], L.threshold);
    }
    __syncthreads();
    if (L.norm != NORM_NONE && warp == 0) {
      if (L.norm == NORM_SOFTMAX) {
        double s = 0; for (int j = 0; j < L.n_out; j++) { double e = exp(nxt[j * NN_E + lane]); nxt[j * NN_E + lane] = e; s += e; }
        for (int j = 0; j < L.n_out; j++) nxt[j * NN_E + lane] /= s;
      } else {
        double s = 0; for (int j = 0; j < L.n_out; j++) s += nxt[j * NN_E + lane];
        for (int j = 0; j < L.n_out; j++) nxt[j * NN_E + lane] /= s;
      }
    }
    __syncthreads();
    double* t = cur; cur = nxt; nxt = t;
  }
  if (warp != 0 || row >= a.n) return;
  int fl = flags[lane];
  if (row_fails_preparation(m, a, row)) fl |= 1;
  if (fl & 1) { write_out(a, row, 0, ST_PREP, -1); return; }
  if (fl & 2) { write_out(a, row, 0, m.target_status ? m.target_status : ST_EVAL, -1); return; }
  if (!nd.classify) {
    double y = cur[nd.out_neuron[0] * NN_E + lane];
    if (nd.out_denorm_n) {
      const double* pts = nd.pts + 2 * nd.out_denorm_off; int n = nd.out_denorm_n, i = 0;
      while (i + 2 < n && y > pts[2 * (i + 1) + 1]) i++;
      double a0 = pts[2 * i], b0 = pts[2 * i + 1], a1 = pts[2 * i + 2], b1 = pts[2 * i + 3];
      y = (y - b0) / (b1 - b0) * (a1 - a0) + a0;
    }
    write_out(a, row, apply_target(m.tx, y), m.target_status, -1);
    return;
  }
  int best = -1; double bp = 0;
  for (int o = 0; o < nd.n_outputs; o++) {
    double p = cur[nd.out_neuron[o] * NN_E + lane];
    if (a.probs && o < a.prob_stride) a.probs[row * (long)a.prob_stride + o] = p;
    if (best < 0 || java_cmp(p, bp) > 0) { best = o; bp = p; }  // first maximal NeuralOutput
  }
  write_label(m, a, row, best, best);
}

// ---- layer-wise register-tiled MLP with streamed weights (config C5-MLP) -------------------------------------------
// CTA = 8 warps x 32 events (lane <-> event).  Activations ping-pong in shared memory, transposed [neuron][event].
// The transposed weight matrix of every layer streams L2 -> shared memory in row chunks through a 2-slot ring of
// bulk async copies (cp.async.bulk + mbarrier), the same producer scheme as the forest kernel.  A warp owns up to
// 4 groups of 8 output neurons (32 running sums in registers); per input neuron it does one conflict-free LDS.64 of the
// activation and 4 broadcast LDS.128 per group, feeding 16 double-precision instructions per group.
// z accumulates from the bias in connection order, unfused, in double (JPMML NeuralNetworkEvaluator order).
struct NeuralPlan { uint32_t off_act1, off_ring, off_bars, off_flags, smem_bytes; int width; };
// rows of one weight chunk times one activation column: NG groups of 8 running sums per thread
constexpr int NT_E = 32, NT_WARPS = 16, NT_G = 2, NT_RING = 32 * 1024;
template <int NG>
__device__ __forceinline__ void nt_rows(double (&z)[NT_G][8], const double* __restrict__ xcol, const double* __restrict__ wbase, int rows, int n_out_pad, int warp) {
#pragma unroll 4
  for (int r = 0; r < rows; r++) {
    const double x = xcol[r * 32];
    const double2* w2 = reinterpret_cast<const double2*>(wbase + (size_t)r * n_out_pad);
#pragma unroll
    for (int g = 0; g < NG; g++) {
      const int jg = warp + g * NT_WARPS;
#pragma unroll
      for (int k = 0; k < 4; k++) { double2 w = w2[jg * 4 + k]; z[g][2 * k] += w.x * x; z[g][2 * k + 1] += w.y * x; }
    }
  }
}
__global__ void __launch_bounds__(NT_E* NT_WARPS, 1) neural_tile_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const NeuralPlan pl) {
  extern __shared__ __align__(128) uint8_t smem[];
  const NeuralDesc& nd = m.nn;
  double* act0 = reinterpret_cast<double*>(smem); double* act1 = reinterpret_cast<double*>(smem + pl.off_act1);
  double* ring = reinterpret_cast<double*>(smem + pl.off_ring);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + pl.off_bars);
  int* flags = reinterpret_cast<int*>(smem + pl.off_flags);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  uint32_t q_issue = 0, q_wait = 0; int il = 0, ic = 0;   // producer cursor (thread 0): layer, first row of the next chunk
  auto rows_per_chunk = [&](const NeuralLayer& L) { int r = NT_RING / (L.n_out_pad * 8); return r < 1 ? 1 : (r > L.n_in ? L.n_in : r); };
  auto issue_next = [&]() {   // thread 0: next chunk of the per-tile sequence into ring slot q_issue & 1; false at the end
    if (il >= nd.n_layers) return false;
    const NeuralLayer& L = nd.layers[il]; int rpc = rows_per_chunk(L); int rows = min(rpc, L.n_in - ic);
    uint32_t bytes = (uint32_t)rows * L.n_out_pad * 8u, slot = q_issue & 1u;
    mbar_expect_tx(&bars[slot], bytes);
    bulk_g2s(ring + (size_t)slot * (NT_RING / 8), nd.weights_t + L.wt_off + (size_t)ic * L.n_out_pad, bytes, &bars[slot]);
    q_issue++; ic += rows; if (ic >= L.n_in) { ic = 0; il++; }
    return true;
  };
  const long ntiles = (a.n + NT_E - 1) / NT_E;
  for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long row = tile * NT_E + lane;
    if (tid == 0) { il = 0; ic = 0; fence_proxy_async(); issue_next(); issue_next(); }
    if (warp == 0) flags[lane] = 0;
    __syncthreads();
    for (int i = warp; i < nd.n_inputs; i += NT_WARPS) {
      NeuralInput in = nd.inputs[i]; double out = 0; int fl = 0;
      if (row < a.n) {
        if (in.kind == NI_CONSTANT) out = in.value;
        else {
          double v; int st = in.slot < 0 ? 1 : load_field(m, a, row, in.slot, v);
          if (st == 2) fl = 1;
          else if (st == 1) { if (in.has_map_missing) out = in.map_missing; else fl = 2; }
          else if (in.kind == NI_FIELDREF) out = v;
          else if (in.kind == NI_NORMDISC) out = java_eq(v, in.value) ? 1.0 : 0.0;
          else { bool miss = false; out = norm_cont(nd.pts + 2 * in.pts_off, in.n_pts, in.outliers, v, miss); if (miss) fl = 2; }
        }
      }
      act0[i * NT_E + lane] = out;
      if (fl) atomicOr(&flags[lane], fl);
    }
    __syncthreads();
    double* cur = act0; double* nxt = act1;
    for (int l = 0; l < nd.n_layers; l++) {
      const NeuralLayer L = nd.layers[l]; const double* B = nd.biases + L.b_off;
      const int groups = L.n_out_pad / 8, rpc = rows_per_chunk(L);
      double z[NT_G][8];
#pragma unroll
      for (int g = 0; g < NT_G; g++) {
        int jg = warp + g * NT_WARPS;
#pragma unroll
        for (int k = 0; k < 8; k++) { int j = jg * 8 + k; z[g][k] = (jg < groups && j < L.n_out) ? B[j] : 0.0; }
      }
      for (int i0 = 0; i0 < L.n_in; i0 += rpc) {
        const int rows = min(rpc, L.n_in - i0);
        uint32_t slot = q_wait & 1u; mbar_wait(&bars[slot], (q_wait >> 1) & 1u); q_wait++;
        const double* wbase = ring + (size_t)slot * (NT_RING / 8);
        const double* xcol = cur + (size_t)i0 * NT_E + lane;
        const int ng = groups > warp ? min(NT_G, (groups - warp + NT_WARPS - 1) / NT_WARPS) : 0;   // warp-uniform
        switch (ng) {  // no branches inside the row loop
          case 2: nt_rows<2>(z, xcol, wbase, rows, L.n_out_pad, warp); break;
          case 1: nt_rows<1>(z, xcol, wbase, rows, L.n_out_pad, warp); break;
          default: break;
        }
        __syncthreads();   // ring slot consumed
        if (tid == 0) { fence_proxy_async(); issue_next(); }
      }
#pragma unroll
      for (int g = 0; g < NT_G; g++) {
        int jg = warp + g * NT_WARPS;
#pragma unroll
        for (int k = 0; k < 8; k++) { int j = jg * 8 + k; if (jg < groups && j < L.n_out) nxt[j * NT_E + lane] = activate(L.act, z[g][k], L.threshold); }
      }
      __syncthreads();
      if (L.norm != NORM_NONE && warp == 0) {
        if (L.norm == NORM_SOFTMAX) {
          double s = 0; for (int j = 0; j < L.n_out; j++) { double ex = exp(nxt[j * NT_E + lane]); nxt[j * NT_E + lane] = ex; s += ex; }
          for (int j = 0; j < L.n_out; j++) nxt[j * NT_E + lane] /= s;
        } else {
          double s = 0; for (int j = 0; j < L.n_out; j++) s += nxt[j * NT_E + lane];
          for (int j = 0; j < L.n_out; j++) nxt[j * NT_E + lane] /= s;
        }
      }
      __syncthreads();
      double* t = cur; cur = nxt; nxt = t;
    }
    if (warp == 0 && row < a.n) {
      int fl = flags[lane];
      if (row_fails_preparation(m, a, row)) fl |= 1;
      if (fl & 1) write_out(a, row, 0, ST_PREP, -1);
      else if (fl & 2) write_out(a, row, 0, m.target_status ? m.target_status : ST_EVAL, -1);
      else if (!nd.classify) {
        double y = cur[nd.out_neuron[0] * NT_E + lane];
        if (nd.out_denorm_n) {
          const double* pts = nd.pts + 2 * nd.out_denorm_off; int n = nd.out_denorm_n, i = 0;
          while (i + 2 < n && y > pts[2 * (i + 1) + 1]) i++;
          double a0 = pts[2 * i], b0 = pts[2 * i + 1], a1 = pts[2 * i + 2], b1 = pts[2 * i + 3];
          y = (y - b0) / (b1 - b0) * (a1 - a0) + a0;
        }
        write_out(a, row, apply_target(m.tx, y), m.target_status, -1);
      } else {
        int best = -1; double bp = 0;
        for (int o = 0; o < nd.n_outputs; o++) {
          double p = cur[nd.out_neuron[o] * NT_E + lane];
          if (a.probs && o < a.prob_stride) a.probs[row * (long)a.prob_stride + o] = p;
          if (best < 0 || java_cmp(p, bp) > 0) { best = o; bp = p; }
        }
        write_label(m, a, row, best, best);
      }
    }
    __syncthreads();   // activations / flags are reused by the next tile
  }
}

// ---- the same streamed-weight scheme with TWO events per thread ------------------------------------------------------
// With one event per thread a row of the weight matrix costs a warp 8 broadcast LDS.128 (two shared-memory wavefronts each)
// + 1 activation load for 32 FP64 instructions: 20 wavefronts against 16 FP64 cycles.  Here lane l owns events l and l + 32
// of a 64-event tile (16 warps x NP_G = 2 groups of 8 output neurons = 256 outputs, as above), so the weight loads are
// shared by twice the arithmetic (8 warps x 4 groups was slower: too few warps to cover the latencies).  One activation buffer [neuron][64]: a layer's outputs stay in registers until every
// warp has finished reading its inputs, then overwrite them in place.  Sums run from the bias in connection order, unfused.
constexpr int NP_E = 64, NP_WARPS = 16, NP_G = 2;
template <int NG>
__device__ __forceinline__ void np_rows(double (&z)[NP_G][2][8], const double* __restrict__ xcol, const double* __restrict__ wbase, int rows, int n_out_pad, int warp) {
  if constexpr (NG <= NP_G)
#pragma unroll 2
  for (int r = 0; r < rows; r++) {
    const double x0 = xcol[r * NP_E], x1 = xcol[r * NP_E + 32];
    const double2* w2 = reinterpret_cast<const double2*>(wbase + (size_t)r * n_out_pad);
#pragma unroll
    for (int g = 0; g < NG; g++) {
      const int jg = warp + g * NP_WARPS;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        double2 w = w2[jg * 4 + k];
        z[g][0][2 * k] += w.x * x0; z[g][0][2 * k + 1] += w.y * x0;
        z[g][1][2 * k] += w.x * x1; z[g][1][2 * k + 1] += w.y * x1;
      }
    }
  }
}
struct NeuralPairPlan { uint32_t off_ring, off_bars, off_flags, smem_bytes; };
__global__ void __launch_bounds__(32 * NP_WARPS, 1) neural_pair_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const NeuralPairPlan pl) {
  extern __shared__ __align__(128) uint8_t smem[];
  const NeuralDesc& nd = m.nn;
  double* act = reinterpret_cast<double*>(smem);                         // [neuron][NP_E]
  double* ring = reinterpret_cast<double*>(smem + pl.off_ring);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + pl.off_bars);
  int* flags = reinterpret_cast<int*>(smem + pl.off_flags);              // NP_E
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  uint32_t q_issue = 0, q_wait = 0; int il = 0, ic = 0;   // producer cursor (thread 0): layer, first row of the next chunk
  auto rows_per_chunk = [&](const NeuralLayer& L) { int r = NT_RING / (L.n_out_pad * 8); return r < 1 ? 1 : (r > L.n_in ? L.n_in : r); };
  auto issue_next = [&]() {   // thread 0: next chunk of the per-tile sequence into ring slot q_issue & 1
    if (il >= nd.n_layers) return;
    const NeuralLayer& L = nd.layers[il]; int rpc = rows_per_chunk(L); int rows = min(rpc, L.n_in - ic);
    uint32_t bytes = (uint32_t)rows * L.n_out_pad * 8u, slot = q_issue & 1u;
    mbar_expect_tx(&bars[slot], bytes);
    bulk_g2s(ring + (size_t)slot * (NT_RING / 8), nd.weights_t + L.wt_off + (size_t)ic * L.n_out_pad, bytes, &bars[slot]);
    q_issue++; ic += rows; if (ic >= L.n_in) { ic = 0; il++; }
  };
  // re-scoring mode (a.row_map): work item i is row row_map[i], *n_dev of them (the events the tensor-core path handed back)
  const long n_items = a.n_dev ? min((long)*a.n_dev, a.n) : a.n;
  auto row_of = [&](long item) { return a.row_map ? (long)a.row_map[item] : item; };
  const long ntiles = (n_items + NP_E - 1) / NP_E;
  for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long row0 = tile * NP_E;
    if (tid == 0) { il = 0; ic = 0; fence_proxy_async(); issue_next(); issue_next(); }
    if (tid < NP_E) flags[tid] = 0;
    __syncthreads();
    for (int i = warp; i < nd.n_inputs; i += NP_WARPS) {
      NeuralInput in = nd.inputs[i];
#pragma unroll
      for (int v = 0; v < 2; v++) {
        const int e = lane + 32 * v; const long item = row0 + e; double out = 0; int fl = 0;
        if (item < n_items) {
          const long row = row_of(item);
          if (in.kind == NI_CONSTANT) out = in.value;
          else {
            double val; int st = in.slot < 0 ? 1 : load_field(m, a, row, in.slot, val);
            if (st == 2) fl = 1;
            else if (st == 1) { if (in.has_map_missing) out = in.map_missing; else fl = 2; }
            else if (in.kind == NI_FIELDREF) out = val;
            else if (in.kind == NI_NORMDISC) out = java_eq(val, in.value) ? 1.0 : 0.0;
            else { bool miss = false; out = norm_cont(nd.pts + 2 * in.pts_off, in.n_pts, in.outliers, val, miss); if (miss) fl = 2; }
          }
        }
        act[i * NP_E + e] = out;
        if (fl) atomicOr(&flags[e], fl);
      }
    }
    __syncthreads();
    for (int l = 0; l < nd.n_layers; l++) {
      const NeuralLayer L = nd.layers[l]; const double* B = nd.biases + L.b_off;
      const int groups = L.n_out_pad / 8, rpc = rows_per_chunk(L);
      double z[NP_G][2][8];
#pragma unroll
      for (int g = 0; g < NP_G; g++) {
        int jg = warp + g * NP_WARPS;
#pragma unroll
        for (int k = 0; k < 8; k++) { int j = jg * 8 + k; double b = (jg < groups && j < L.n_out) ? B[j] : 0.0; z[g][0][k] = b; z[g][1][k] = b; }
      }
      const int ng = groups > warp ? min(NP_G, (groups - warp + NP_WARPS - 1) / NP_WARPS) : 0;   // warp-uniform
      for (int i0 = 0; i0 < L.n_in; i0 += rpc) {
        const int rows = min(rpc, L.n_in - i0);
        uint32_t slot = q_wait & 1u; mbar_wait(&bars[slot], (q_wait >> 1) & 1u); q_wait++;
        const double* wbase = ring + (size_t)slot * (NT_RING / 8);
        const double* xcol = act + (size_t)i0 * NP_E + lane;
        switch (ng) {  // no branches inside the row loop
          case 4: np_rows<4>(z, xcol, wbase, rows, L.n_out_pad, warp); break;
          case 3: np_rows<3>(z, xcol, wbase, rows, L.n_out_pad, warp); break;
          case 2: np_rows<2>(z, xcol, wbase, rows, L.n_out_pad, warp); break;
          case 1: np_rows<1>(z, xcol, wbase, rows, L.n_out_pad, warp); break;
          default: break;
        }
        __syncthreads();   // ring slot consumed (after the last chunk: every warp is done reading the layer's inputs)
        if (tid == 0) { fence_proxy_async(); issue_next(); }
      }
#pragma unroll
      for (int g = 0; g < NP_G; g++) {
        int jg = warp + g * NP_WARPS;
#pragma unroll
        for (int k = 0; k < 8; k++) {
          int j = jg * 8 + k;
          if (jg < groups && j < L.n_out) { act[j * NP_E + lane] = activate(L.act, z[g][0][k], L.threshold); act[j * NP_E + lane + 32] = activate(L.act, z[g][1][k], L.threshold); }
        }
      }
      __syncthreads();
      if (L.norm != NORM_NONE && warp < 2) {
        double* col = act + warp * 32 + lane;
        if (L.norm == NORM_SOFTMAX) {
          double s = 0; for (int j = 0; j < L.n_out; j++) { double ex = exp(col[j * NP_E]); col[j * NP_E] = ex; s += ex; }
          for (int j = 0; j < L.n_out; j++) col[j * NP_E] /= s;
        } else {
          double s = 0; for (int j = 0; j < L.n_out; j++) s += col[j * NP_E];
          for (int j = 0; j < L.n_out; j++) col[j * NP_E] /= s;
        }
      }
      __syncthreads();
    }
    if (warp < 2) {
      const int e = warp * 32 + lane; const long item = row0 + e; const double* col = act + e;
      if (item < n_items) {
        const long row = row_of(item);
        int fl = flags[e];
        if (row_fails_preparation(m, a, row)) fl |= 1;
        if (fl & 1) write_out(a, row, 0, ST_PREP, -1);
        else if (fl & 2) write_out(a, row, 0, m.target_status ? m.target_status : ST_EVAL, -1);
        else if (!nd.classify) {
          double y = col[nd.out_neuron[0] * NP_E];
          if (nd.out_denorm_n) {
            const double* pts = nd.pts + 2 * nd.out_denorm_off; int n = nd.out_denorm_n, i = 0;
            while (i + 2 < n && y > pts[2 * (i + 1) + 1]) i++;
            double a0 = pts[2 * i], b0 = pts[2 * i + 1], a1 = pts[2 * i + 2], b1 = pts[2 * i + 3];
            y = (y - b0) / (b1 - b0) * (a1 - a0) + a0;
          }
          write_out(a, row, apply_target(m.tx, y), m.target_status, -1);
        } else {
          int best = -1; double bp = 0;
          for (int o = 0; o < nd.n_outputs; o++) {
            double p = col[nd.out_neuron[o] * NP_E];
            if (a.probs && o < a.prob_stride) a.probs[row * (long)a.prob_stride + o] = p;
            if (best < 0 || java_cmp(p, bp) > 0) { best = o; bp = p; }
          }
          write_label(m, a, row, best, best);
        }
      }
    }
    __syncthreads();   // activations / flags are reused by the next tile
  }
}

// ================================================================================================ launchers
#include "forest_compact.cuh"

// Device facts the launch plans depend on.  One process may drive several GPUs from several host threads (one subtask per
// GPU, fjp_model_replicate), so the cache is per device ordinal and the values a call works with are thread-local copies
// refreshed from it at every entry point.
static thread_local int g_smem_optin = 0, g_num_sms = 0;
static void query_device() {
  struct Facts { std::atomic<int> ready{0}; int smem_optin = 0, num_sms = 0; };
  static Facts facts[64];
  int dev = 0; cudaGetDevice(&dev);
  Facts& f = facts[dev & 63];
  if (!f.ready.load(std::memory_order_acquire)) {
    static std::mutex mu; std::lock_guard<std::mutex> lk(mu);
    if (!f.ready.load(std::memory_order_relaxed)) {
      cudaDeviceGetAttribute(&f.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
      cudaDeviceGetAttribute(&f.num_sms, cudaDevAttrMultiProcessorCount, dev);
      f.ready.store(1, std::memory_order_release);
    }
  }
  g_smem_optin = f.smem_optin; g_num_sms = f.num_sms;
}

template <int D, int TI, int MODE, bool DBG, bool SLICE = false>
static cudaError_t launch_fc(const DeviceModel& m, const BatchArgs& a, const ForestLaunch& fl, int grid, int block, cudaStream_t s) {
  auto k = forest_complete_kernel<D, TI, MODE, DBG, SLICE>;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fl.smem_bytes);
  if (e != cudaSuccess) return e;
  k<<<dim3(grid, SLICE ? fl.slices : 1), block, fl.smem_bytes, s>>>(m, a, fl);
  return cudaGetLastError();
}

int env_int(const char* name, int dflt) { const char* v = getenv(name); return v && *v ? atoi(v) : dflt; }
void device_facts(int* num_sms, int* smem_optin) { query_device(); *num_sms = g_num_sms; *smem_optin = g_smem_optin; }

// Picks the tile geometry for the on-chip forest kernel; returns false when no tile of >= 64 events fits.
// FJP_FOREST_E / FJP_FOREST_G / FJP_FOREST_TI override the choice (tuning aid).
bool plan_forest_complete(const DeviceModel& m, ForestLaunch* out) {
  query_device();
  const ForestDesc& fd = m.forest;
  if (fd.format != FOREST_COMPLETE) return false;
  if (fd.classify && (fd.method != FM_VOTE && fd.method != FM_SINGLE)) return false;
  if (fd.classify && m.n_labels > 64) return false;
  const size_t budget = (size_t)g_smem_optin - 1024;
  const size_t blob = (size_t)(2u << fd.depth) * 8;
  const int S = m.F | 1;  // odd row stride (in doubles): conflict-free column reads of the row stage
  const size_t rows = (size_t)32 * S * 8;
  int ti = env_int("FJP_FOREST_TI", 8); ti = ti >= 12 ? 12 : (ti >= 8 ? 8 : (ti >= 6 ? 6 : 4));
  const int force_e = env_int("FJP_FOREST_E", 0), force_g = env_int("FJP_FOREST_G", 0);
  const int t_round = ((fd.n_trees + ti - 1) / ti) * ti;
  for (int E = 512; E >= 64; E -= 32) {
    if (force_e && E != force_e) continue;
    size_t keys = ((size_t)m.F * E * 4 + 127) & ~size_t(127);
    size_t votes = fd.classify ? (size_t)2 * m.n_labels * E * 2 : (fd.method == FM_CHAIN ? (size_t)fd.n_channels * E * 8 : 0);
    size_t fixed = keys + votes + 64 + E + 256;
    if (fixed + 2 * ti * blob > budget) continue;
    int G = (int)((budget - fixed) / (2 * blob)); G = (G / ti) * ti; if (G > 16) G = 16 / ti * ti; if (G > 2 * ti) G = 2 * ti;
    if (force_g) G = force_g;
    if (G > t_round) G = t_round;
    if (G < ti || fixed + 2 * (size_t)G * blob > budget) continue;
    size_t ring = 2 * (size_t)G * blob;
    bool alias = rows <= ring;
    if (!alias && fixed + ring + rows > budget) continue;
    // prefer more resident warps, but never fewer than 8 trees in the ring
    ForestLaunch fl{}; fl.events_per_cta = E; fl.trees_per_stage = G; fl.row_stride = S; fl.trees_interleaved = ti;
    size_t off = keys;
    fl.off_ring = (uint32_t)off; off += ring;
    if (alias) fl.off_rows = fl.off_ring; else { fl.off_rows = (uint32_t)off; off += rows; }
    off = (off + 15) & ~size_t(15); fl.off_votes = (uint32_t)off; off += votes; off = (off + 15) & ~size_t(15);
    fl.off_bars = (uint32_t)off; off += 64;
    fl.off_fail = (uint32_t)off; off += E;
    fl.smem_bytes = (uint32_t)((off + 127) & ~size_t(127));
    *out = fl; return true;
  }
  return false;
}

// Work area the small-batch forest path wants for n events (0: path not applicable)
size_t forest_compact_scratch_bytes(const DeviceModel& m, long n);
size_t dense_fma_scratch_bytes(const DeviceModel& m, const BatchArgs& a);
size_t forest_slice_scratch_bytes(const DeviceModel& m, long n) {
  query_device();
  if (size_t c = forest_compact_scratch_bytes(m, n)) return c;   // large batch: the work area holds the rank-key tiles instead
  if (m.kind != MK_FOREST || (m.forest.method != FM_SUM && m.forest.method != FM_AVERAGE)) return 0;
  ForestLaunch fl;
  if (!plan_forest_complete(m, &fl)) return 0;
  long ntiles = (n + fl.events_per_cta - 1) / fl.events_per_cta;
  if (ntiles * 2 > g_num_sms) return 0;                       // enough tiles to fill the chip: the one-kernel walk is the right one
  size_t bytes = (size_t)n * m.forest.n_trees * 8;
  return bytes <= (size_t(32) << 20) ? bytes : 0;
}

// ---- throughput path (forest_compact.cuh): geometry, work area, launch --------------------------------------
// FJP_FOREST_COMPACT=0 switches the path off, =2 takes it for every batch size; FJP_FOREST_CE / FJP_FOREST_CG / FJP_FOREST_CTI
// override the tile geometry (tuning aids, read at every launch).
// pin_e > 0: tile size to use (else FJP_FOREST_CE, else the first that fits)
static bool plan_forest_compact_e(const DeviceModel& m, ForestLaunch* out, int pin_e) {
  query_device();
  const ForestDesc& fd = m.forest;
  const bool sparse = fd.compact.sparse != 0;   // explicit-children trees: one group of trees per ring slot
  if (!fd.compact.enabled || (fd.format != FOREST_COMPLETE && !sparse) || !env_int("FJP_FOREST_COMPACT", 1)) return false;
  const bool avg_prob = fd.classify && (fd.method == FM_AVG_PROB || fd.method == FM_WAVG_PROB || fd.method == FM_WVOTE) && fd.compact.leaf_bytes > 8 && m.n_labels <= 16;
  if (fd.classify ? (fd.method != FM_VOTE && !avg_prob) : (fd.method != FM_SUM && fd.method != FM_AVERAGE && fd.method != FM_WSUM && fd.method != FM_WAVERAGE && fd.method != FM_CHAIN)) return false;
  if (fd.classify && m.n_labels > 64) return false;
  const size_t budget = (size_t)g_smem_optin - 1024, TB = sparse ? (size_t)fd.compact.group_max_bytes : (size_t)fd.compact.tree_bytes;
  int ti = env_int("FJP_FOREST_CTI", sparse ? 4 : 8); ti = ti >= 8 ? 8 : 4;
  if (sparse) {
    const int force_e = pin_e ? pin_e : env_int("FJP_FOREST_CE", 0);
    static const int kEs[] = {640, 768, 1024, 576, 512, 448, 384, 320, 256, 192, 128, 64};
    for (int E : kEs) {
      if (force_e && E != force_e) continue;
      if (E == 1024) ti = 4;
      size_t keys = ((size_t)((m.F + 1) / 2) * E * 4 + 127) & ~size_t(127);
      size_t votes = avg_prob ? (size_t)m.n_labels * E * 8 : (fd.classify ? (size_t)2 * m.n_labels * E * 2 : (fd.method == FM_CHAIN ? (size_t)fd.n_channels * E * 8 : 0));
      size_t off = keys + 2 * TB; if (off + votes + 64 + 256 > budget) continue;
      ForestLaunch fl{}; fl.events_per_cta = E; fl.trees_per_stage = 1; fl.trees_interleaved = ti;
      fl.off_ring = (uint32_t)keys; off = (off + 15) & ~size_t(15); fl.off_votes = (uint32_t)off; off += votes; off = (off + 15) & ~size_t(15);
      fl.off_bars = (uint32_t)off; off += 64; fl.smem_bytes = (uint32_t)((off + 127) & ~size_t(127));
      *out = fl; return true;
    }
    return false;
  }
  const int force_e = pin_e ? pin_e : env_int("FJP_FOREST_CE", 0), force_g = env_int("FJP_FOREST_CG", 0);
  const int t_round = ((fd.n_trees + ti - 1) / ti) * ti;
  static const int kE[] = {640, 768, 1024, 576, 512, 448, 384, 320, 256, 192, 128, 64};
  for (int E : kE) {
    if (force_e && E != force_e) continue;
    if (E == 1024) ti = 4;   // 64 registers per thread
    size_t keys = ((size_t)((m.F + 1) / 2) * E * 4 + 127) & ~size_t(127);
    size_t votes = avg_prob ? (size_t)m.n_labels * E * 8 : (fd.classify ? (size_t)2 * m.n_labels * E * 2 : (fd.method == FM_CHAIN ? (size_t)fd.n_channels * E * 8 : 0));
    size_t fixed = keys + votes + 64 + 256;
    if (fixed + 2 * ti * TB > budget) continue;
    int G = (int)((budget - fixed) / (2 * TB)); G = (G / ti) * ti; if (G > 2 * ti) G = 2 * ti;
    if (force_g) G = force_g;
    if (G > t_round) G = t_round;
    if (G < ti || fixed + 2 * (size_t)G * TB > budget) continue;
    ForestLaunch fl{}; fl.events_per_cta = E; fl.trees_per_stage = G; fl.trees_interleaved = ti;
    size_t off = keys; fl.off_ring = (uint32_t)off; off += 2 * (size_t)G * TB;
    off = (off + 15) & ~size_t(15); fl.off_votes = (uint32_t)off; off += votes; off = (off + 15) & ~size_t(15);
    fl.off_bars = (uint32_t)off; off += 64;
    fl.smem_bytes = (uint32_t)((off + 127) & ~size_t(127));
    *out = fl; return true;
  }
  return false;
}
// n = batch size.  Ensembles the one-kernel walk cannot take (explicit-children trees, averaged class distributions) pick,
// for mid-size batches, the largest tile whose tiles still cover the SMs instead of falling back to the global-memory walk.
static bool plan_forest_compact(const DeviceModel& m, ForestLaunch* out, long n) {
  query_device();
  ForestLaunch probe;
  const bool one_kernel = m.forest.method != FM_SINGLE && plan_forest_complete(m, &probe);
  if (one_kernel || env_int("FJP_FOREST_CE", 0)) return plan_forest_compact_e(m, out, 0);
  static const int kFit[] = {640, 512, 384, 320, 256, 192, 128, 64};
  for (int E : kFit) if (((n + E - 1) / E >= g_num_sms || E == 64) && plan_forest_compact_e(m, out, E)) return true;
  return false;
}
// tiles scored per pass over the work area: a multiple of the SM count, at most ~64 MB of keys -- written by the pre-pass
// and read back by the walk kernel while still resident in the 126 MB L2, so they cost no HBM traffic
static long compact_tiles_per_pass(const DeviceModel& m, const ForestLaunch& fl, long ntiles) {
  const size_t tile_bytes = (size_t)((m.F + 1) / 2) * fl.events_per_cta * 4;
  long cap = (long)((size_t(64) << 20) / tile_bytes); cap = cap / g_num_sms * g_num_sms; if (cap < g_num_sms) cap = g_num_sms;
  if (int per_sm = env_int("FJP_FOREST_CPASS", 0)) cap = (long)per_sm * g_num_sms;   // tuning aid: tiles per SM and pass
  return ntiles < cap ? ntiles : cap;
}
size_t forest_compact_scratch_bytes(const DeviceModel& m, long n) {
  if (m.kind != MK_FOREST) return 0;
  ForestLaunch fl;
  if (!plan_forest_compact(m, &fl, n)) return 0;
  const long ntiles = (n + fl.events_per_cta - 1) / fl.events_per_cta;
  ForestLaunch probe; const bool one_kernel = m.forest.method != FM_SINGLE && plan_forest_complete(m, &probe);
  // too few tiles to fill the chip: the one-kernel walk / the sliced latency path take it (2 = always, for tests); ensembles
  // without a one-kernel walk come here from 16 tiles on
  if (ntiles < (one_kernel ? g_num_sms : 16) && env_int("FJP_FOREST_COMPACT", 1) != 2) return 0;
  const long pass = compact_tiles_per_pass(m, fl, ntiles);
  return (size_t)pass * ((m.F + 1) / 2) * fl.events_per_cta * 4 + (size_t)pass * fl.events_per_cta;
}
template <int D, int TI, int MODE, int MAXT>
static cudaError_t launch_fcc(const DeviceModel& m, const BatchArgs& a, const ForestLaunch& fl, const KeyTiles& kt, int grid, cudaStream_t s) {
  auto k = forest_compact_kernel<D, TI, MODE, MAXT>;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fl.smem_bytes);
  if (e != cudaSuccess) return e;
  k<<<grid, fl.events_per_cta, fl.smem_bytes, s>>>(m, a, fl, kt);
  return cudaGetLastError();
}
template <int TI, int MODE, int MAXT>
static cudaError_t launch_fsp(const DeviceModel& m, const BatchArgs& a, const ForestLaunch& fl, const KeyTiles& kt, int grid, cudaStream_t s) {
  auto k = forest_sparse_kernel<TI, MODE, MAXT>;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fl.smem_bytes);
  if (e != cudaSuccess) return e;
  k<<<grid, fl.events_per_cta, fl.smem_bytes, s>>>(m, a, fl, kt);
  return cudaGetLastError();
}
static cudaError_t launch_forest_compact(const DeviceModel& m, const BatchArgs& a, const ForestLaunch& fl, void* scratch, cudaStream_t s, int* launches) {
  const CompactDesc& cd = m.forest.compact;
  const int E = fl.events_per_cta; const long ntiles = (a.n + E - 1) / E, pass = compact_tiles_per_pass(m, fl, ntiles);
  KeyTiles kt{}; kt.keys = static_cast<uint32_t*>(scratch); kt.E = E; kt.tile_bytes = (uint32_t)((size_t)((m.F + 1) / 2) * E * 4);
  kt.uniform_mode = m.forest.compact.uniform_mode;
  uint8_t* fail = static_cast<uint8_t*>(scratch) + (size_t)pass * kt.tile_bytes;
  const size_t rk_smem = (size_t)cd.block_smem;   // the feature block's tables
  const int mode = m.forest.classify ? (m.forest.method == FM_VOTE ? 2 : 4) : (m.forest.method == FM_CHAIN ? 3 : ((m.forest.method == FM_WSUM || m.forest.method == FM_WAVERAGE) ? 1 : 0));
  const int d = m.forest.depth;
  for (long t0 = 0; t0 < ntiles; t0 += pass) {
    kt.row_begin = t0 * E; kt.ntiles = ntiles - t0 < pass ? ntiles - t0 : pass;
    kt.fail = m.prep_simple ? nullptr : fail;
    cudaError_t e;
    if (kt.fail) { e = cudaMemsetAsync(fail, 0, (size_t)kt.ntiles * E, s); if (e != cudaSuccess) return e; }
    int gx = RK_CTAS * g_num_sms / cd.n_blocks; if (gx < 1) gx = 1; if (gx > kt.ntiles) gx = (int)kt.ntiles;   // one wave of resident CTAs
#define FJP_RK(SS, WW) do { cudaFuncSetAttribute(forest_rank_keys_kernel<SS, WW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rk_smem); \
                            forest_rank_keys_kernel<SS, WW><<<dim3(gx, cd.n_blocks), RK_THREADS, rk_smem, s>>>(m, a, kt); } while (0)
    if (m.prep_simple) { if (cd.wide) FJP_RK(true, true); else FJP_RK(true, false); }
    else { if (cd.wide) FJP_RK(false, true); else FJP_RK(false, false); }
#undef FJP_RK
    e = cudaGetLastError(); if (e != cudaSuccess) return e;
    if (launches) (*launches)++;
    const int grid = (int)(kt.ntiles < g_num_sms ? kt.ntiles : g_num_sms);
#define FJP_CC3(DD, TT, MM, XX) launch_fcc<DD, TT, MM, XX>(m, a, fl, kt, grid, s)
#define FJP_CC2(DD, MM) (E > 768 ? FJP_CC3(DD, 4, MM, 1024) : (E > 640 ? (fl.trees_interleaved >= 8 ? FJP_CC3(DD, 8, MM, 768) : FJP_CC3(DD, 4, MM, 768)) \
                                                                         : (fl.trees_interleaved >= 8 ? FJP_CC3(DD, 8, MM, 640) : FJP_CC3(DD, 4, MM, 640))))
#define FJP_CC1(DD) (mode == 2 ? FJP_CC2(DD, 2) : (mode == 4 ? FJP_CC2(DD, 4) : (mode == 3 ? FJP_CC2(DD, 3) : (mode == 1 ? FJP_CC2(DD, 1) : FJP_CC2(DD, 0)))))
    if (cd.sparse) {
#define FJP_SP3(TT, MM, XX) launch_fsp<TT, MM, XX>(m, a, fl, kt, grid, s)
#define FJP_SP2(MM) (E > 768 ? FJP_SP3(4, MM, 1024) : (E > 640 ? FJP_SP3(4, MM, 768) : (fl.trees_interleaved >= 8 ? FJP_SP3(8, MM, 640) : FJP_SP3(4, MM, 640))))
      e = mode == 2 ? FJP_SP2(2) : (mode == 4 ? FJP_SP2(4) : (mode == 3 ? FJP_SP2(3) : (mode == 1 ? FJP_SP2(1) : FJP_SP2(0))));
#undef FJP_SP2
#undef FJP_SP3
    } else
    e = d == 8 ? FJP_CC1(8) : FJP_CC1(0);
#undef FJP_CC1
#undef FJP_CC2
#undef FJP_CC3
    if (e != cudaSuccess) return e;
    if (launches && t0 + pass < ntiles) (*launches)++;
  }
  return cudaSuccess;
}

// ---- dense tile kernel (RegressionModel tables / k-means centres): geometry + dispatch ----------------------------------
// fast: 0 = the reference's unfused arithmetic, 1 = FMA + margin (softmax classification), 2 = FMA + margin (nearest centre)
static cudaError_t launch_dense_tile(const DeviceModel& m, const BatchArgs& a, const DenseView& dv, bool hasmiss, int fast, cudaStream_t s) {
  // tables per thread: 8, 10 or 12, whichever pads the table count least (20 tables -> 2 groups of 10); the same rule lays out wt
  int tc = 8, best_pad = 1 << 30;
  for (int cand : {8, 10, 12}) { int pad = (dv.C + cand - 1) / cand * cand - dv.C; if (pad < best_pad) { best_pad = pad; tc = cand; } }
  DensePlan pl{}; pl.kp = ((dv.C + tc - 1) / tc) * tc; pl.groups = pl.kp / tc;
  // two events per thread once there are enough fields for the coefficient traffic to matter
  // nearest centre (fast >= 2): four events per thread -- 8 weights serve 32 fused multiply-adds, so the coefficient loads stop pacing the FP64 pipe
  const int ept = fast >= 2 ? env_int("FJP_KM_EPT", 4) : (fast ? 2 : (env_int("FJP_REG_EPT", m.F >= 64 ? 2 : 1) >= 2 ? 2 : 1)), per_sm = ept >= 2 ? 2 : 3;
  pl.event_blocks = pl.groups >= 8 ? 1 : 8 / pl.groups; pl.warps = pl.event_blocks * pl.groups; pl.events = 32 * ept * pl.event_blocks;
  const int EP = pl.events + 1;
  int chunk = 64; while (chunk > m.F && chunk > 1) chunk /= 2;   // power of two: cheap row/column split
  const unsigned smem_cap = (unsigned)env_int("FJP_REG_SMEM_KB", ept >= 2 ? 110 : 72) * 1024u;   // per_sm CTAs per SM: one loads while the others compute
  for (;; chunk /= 2) {
    size_t off = ((size_t)2 * chunk * pl.kp * 8 + 127) & ~size_t(127); pl.off_x = (uint32_t)off;
    const size_t xbytes = (size_t)2 * chunk * EP * 8, ybytes = (size_t)pl.kp * EP * 8;
    pl.off_y = (uint32_t)off; off += xbytes > ybytes ? xbytes : ybytes;   // ys (epilogue only) reuses the input buffers
    pl.off_st = (uint32_t)off; if (hasmiss) off += (size_t)2 * chunk * (pl.events + 4); off = (off + 15) & ~size_t(15);
    pl.off_flags = (uint32_t)off; off += (size_t)pl.events * (fast == 3 ? 20 : 4); pl.smem_bytes = (uint32_t)((off + 127) & ~size_t(127));
    if (pl.smem_bytes <= smem_cap || chunk == 1) break;
  }
  pl.chunk = chunk;
  if (pl.smem_bytes > (uint32_t)g_smem_optin || pl.warps * 32 > 256) return cudaErrorInvalidConfiguration;
  long ntiles = (a.n + pl.events - 1) / pl.events; int grid = (int)(ntiles < (long)per_sm * g_num_sms ? ntiles : (long)per_sm * g_num_sms);
#define FJP_RT(HM, TCV, EV, FV) do { cudaFuncSetAttribute(regression_tile_kernel<HM, TCV, EV, FV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes); \
                                     regression_tile_kernel<HM, TCV, EV, FV><<<grid, pl.warps * 32, pl.smem_bytes, s>>>(m, a, pl, dv); } while (0)
#define FJP_RT2(HM, TCV) do { if (ept == 2) FJP_RT(HM, TCV, 2, 0); else FJP_RT(HM, TCV, 1, 0); } while (0)
  if (fast == 1) { if (tc == 8) FJP_RT(false, 8, 2, 1); else if (tc == 10) FJP_RT(false, 10, 2, 1); else FJP_RT(false, 12, 2, 1); }
  else if (fast == 2 && ept == 4) { if (tc == 8) FJP_RT(false, 8, 4, 2); else if (tc == 10) FJP_RT(false, 10, 4, 2); else FJP_RT(false, 12, 4, 2); }
  else if (fast == 3 && ept == 4) { if (tc == 8) FJP_RT(false, 8, 4, 3); else if (tc == 10) FJP_RT(false, 10, 4, 3); else FJP_RT(false, 12, 4, 3); }
  else if (fast == 2) { if (tc == 8) FJP_RT(false, 8, 2, 2); else if (tc == 10) FJP_RT(false, 10, 2, 2); else FJP_RT(false, 12, 2, 2); }
  else if (fast == 3) { if (tc == 8) FJP_RT(false, 8, 2, 3); else if (tc == 10) FJP_RT(false, 10, 2, 3); else FJP_RT(false, 12, 2, 3); }
  else if (hasmiss) { if (tc == 8) FJP_RT2(true, 8); else if (tc == 10) FJP_RT2(true, 10); else FJP_RT2(true, 12); }
  else { if (tc == 8) FJP_RT2(false, 8); else if (tc == 10) FJP_RT2(false, 10); else FJP_RT2(false, 12); }
#undef FJP_RT2
#undef FJP_RT
  return cudaGetLastError();
}
// FMA paths of the dense tile kernel: softmax classification (RegressionModel) and nearest centre (ClusteringModel) over plain
// dense rows, batches that fill the chip.  Work area: the re-scoring counter + one row index per event.  FJP_DENSE_FMA=0 switches
// them off, =2 takes them for every batch size.
size_t dense_fma_scratch_bytes(const DeviceModel& m, const BatchArgs& a) {
  if (sliced_dense_takes(m, a)) return 256 + (((size_t)a.n * 4 + 255) & ~size_t(255));   // the fused tensor-core layer keeps the same re-scoring list
  const int mode = env_int("FJP_DENSE_FMA", 1);
  const bool masked_centres = a.present && m.kind == MK_CLUSTERING && m.F <= 128 && m.F >= 32;   // FAST = 3 of the tile kernel
  if (!mode || (a.present && !masked_centres) || a.replace || !m.prep_identity || a.row_map || a.node_ids) return 0;
  query_device();
  if (mode != 2 && a.n < 64L * g_num_sms) return 0;
  if (m.kind == MK_REGRESSION) { if (!(m.reg.dense && m.reg.classify && m.reg.norm == NORM_SOFTMAX && m.reg.n_tables <= REG_MAXC)) return 0; }
  else if (m.kind == MK_CLUSTERING) { if (!(m.cluster.fast && m.cluster.slots_identity && m.cluster.fast_wt && m.cluster.k <= REG_MAXC && !a.probs)) return 0; }
  else return 0;
  return 256 + (((size_t)a.n * 4 + 255) & ~size_t(255));
}

cudaError_t launch_score(const DeviceModel& m, const BatchArgs& a, cudaStream_t s, int* launches, void* slice_scratch) {
  query_device();
  if (a.n <= 0) return cudaSuccess;
  if (launches) (*launches)++;
  switch (m.kind) {
    case MK_FOREST: {
      ForestLaunch fl;
      if (slice_scratch && !a.node_ids && m.forest.method != FM_SINGLE && forest_compact_scratch_bytes(m, a.n) && plan_forest_compact(m, &fl, a.n))
        return launch_forest_compact(m, a, fl, slice_scratch, s, launches);
      // single trees go through the general path (it reports node ids in aux); ensembles through the on-chip path
      if (m.forest.method != FM_SINGLE && plan_forest_complete(m, &fl)) {
        long ntiles = (a.n + fl.events_per_cta - 1) / fl.events_per_cta;
        int grid = (int)(ntiles < g_num_sms ? ntiles : g_num_sms);
        if (slice_scratch && !a.node_ids && forest_slice_scratch_bytes(m, a.n)) {   // small batch: spread the trees over the idle SMs
          const int NG = (m.forest.n_trees + fl.trees_per_stage - 1) / fl.trees_per_stage;
          int slices = (int)(g_num_sms / ntiles); if (slices > NG) slices = NG;
          if (slices > 1) {
            fl.slices = slices; fl.leaf_out = static_cast<double*>(slice_scratch);
            cudaError_t e1 = m.forest.depth == 8 ? (fl.trees_interleaved >= 8 ? launch_fc<8, 8, 0, false, true>(m, a, fl, grid, fl.events_per_cta, s) : launch_fc<8, 4, 0, false, true>(m, a, fl, grid, fl.events_per_cta, s))
                                                  : (fl.trees_interleaved >= 8 ? launch_fc<0, 8, 0, false, true>(m, a, fl, grid, fl.events_per_cta, s) : launch_fc<0, 4, 0, false, true>(m, a, fl, grid, fl.events_per_cta, s));
            if (e1 != cudaSuccess) return e1;
            if (launches) (*launches)++;
            forest_ordered_sum_kernel<<<(unsigned)((a.n + 127) / 128), 128, 0, s>>>(m, a, fl.leaf_out);
            return cudaGetLastError();
          }
        }
        const bool c = m.forest.classify; const int d = m.forest.depth;
#define FJP_FC2(DD, TT, MM) (a.node_ids ? launch_fc<DD, 4, MM, true>(m, a, fl, grid, fl.events_per_cta, s) : launch_fc<DD, TT, MM, false>(m, a, fl, grid, fl.events_per_cta, s))
#define FJP_FC1(DD, MM) (fl.trees_interleaved >= 12 ? FJP_FC2(DD, 12, MM) : fl.trees_interleaved >= 8 ? FJP_FC2(DD, 8, MM) : (fl.trees_interleaved >= 6 ? FJP_FC2(DD, 6, MM) : FJP_FC2(DD, 4, MM)))
#define FJP_FC(DD) (c ? FJP_FC1(DD, 2) : (m.forest.method == FM_CHAIN ? FJP_FC1(DD, 3) : ((m.forest.method == FM_WSUM || m.forest.method == FM_WAVERAGE) ? FJP_FC1(DD, 1) : FJP_FC1(DD, 0))))
        if (d == 8) return FJP_FC(8);
        if (d == 6) return FJP_FC(6);
        return FJP_FC(0);
#undef FJP_FC2
#undef FJP_FC1
#undef FJP_FC
      }
      if (m.forest.classify && m.n_labels > 64) return cudaErrorInvalidValue;
      forest_general_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(m, a);
      return cudaGetLastError();
    }
    case MK_CLUSTERING: {
      const int nf = m.cluster.nf;
      if (slice_scratch && dense_fma_scratch_bytes(m, a)) {
        // |x - c|^2 = |x|^2 + (|c|^2 - 2 x.c): one fused multiply-add per term instead of sub, mul, add; the nearest centre is decided
        // here when the runner-up is farther than the rounding bound, the other events go to the exact kernel below
        DenseView dv{}; dv.wt = m.cluster.fast_wt; dv.icpt = m.cluster.fast_icpt; dv.C = m.cluster.k; dv.amax = m.cluster.cmax; dv.csq = m.cluster.fast_csq;
        dv.fb_count = static_cast<int32_t*>(slice_scratch); dv.fb_rows = dv.fb_count + 64;
        cudaError_t e = cudaMemsetAsync(dv.fb_count, 0, 4, s); if (e != cudaSuccess) return e;
        e = launch_dense_tile(m, a, dv, false, a.present ? 3 : 2, s); if (e != cudaSuccess) return e;
        if (launches) (*launches)++;
        BatchArgs fb = a; fb.row_map = dv.fb_rows; fb.n_dev = dv.fb_count;
        size_t sm = (size_t)nf * CL_XS * 8 + (((size_t)nf * CL_SS + 15) & ~size_t(15)) + CL_G * CL_E * 12;
        if (sm > (size_t)g_smem_optin) return cudaErrorInvalidValue;
        unsigned grid = (unsigned)std::min<long>((a.n + CL_E - 1) / CL_E, 4L * g_num_sms);
        cudaFuncSetAttribute(cluster_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cluster_kernel<true><<<grid, CL_E * CL_G, sm, s>>>(m, fb);
        return cudaGetLastError();
      }
      if (m.cluster.fast && (m.cluster.measure == CM_SQEUCLIDEAN || m.cluster.measure == CM_EUCLIDEAN)) {
        ClusterPlan pl{}; pl.kp = ((m.cluster.k + CK_TC - 1) / CK_TC) * CK_TC; pl.groups = pl.kp / CK_TC;
        pl.event_blocks = pl.groups >= CK_WARPS ? 1 : CK_WARPS / pl.groups;
        while (CK_WARPS % pl.event_blocks) pl.event_blocks--;
        pl.events = 32 * pl.event_blocks;
        size_t cent = ((size_t)nf * pl.kp * 8 + 127) & ~size_t(127);
        size_t room = (size_t)g_smem_optin - 2048 - cent - (size_t)(CK_WARPS / pl.event_blocks) * pl.events * 12;
        if (cent + 16384 < (size_t)g_smem_optin) {
          int chunk = (int)(room / ((size_t)(pl.events + 1) * 8 + pl.events + 4)); if (chunk > nf) chunk = nf; if (chunk > 128) chunk = 128;
          {  // two events per thread when the rows are plain doubles and the cluster groups fit 16 warps
            const bool plain = a.replace == nullptr && m.prep_identity && (a.present == nullptr || (m.cluster.slots_identity && a.present_stride <= 4));
            if (plain && pl.groups <= 16 && env_int("FJP_CLUSTER_PAIR", 1)) {
              // the group count is rounded up to a divisor of 16 (padding clusters never win): 3 groups run as 4, 5..7 as 8
              ClusterPairPlan pp{}; pp.groups = 1; while (pp.groups < pl.groups) pp.groups *= 2;
              pp.kp = pp.groups * CK_TC; pp.event_blocks = 16 / pp.groups; pp.events = 64 * pp.event_blocks;
              const size_t cent = ((size_t)nf * pp.kp * 8 + 127) & ~size_t(127);   // shadows the one-event geometry's figure
              const size_t fixed = cent + (size_t)2 * pp.events * 16 + (size_t)pp.groups * pp.events * 12 + 1024;
              for (pp.nchunks = 1; pp.nchunks <= 8; pp.nchunks++) {   // fewest chunks (of a multiple of 32 fields) whose two buffers fit
                pp.chunk = pp.nchunks == 1 ? nf : (((nf + pp.nchunks - 1) / pp.nchunks + 31) / 32) * 32;
                if (fixed + (size_t)2 * pp.chunk * (pp.events + 1) * 8 <= (size_t)g_smem_optin) break;
              }
              if (pp.nchunks <= 8 && (long)pp.chunk * (pp.nchunks - 1) < nf) {
                size_t off = cent; pp.off_x = (uint32_t)off; off += (size_t)2 * pp.chunk * (pp.events + 1) * 8;
                pp.off_p = (uint32_t)off; off += (size_t)2 * pp.events * 16;
                pp.off_best = (uint32_t)off; off += (size_t)pp.groups * pp.events * 12; pp.smem_bytes = (uint32_t)((off + 127) & ~size_t(127));
                long ntiles = (a.n + pp.events - 1) / pp.events; int grid = (int)(ntiles < g_num_sms ? ntiles : g_num_sms);
                if (a.present) { cudaFuncSetAttribute(cluster_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pp.smem_bytes); cluster_pair_kernel<true><<<grid, 512, pp.smem_bytes, s>>>(m, a, pp); }
                else { cudaFuncSetAttribute(cluster_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pp.smem_bytes); cluster_pair_kernel<false><<<grid, 512, pp.smem_bytes, s>>>(m, a, pp); }
                return cudaGetLastError();
              }
            }
          }
          if (chunk >= 1) {
            // plain double rows (with or without present bits) whose fields fit one chunk twice over, one cluster group per
            // warp: double-buffered tiles.  With present bits: fields = vector slots, at most 128 of them.
            const bool dense = a.replace == nullptr && m.prep_identity &&
                               (a.present == nullptr || (m.cluster.slots_identity && a.present_stride <= 4 && (size_t)nf * (pl.events + 4) >= (size_t)2 * pl.events * 16));   // the status area doubles as the present-word buffers
            pl.pipelined = dense && chunk >= nf && pl.groups * pl.event_blocks == CK_WARPS && (size_t)2 * nf * (pl.events + 1) * 8 + (size_t)nf * (pl.events + 4) <= room;
            if (pl.pipelined) chunk = nf;
            pl.chunk = chunk; size_t off = cent; pl.off_x = (uint32_t)off; off += (size_t)(pl.pipelined ? 2 : 1) * chunk * (pl.events + 1) * 8;
            pl.off_st = (uint32_t)off; off += (size_t)chunk * (pl.events + 4); off = (off + 15) & ~size_t(15);
            pl.off_best = (uint32_t)off; off += (size_t)(CK_WARPS / pl.event_blocks) * pl.events * 12;
            pl.smem_bytes = (uint32_t)((off + 127) & ~size_t(127));
            long ntiles = (a.n + pl.events - 1) / pl.events; int grid = (int)(ntiles < g_num_sms ? ntiles : g_num_sms);
            const bool hasmiss = a.present != nullptr || a.replace != nullptr || !m.prep_identity;
            if (hasmiss) { cudaFuncSetAttribute(cluster_fast_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes); cluster_fast_kernel<true><<<grid, CK_WARPS * 32, pl.smem_bytes, s>>>(m, a, pl); }
            else { cudaFuncSetAttribute(cluster_fast_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes); cluster_fast_kernel<false><<<grid, CK_WARPS * 32, pl.smem_bytes, s>>>(m, a, pl); }
            return cudaGetLastError();
          }
        }
      }
      size_t sm = (size_t)nf * CL_XS * 8 + (((size_t)nf * CL_SS + 15) & ~size_t(15)) + CL_G * CL_E * 12;
      if (sm > (size_t)g_smem_optin) return cudaErrorInvalidValue;
      unsigned grid = (unsigned)((a.n + CL_E - 1) / CL_E);
      if (m.cluster.fast) {
        cudaFuncSetAttribute(cluster_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cluster_kernel<true><<<grid, CL_E * CL_G, sm, s>>>(m, a);
      } else {
        cudaFuncSetAttribute(cluster_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cluster_kernel<false><<<grid, CL_E * CL_G, sm, s>>>(m, a);
      }
      return cudaGetLastError();
    }
    case MK_REGRESSION: {
      if (m.reg.n_tables > REG_MAXC) return cudaErrorInvalidValue;
      if (m.reg.dense) {
        DenseView dv{}; dv.wt = m.reg.dense_wt; dv.icpt = m.reg.intercept; dv.C = m.reg.n_tables; dv.amax = m.reg.dense_amax; dv.bmax = m.reg.dense_bmax;
        const bool hasmiss = a.present != nullptr || a.replace != nullptr || !m.prep_identity;
        // wide softmax models, labels only: the fused tensor-core layer (sliced_gemm.cuh); what it cannot decide is re-scored below
        if (slice_scratch && dense_fma_scratch_bytes(m, a) && sliced_dense_takes(m, a)) {
          dv.fb_count = static_cast<int32_t*>(slice_scratch); dv.fb_rows = dv.fb_count + 64;
          cudaError_t e = cudaMemsetAsync(dv.fb_count, 0, 4, s); if (e != cudaSuccess) return e;
          e = launch_sliced_dense(m, a, dv.fb_count, dv.fb_rows, s); if (e != cudaSuccess) return e;
          if (launches) (*launches)++;
          BatchArgs fb = a; fb.row_map = dv.fb_rows; fb.n_dev = dv.fb_count;
          return launch_dense_tile(m, fb, dv, true, 0, s);
        }
        // softmax classification over plain dense rows: fused multiply-add + margin test, the undecided events re-scored below
        if (slice_scratch && dense_fma_scratch_bytes(m, a)) {
          dv.fb_count = static_cast<int32_t*>(slice_scratch); dv.fb_rows = dv.fb_count + 64;
          cudaError_t e = cudaMemsetAsync(dv.fb_count, 0, 4, s); if (e != cudaSuccess) return e;
          e = launch_dense_tile(m, a, dv, false, 1, s); if (e != cudaSuccess) return e;
          if (launches) (*launches)++;
          BatchArgs fb = a; fb.row_map = dv.fb_rows; fb.n_dev = dv.fb_count;
          return launch_dense_tile(m, fb, dv, true, 0, s);
        }
        return launch_dense_tile(m, a, dv, hasmiss || a.row_map != nullptr, 0, s);
      }
      regression_kernel<<<(unsigned)((a.n + 127) / 128), 128, 0, s>>>(m, a);
      return cudaGetLastError();
    }
    case MK_GENERAL_REGRESSION: {
      general_regression_kernel<<<(unsigned)((a.n + 127) / 128), 128, 0, s>>>(m, a);
      return cudaGetLastError();
    }
    case MK_NEURAL: {
      { cudaError_t e; if (slice_scratch && launch_sliced(m, a, slice_scratch, s, launches, &e)) return e; }
      bool tiled = true;   // the streamed path keeps <= 4 groups of 8 outputs per warp: layers up to 256 neurons
      for (int l = 0; l < m.nn.n_layers; l++) tiled = tiled && m.nn.layers[l].n_out_pad <= 8 * NT_G * NT_WARPS && m.nn.layers[l].n_out_pad * 8 <= NT_RING
                                                           && m.nn.layers[l].act != ACT_RADIAL;   // radial basis layers: the general kernel only
      if (tiled && env_int("FJP_NN_PAIR", 1)) {   // two events per thread, one in-place activation buffer
        cudaError_t e = launch_neural_pair(m, a, s);
        if (e != cudaErrorInvalidConfiguration) return e;
      }
      if (tiled) {
        NeuralPlan pl{}; pl.width = m.nn.max_width; size_t act = ((size_t)m.nn.max_width * NT_E * 8 + 127) & ~size_t(127);
        pl.off_act1 = (uint32_t)act; pl.off_ring = (uint32_t)(2 * act); pl.off_bars = pl.off_ring + 2 * NT_RING; pl.off_flags = pl.off_bars + 64;
        pl.smem_bytes = pl.off_flags + 128;
        if (pl.smem_bytes <= (uint32_t)g_smem_optin) {
          long ntiles = (a.n + NT_E - 1) / NT_E; int per_sm = (int)((size_t)g_smem_optin / pl.smem_bytes); if (per_sm < 1) per_sm = 1; if (per_sm > 4) per_sm = 4;
          int grid = (int)(ntiles < (long)per_sm * g_num_sms ? ntiles : (long)per_sm * g_num_sms);
          cudaFuncSetAttribute(neural_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes);
          neural_tile_kernel<<<grid, NT_E * NT_WARPS, pl.smem_bytes, s>>>(m, a, pl);
          return cudaGetLastError();
        }
      }
      size_t sm = (size_t)2 * m.nn.max_width * NN_E * 8;
      if (sm > (size_t)g_smem_optin) return cudaErrorInvalidValue;
      cudaFuncSetAttribute(neural_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
      neural_kernel<<<(unsigned)((a.n + NN_E - 1) / NN_E), NN_E * NN_WARPS, sm, s>>>(m, a);
      return cudaGetLastError();
    }
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_neural_pair(const DeviceModel& m, const BatchArgs& a, cudaStream_t s) {
  query_device();
  NeuralPairPlan pp{}; size_t act = ((size_t)m.nn.max_width * NP_E * 8 + 127) & ~size_t(127);
  pp.off_ring = (uint32_t)act; pp.off_bars = pp.off_ring + 2 * NT_RING; pp.off_flags = pp.off_bars + 64; pp.smem_bytes = pp.off_flags + NP_E * 4;
  if (pp.smem_bytes > (uint32_t)g_smem_optin) return cudaErrorInvalidConfiguration;
  long ntiles = (a.n + NP_E - 1) / NP_E; int grid = (int)(ntiles < g_num_sms ? ntiles : g_num_sms);
  cudaFuncSetAttribute(neural_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pp.smem_bytes);
  neural_pair_kernel<<<grid, 32 * NP_WARPS, pp.smem_bytes, s>>>(m, a, pp);
  return cudaGetLastError();
}

// ================================================================================================ sparse ingest
// SparseVector rows arrive as CSR (row_ptr, indices, values) and are expanded ON THE DEVICE into the dense row +
// present-bit form every scoring kernel reads (absent index = cleared bit = missing value; an explicit 0.0 is present:
// api/converter/VectorConverter.scala:80-86, 100-103).  One warp per row; the expanded rows live in scratch memory that
// the caller provides (n x F doubles + n x ceil(F/32) words).  Indices outside [0, F) match no slot and are dropped, as
// in the reference's `(0 until size).map(indices.contains)`.
__global__ void __launch_bounds__(256) expand_sparse_kernel(const long long* __restrict__ row_ptr, const int32_t* __restrict__ idx,
                                                            const double* __restrict__ val, long n, int F, int W,
                                                            double* __restrict__ X, uint32_t* __restrict__ present) {
  const int lane = threadIdx.x & 31; const long row = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
  if (row >= n) return;
  double* xr = X + row * (long)F; uint32_t* pr = present + row * (long)W;
  for (int f = lane; f < F; f += 32) xr[f] = 0.0;
  for (int w = lane; w < W; w += 32) pr[w] = 0u;
  __syncwarp();
  const long long lo = row_ptr[row], hi = row_ptr[row + 1];
  for (long long k = lo + lane; k < hi; k += 32) {
    int i = idx[k];
    if (i >= 0 && i < F) { xr[i] = val[k]; atomicOr(&pr[i >> 5], 1u << (i & 31)); }
  }
}
cudaError_t launch_expand_sparse(const long long* row_ptr, const int32_t* idx, const double* val, long n, int F,
                                 double* X, uint32_t* present, cudaStream_t s, int* launches) {
  if (n <= 0) return cudaSuccess;
  if (launches) (*launches)++;
  expand_sparse_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, s>>>(row_ptr, idx, val, n, F, (F + 31) / 32, X, present);
  return cudaGetLastError();
}

__device__ double norm_cont_dev(const double* __restrict__ pts, int n, int outliers, double x, bool& missing) { return norm_cont(pts, n, outliers, x, missing); }

// float rows (fjp_score_batch_f32): widened on the device into the double rows every kernel reads (exact)
__global__ void __launch_bounds__(256) widen_f32_kernel(const float* __restrict__ src, double* __restrict__ dst, long count) {
  const long i = (blockIdx.x * (long)blockDim.x + threadIdx.x) * 4;
  if (i + 3 < count) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(src + i));
    reinterpret_cast<double2*>(dst + i)[0] = make_double2((double)v.x, (double)v.y); reinterpret_cast<double2*>(dst + i)[1] = make_double2((double)v.z, (double)v.w);
  } else for (long k = i; k < count; k++) dst[k] = (double)src[k];
}
cudaError_t launch_widen_f32(const float* src, double* dst, long count, cudaStream_t s, int* launches) {
  if (count <= 0) return cudaSuccess;
  if (launches) (*launches)++;
  widen_f32_kernel<<<(unsigned)((count / 4 + 256) / 256), 256, 0, s>>>(src, dst, count);
  return cudaGetLastError();
}

// ================================================================================================ derived fields
// Pre-pass for models with a TransformationDictionary / LocalTransformations: one thread per event prepares the F_in raw
// slots (the same preparation every kernel fuses otherwise), runs the DerivedField programs and writes the augmented row
// (F = F_in + n_derived values + present bits) into scratch memory; the family kernel then scores the augmented rows with
// an identity preparation, and fixup_kernel puts the pre-pass failures (preparation 2, evaluation 3) back.
__device__ bool dx_run(const DerivedDesc& dd, int d, const double* __restrict__ row, const uint32_t* __restrict__ bits, double& out) {
  double sv[DX_STACK]; bool sp[DX_STACK]; int top = 0;
  auto field = [&](int slot, double& v) { v = row[slot]; return ((bits[slot >> 5] >> (slot & 31)) & 1u) != 0; };
  for (int i = dd.prog_off[d]; i < dd.prog_off[d + 1]; i++) {
    const DxOp o = dd.ops[i]; double v = 0; bool p = false;
    switch (o.op) {
      case DX_CONST: v = o.x; p = true; break;
      case DX_FIELD: p = field(o.a, v); if (!p && (o.flags & 1)) { v = o.x; p = true; } break;
      case DX_NORMDISC: { double f; p = field(o.a, f); if (p) v = java_eq(f, o.z) ? 1.0 : 0.0; else if (o.flags & 1) { v = o.x; p = true; } break; }
      case DX_NORMCONT: {
        double f; p = field(o.a, f);
        if (p) { bool miss = false; v = norm_cont_dev(dd.pool + o.b, o.c, o.outliers, f, miss); p = !miss; }
        else if (o.flags & 1) { v = o.x; p = true; }
        break;
      }
      case DX_DISCRETIZE: {
        double f; p = field(o.a, f);
        if (!p) { if (o.flags & 1) { v = o.x; p = true; } break; }
        p = false;
        for (int k = 0; k < o.c && !p; k++) {
          const double* bn = dd.pool + o.b + 4 * k; int fl = (int)bn[2];
          bool lo = !(fl & 1) || ((fl & 4) ? f > bn[0] : f >= bn[0]), hi = !(fl & 2) || ((fl & 8) ? f < bn[1] : f <= bn[1]);
          if (lo && hi) { v = bn[3]; p = true; }
        }
        if (!p && (o.flags & 2)) { v = o.y; p = true; }
        break;
      }
      case DX_MAPVALUES: {
        double f; p = field(o.a, f);
        if (!p) { if (o.flags & 1) { v = o.x; p = true; } break; }
        p = false;
        for (int k = 0; k < o.c && !p; k++) if (java_eq(f, dd.pool[o.b + 2 * k])) { v = dd.pool[o.b + 2 * k + 1]; p = true; }
        if (!p && (o.flags & 2)) { v = o.y; p = true; }
        break;
      }
      default: {  // DX_APPLY: pops o.b arguments
        const int n = o.b; const double* a = sv + top - n; const bool* ap = sp + top - n; top -= n;
        if (o.a == DF_ISMISSING || o.a == DF_ISNOTMISSING) { v = (ap[0] == (o.a == DF_ISNOTMISSING)) ? 1.0 : 0.0; p = true; break; }
        if (o.a == DF_IF) {
          if (!ap[0]) { if (o.flags & 1) { v = o.x; p = true; } break; }
          if (a[0] != 0.0) { v = a[1]; p = ap[1]; } else if (n > 2) { v = a[2]; p = ap[2]; }
          if (!p && (o.flags & 2)) { v = o.y; p = true; }
          break;
        }
        if (o.a == DF_COALESCE) { v = a[0]; p = ap[0]; if (!p) { v = o.x; p = true; } break; }
        if (o.a == DF_USERCALL) {   // a[0..n-2]: the arguments (present only when mapMissingTo looks at them), a[n-1]: the body's value
          bool args_ok = true; for (int k = 0; k + 1 < n; k++) args_ok = args_ok && ap[k];
          if (!args_ok) { v = o.x; p = true; break; }
          v = a[n - 1]; p = ap[n - 1];
          if (p && o.outliers) v = (double)(float)v;
          if (!p && (o.flags & 2)) { v = o.y; p = true; }
          break;
        }
        bool all = true; for (int k = 0; k < n; k++) all = all && ap[k];
        if (!all) { if (o.flags & 1) { v = o.x; p = true; } break; }
        p = true;
        switch (o.a) {
          case DF_ADD: v = a[0] + a[1]; break; case DF_SUB: v = a[0] - a[1]; break; case DF_MUL: v = a[0] * a[1]; break; case DF_DIV: v = a[0] / a[1]; break;
          case DF_MIN: v = a[0]; for (int k = 1; k < n; k++) if (java_cmp(a[k], v) < 0) v = a[k]; break;
          case DF_MAX: v = a[0]; for (int k = 1; k < n; k++) if (java_cmp(a[k], v) > 0) v = a[k]; break;
          case DF_SUM: v = 0; for (int k = 0; k < n; k++) v += a[k]; break;
          case DF_AVG: v = 0; for (int k = 0; k < n; k++) v += a[k]; v /= (double)n; break;
          case DF_PRODUCT: v = 1; for (int k = 0; k < n; k++) v *= a[k]; break;
          case DF_POW: v = pow(a[0], a[1]); break; case DF_EXP: v = exp(a[0]); break; case DF_LN: v = log(a[0]); break;
          case DF_LOG10: v = log10(a[0]); break; case DF_SQRT: v = sqrt(a[0]); break; case DF_ABS: v = fabs(a[0]); break;
          case DF_FLOOR: v = floor(a[0]); break; case DF_CEIL: v = ceil(a[0]); break; case DF_ROUND: v = floor(a[0] + 0.5); break;
          case DF_THRESHOLD: v = a[0] > a[1] ? 1.0 : 0.0; break;
          case DF_EQ: v = java_cmp(a[0], a[1]) == 0; break; case DF_NE: v = java_cmp(a[0], a[1]) != 0; break;
          case DF_LT: v = java_cmp(a[0], a[1]) < 0; break; case DF_LE: v = java_cmp(a[0], a[1]) <= 0; break;
          case DF_GT: v = java_cmp(a[0], a[1]) > 0; break; case DF_GE: v = java_cmp(a[0], a[1]) >= 0; break;
          case DF_AND: v = 1; for (int k = 0; k < n; k++) if (a[k] == 0.0) v = 0; break;
          case DF_OR: v = 0; for (int k = 0; k < n; k++) if (a[k] != 0.0) v = 1; break;
          case DF_NOT: v = a[0] == 0.0 ? 1.0 : 0.0; break;
          default: { bool in = false; for (int k = 1; k < n; k++) if (java_cmp(a[0], a[k]) == 0) in = true; v = (in == (o.a == DF_ISIN)) ? 1.0 : 0.0; }
        }
      }
    }
    sv[top] = v; sp[top] = p; top++;
  }
  out = sv[0];
  return sp[0];
}

__global__ void __launch_bounds__(128) derive_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, double* __restrict__ X2,
                                                    uint32_t* __restrict__ P2, uint8_t* __restrict__ pre_status) {
  long row = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (row >= a.n) return;
  const int F2 = m.F, W2 = (F2 + 31) / 32;
  double* xr = X2 + row * (long)F2; uint32_t* pr = P2 + row * (long)W2;
  for (int w = 0; w < W2; w++) pr[w] = 0u;
  int status = 0;
  double rep = 0; bool has_rep = false;
  if (a.replace) { rep = a.replace[row]; has_rep = rep == rep; }
  for (int f = 0; f < m.F_in; f++) {
    double raw = a.X[row * (long)m.F_in + f];
    bool present = a.present ? ((a.present[row * (long)a.present_stride + (f >> 5)] >> (f & 31)) & 1u) : true;
    if (!present && has_rep) { raw = rep; present = true; }
    double v = 0; int st = prep_value(m.prep_in[f], m.prep_in_pool, present, raw, v);
    if (st == 2) status = ST_PREP;
    xr[f] = v; if (st == 0) pr[f >> 5] |= 1u << (f & 31);
  }
  if (status == 0) {
    for (int d = 0; d < m.derived.n_derived; d++) {
      double v; bool p = dx_run(m.derived, d, xr, pr, v); const int slot = m.F_in + d;
      if (p) {
        const int dt = m.derived.dtype[d];
        if (dt == DT_FLOAT) v = (double)(float)v;
        else if (dt == DT_INTEGER && !(isfinite(v) && v == floor(v))) { status = ST_EVAL; break; }
        xr[slot] = v; pr[slot >> 5] |= 1u << (slot & 31);
      } else xr[slot] = 0.0;
    }
  }
  pre_status[row] = (uint8_t)status;
}
__global__ void fixup_kernel(BatchArgs a, const uint8_t* __restrict__ pre_status) {
  long row = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (row < a.n && pre_status[row]) write_out(a, row, 0, pre_status[row], -1);
}

size_t derived_scratch_bytes(const DeviceModel& m, long n) {
  if (m.derived.n_derived == 0) return 0;
  size_t x = ((size_t)n * m.F * 8 + 255) & ~size_t(255), p = ((size_t)n * ((m.F + 31) / 32) * 4 + 255) & ~size_t(255);
  return x + p + (((size_t)n + 255) & ~size_t(255));
}
// launch_score for any model: with derived fields, `scratch` must hold derived_scratch_bytes(m, a.n) bytes
cudaError_t launch_score_any(const DeviceModel& m, const BatchArgs& a, void* scratch, cudaStream_t s, int* launches) {
  if (m.derived.n_derived == 0) return launch_score(m, a, s, launches, scratch);   // scratch (if any) serves the small-batch forest path
  if (a.n <= 0) return cudaSuccess;
  if (!scratch) return cudaErrorInvalidValue;
  uint8_t* base = static_cast<uint8_t*>(scratch);
  size_t x = ((size_t)a.n * m.F * 8 + 255) & ~size_t(255), p = ((size_t)a.n * ((m.F + 31) / 32) * 4 + 255) & ~size_t(255);
  double* X2 = reinterpret_cast<double*>(base); uint32_t* P2 = reinterpret_cast<uint32_t*>(base + x); uint8_t* pre = base + x + p;
  if (launches) (*launches)++;
  derive_kernel<<<(unsigned)((a.n + 127) / 128), 128, 0, s>>>(m, a, X2, P2, pre);
  cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) return e;
  BatchArgs b = a; b.X = X2; b.present = P2; b.present_stride = (m.F + 31) / 32; b.replace = nullptr;
  e = launch_score(m, b, s, launches, nullptr); if (e != cudaSuccess) return e;
  if (launches) (*launches)++;
  fixup_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(a, pre);
  return cudaGetLastError();
}

// Fills status for a whole batch without looking at the data: size mismatch (1) and unknown model (5).
__global__ void fill_status_kernel(BatchArgs a, int status) {
  long row = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (row < a.n) write_out(a, row, 0, status, -1);
}
cudaError_t launch_fill_status(const BatchArgs& a, int status, cudaStream_t s, int* launches) {
  if (a.n <= 0) return cudaSuccess;
  if (launches) (*launches)++;
  fill_status_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(a, status);
  return cudaGetLastError();
}

}  // namespace fjp
