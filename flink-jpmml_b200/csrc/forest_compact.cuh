// forest_compact.cuh -- throughput path for complete forests (included by kernels.cu after the shared device helpers).
//
// Two kernels per (sub-)batch, see model.h CompactDesc:
//   forest_rank_keys_kernel   X[n][F] doubles (+ present bits, replaceNan, input preparation) -> 16-bit rank keys, written as
//                             ready-made shared-memory images of the walk kernel's key tiles:  tile t = word[f / 2][e], e < E,
//                             each 32-bit word holding the keys of features 2p (low half) and 2p + 1 (high half) of one event.
//                             (Lane e of a warp then always reads bank e whatever feature its node tests; with two EVENTS per
//                             word, lanes 2i and 2i+1 would collide whenever they test different features.)
//   forest_compact_kernel     one CTA = one tile of E events (one per thread): the key tile arrives with one bulk copy, tree
//                             blobs (4-byte nodes) stream through the same 2-stage ring as forest_complete_kernel, TI trees
//                             interleaved per thread; per level LDS.32 node, LDS.U16 key, two LOP3, carry-chain index update.
// Results are bit-identical to forest_complete_kernel: same decisions (rank keys order values exactly like the float images),
// same leaf values, same order of the double additions.
#pragma once

struct KeyTiles {
  uint32_t* keys;        // tiles x (ceil(F/2) x E) words of two rank keys
  uint8_t* fail;         // tiles x E: 1 = input preparation refused the row (returnInvalid); NULL when that cannot happen
  long row_begin;        // first batch row of tile 0
  long ntiles;
  int E;                 // events per tile
  uint32_t tile_bytes;   // ceil(F/2) * E * 4
  int uniform_mode;      // the rounding every feature's keys want (1: up, 2: down, 0: cast) when they all agree, else -1
};

constexpr int RK_THREADS = 512;
constexpr int RK_CTAS = 2;       // resident CTAs per SM the pre-pass is built for (3 spills and is slower)
constexpr int RK_BATCH = 5;      // values a thread has in flight (E = 640: two batches per tile)

// Rank keys of RK_BATCH prepared values of one feature, the chains interleaved by hand: bucket -> first candidate threshold
// -> three forward steps (a bucket holds about half a threshold on average; the threshold array ends with 0xFFFFFFFF
// sentinels, so the scan needs no bound checks) -> a loop only for crowded buckets.
// th_addr / tab_addr: shared-memory byte addresses of the feature's threshold images and bucket table.
__device__ __forceinline__ void ck_rank_keys(uint32_t th_addr, uint32_t tab_addr, const CkFeat& cf, int mode, const double (&val)[RK_BATCH], uint32_t (&key)[RK_BATCH]) {
  uint32_t img[RK_BATCH], r[RK_BATCH], t[RK_BATCH];   // r: byte address of the candidate threshold
  const float top = (float)(cf.nb - 1u);
  const uint32_t flip = mode == 2 ? 0x80000000u : 0u;  // RD(x) = -RU(-x)
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) {
    float xf;
    if (mode == 0) xf = (float)val[k];                  // float field: JPMML's cast (round to nearest); uniform per feature
    else {
      const double sx = __hiloint2double(__double2hiint(val[k]) ^ (int)flip, __double2loint(val[k]));
      xf = __uint_as_float(__float_as_uint(__double2float_ru(sx)) ^ flip);
    }
    const uint32_t u = __float_as_uint(xf);
    img[k] = u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);   // order-preserving image (ordkey); NaN handled below
    // same bucket as model.h ck_bucket: clamp((x - lo) * scale) to [0, nb - 1], NaN -> 0
    const uint32_t b = __float2uint_rz(fminf(fmaxf((xf - cf.lo) * cf.scale, 0.0f), top));
    uint32_t first; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(first) : "r"(tab_addr + 2u * b));   // thresholds in lower buckets are all below x
    r[k] = th_addr + 4u * (xf != xf ? cf.n_thr : first);         // NaN is the greatest value (Double.compareTo)
  }
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t[k]) : "r"(r[k]));
#pragma unroll
  for (int s = 0; s < 3; s++) {
#pragma unroll
    for (int k = 0; k < RK_BATCH; k++) { r[k] += t[k] < img[k] ? 4u : 0u; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t[k]) : "r"(r[k])); }
  }
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) {
    while (t[k] < img[k]) { r[k] += 4u; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t[k]) : "r"(r[k])); }   // crowded bucket (rare)
    key[k] = ((r[k] - th_addr) >> 1) + (t[k] == img[k] ? 2u : 1u);   // 2 * rank + 1 (+1 on equality)
  }
}

// Direct table (model.h CkFeat): ONE 8-byte shared-memory load per value -- the bucket's first threshold image, how many thresholds lie
// in lower buckets, how many in this one.  A bucket with more than one threshold (rare at ~4 buckets per threshold) scans on.
__device__ __forceinline__ void ck_rank_keys_direct(uint32_t th_addr, uint32_t ent_addr, const CkFeat& cf, int mode, int uniform_mode, const double (&val)[RK_BATCH], uint32_t (&key)[RK_BATCH]) {
  uint32_t img[RK_BATCH]; uint2 en[RK_BATCH]; bool nan[RK_BATCH];
  const float top = (float)(cf.nb_e - 1u);
  const uint32_t flip = mode == 2 ? 0x80000000u : 0u;  // RD(x) = -RU(-x)
  float xfv[RK_BATCH];
  if (uniform_mode == 1) {          // the whole launch rounds up (double fields tested with <= / >): no per-lane choice of conversion
#pragma unroll
    for (int k = 0; k < RK_BATCH; k++) xfv[k] = __double2float_ru(val[k]);
  } else {
#pragma unroll
    for (int k = 0; k < RK_BATCH; k++) {
      if (mode == 0) xfv[k] = (float)val[k];
      else {
        const double sx = __hiloint2double(__double2hiint(val[k]) ^ (int)flip, __double2loint(val[k]));
        xfv[k] = __uint_as_float(__float_as_uint(__double2float_ru(sx)) ^ flip);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) {
    const float xf = xfv[k];
    const uint32_t u = __float_as_uint(xf);
    img[k] = u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);
    nan[k] = xf != xf;
    const uint32_t b = __float2uint_rz(fminf(fmaxf((xf - cf.lo) * cf.scale_e, 0.0f), top));     // model.h ck_bucket
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(en[k].x), "=r"(en[k].y) : "r"(ent_addr + 8u * b));
  }
  bool crowded[RK_BATCH], any_crowded = false;
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) {
    const uint32_t first = en[k].y & 0xFFFFu;
    key[k] = 2u * first + 1u + (img[k] > en[k].x ? 2u : (img[k] == en[k].x ? 1u : 0u));
    crowded[k] = en[k].y > 0x1FFFFu && img[k] > en[k].x;   // more than one threshold in the bucket and the first one is below the value
    any_crowded = any_crowded || crowded[k];
  }
  if (any_crowded) {   // rare (about four buckets per threshold): one branch for the whole batch
#pragma unroll
    for (int k = 0; k < RK_BATCH; k++) if (crowded[k]) {   // scan the threshold array from the bucket's second one on (sentinels end it)
      uint32_t r = th_addr + 4u * ((en[k].y & 0xFFFFu) + 1u), t;
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t) : "r"(r));
      while (t < img[k]) { r += 4u; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t) : "r"(r)); }
      key[k] = ((r - th_addr) >> 1) + (t == img[k] ? 2u : 1u);
    }
  }
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) key[k] = nan[k] ? 2u * cf.n_thr + 1u : key[k];      // NaN is the greatest value (Double.compareTo)
}

// The same for thresholds that are not floats: values are compared as doubles (64-bit images, Double.compareTo order);
// the bucket still comes from the value rounded to float, which is monotone.
__device__ __forceinline__ void ck_rank_keys_wide(uint32_t th_addr, uint32_t tab_addr, const CkFeat& cf, const double (&val)[RK_BATCH], uint32_t (&key)[RK_BATCH]) {
  uint64_t img[RK_BATCH], t[RK_BATCH]; uint32_t r[RK_BATCH];
  const float top = (float)(cf.nb - 1u);
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) {
    const uint64_t u = (uint64_t)__double_as_longlong(val[k]);
    img[k] = u ^ ((uint64_t)((long long)u >> 63) | 0x8000000000000000ull);
    const float xf = (float)val[k];
    const uint32_t b = __float2uint_rz(fminf(fmaxf((xf - cf.lo) * cf.scale, 0.0f), top));
    uint32_t first; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(first) : "r"(tab_addr + 2u * b));
    r[k] = th_addr + 8u * (val[k] != val[k] ? cf.n_thr : first);
  }
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) asm volatile("ld.shared.u64 %0, [%1];" : "=l"(t[k]) : "r"(r[k]));
#pragma unroll
  for (int s = 0; s < 3; s++) {
#pragma unroll
    for (int k = 0; k < RK_BATCH; k++) { r[k] += t[k] < img[k] ? 8u : 0u; asm volatile("ld.shared.u64 %0, [%1];" : "=l"(t[k]) : "r"(r[k])); }
  }
#pragma unroll
  for (int k = 0; k < RK_BATCH; k++) {
    while (t[k] < img[k]) { r[k] += 8u; asm volatile("ld.shared.u64 %0, [%1];" : "=l"(t[k]) : "r"(r[k])); }
    key[k] = ((r[k] - th_addr) >> 2) + (t[k] == img[k] ? 2u : 1u);
  }
}

// grid = (x, feature blocks); a CTA keeps the tables of its CK_FB features in shared memory and runs over the tiles
// x, x + gridDim.x, ...  Element i of a tile is (event, feature) = (i / 8, i % 8): a warp reads four 64-byte runs of X.
// Loads are never predicated off: the idle lanes of a short last block redo feature f0 (and store nothing), rows past the end
// of the batch redo the last row (the walk kernel never reports them).  Keys go straight to their place in the tile image,
// word[f / 2][event] half f & 1, as 2-byte stores that the L2 merges (a shared-memory staging tile with coalesced write-out
// cost two barriers per tile and was 15 % slower).
// Shared memory: the feature block's tables.
template <bool SIMPLE, bool WIDE>
__global__ void __launch_bounds__(RK_THREADS, RK_CTAS) forest_rank_keys_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const KeyTiles kt) {
  extern __shared__ __align__(16) uint8_t rk_smem[];
  const ForestDesc& fd = m.forest; const CompactDesc& cd = fd.compact;
  const int F = m.F, E = kt.E, tid = threadIdx.x;
  const int f0 = blockIdx.y * CK_FB, nf = min(CK_FB, F - f0);
  const int total = E * CK_FB;
  {
    const uint32_t t0 = cd.block_off[blockIdx.y], tbytes = cd.block_off[blockIdx.y + 1] - t0;
    const uint4* src = reinterpret_cast<const uint4*>(cd.tables + t0); uint4* dst = reinterpret_cast<uint4*>(rk_smem);
    for (uint32_t i = tid; i < tbytes / 16; i += RK_THREADS) dst[i] = __ldg(src + i);
  }
  const int j = tid & (CK_FB - 1), f = j < nf ? f0 + j : f0;   // this thread's feature never changes (RK_THREADS is a multiple of CK_FB)
  const bool owner = j < nf;
  const CkFeat cf = cd.feat[f]; const int mode = fd.feat_mode[f]; const bool float_field = m.prep[f].dtype == DT_FLOAT;
  const uint32_t th_addr = smem_u32(rk_smem) + cf.thr_off, tab_addr = smem_u32(rk_smem) + cf.tab_off, ent_addr = smem_u32(rk_smem) + cf.ent_off;
  const uint32_t pbit = 1u << (f & 31); const int pword = f0 >> 5;   // a block's 8 features share one present word
  __syncthreads();
  for (long tile = blockIdx.x; tile < kt.ntiles; tile += gridDim.x) {
    const long row0 = kt.row_begin + tile * E;
    const int last = (int)min((long)E, a.n - row0) - 1;          // last live row of this tile (>= 0)
    const double* xt = a.X + row0 * (long)F + f;
    const uint32_t* pt = a.present ? a.present + row0 * (long)a.present_stride + pword : nullptr;
    uint16_t* const gout = reinterpret_cast<uint16_t*>(kt.keys + (size_t)tile * (kt.tile_bytes / 4) + (size_t)(f >> 1) * E) + (f & 1);
    // full tiles whose element count is a whole number of thread batches (E = 640: two) need none of the clamps, and their
    // addresses advance by constants: event slot of (iteration it, value k) = tid / 8 + (it * RK_BATCH + k) * RK_THREADS / 8
    auto body = [&](auto full_c) {
      constexpr bool FULL = decltype(full_c)::value;
      constexpr int EV_STEP = RK_THREADS / 8;                               // events between two values of a thread
      const double* xp = xt + (long)(tid >> 3) * F; const long xstep = (long)EV_STEP * F;
      const uint32_t* pp = pt ? pt + (long)(tid >> 3) * a.present_stride : nullptr; const long pstep = (long)EV_STEP * a.present_stride;
      uint16_t* gp = gout + 2 * (tid >> 3);
      for (int i0 = tid; i0 < total; i0 += RK_THREADS * RK_BATCH) {
        double val[RK_BATCH]; uint32_t pw[RK_BATCH], key[RK_BATCH]; int st[RK_BATCH], e[RK_BATCH]; bool present[RK_BATCH];
#pragma unroll
        for (int k = 0; k < RK_BATCH; k++) {
          if (FULL) {
            val[k] = __ldcs(xp + k * xstep);
            pw[k] = pp ? __ldg(pp + k * pstep) : pbit;
          } else {
            e[k] = min(min(i0 + k * RK_THREADS, total - 1) >> 3, last);   // clamped: duplicates recompute a live element
            val[k] = __ldcs(xt + e[k] * F);
            pw[k] = pt ? __ldg(pt + e[k] * a.present_stride) : pbit;
          }
        }
#pragma unroll
        for (int k = 0; k < RK_BATCH; k++) present[k] = (pw[k] & pbit) != 0;
        if (a.replace) {
#pragma unroll
          for (int k = 0; k < RK_BATCH; k++) if (!present[k]) {
            const int ek = FULL ? (i0 + k * RK_THREADS) >> 3 : e[k];
            double rep = a.replace[row0 + ek]; if (rep == rep) { val[k] = rep; present[k] = true; }
          }
        }
#pragma unroll
        for (int k = 0; k < RK_BATCH; k++) {
          if (SIMPLE) { st[k] = present[k] ? 0 : 1; if (WIDE && float_field) val[k] = (double)(float)val[k]; }   // JPMML casts to the field's dataType
          else { double raw = val[k]; val[k] = 0.0; st[k] = prep_value(m.prep[f], m.prep_pool, present[k], raw, val[k]); }
        }
        if (WIDE) ck_rank_keys_wide(th_addr, tab_addr, cf, val, key);
        else if (cf.nb_e) ck_rank_keys_direct(th_addr, ent_addr, cf, mode, kt.uniform_mode, val, key);
        else ck_rank_keys(th_addr, tab_addr, cf, mode, val, key);
#pragma unroll
        for (int k = 0; k < RK_BATCH; k++) {
          const int ee = FULL ? (i0 + k * RK_THREADS) >> 3 : min(i0 + k * RK_THREADS, total - 1) >> 3;      // the unclamped event slot of the tile
          if (!SIMPLE && st[k] == 2 && ee <= last) kt.fail[tile * E + ee] = 1;
          const uint16_t kv = (uint16_t)(st[k] == 0 ? key[k] : 0x8000u);   // 0x8000 = missing
          if (owner) { if (FULL) gp[2 * EV_STEP * k] = kv; else gout[2 * ee] = kv; }
        }
        if (FULL) { xp += RK_BATCH * xstep; if (pp) pp += RK_BATCH * pstep; gp += 2 * EV_STEP * RK_BATCH; }
      }
    };
    if (last == E - 1 && total % (RK_THREADS * RK_BATCH) == 0) body(std::true_type{}); else body(std::false_type{});
  }
}

// MODE: 0 sum / average, 1 weighted sum / average, 2 majority vote, 3 modelChain (per-channel sums -> RegressionTables),
// 4 (weighted) average of the leaves' class distributions (scikit-learn style random forests) and weighted majority votes.
template <int D, int TI, int MODE, int MAXT>
__global__ void __launch_bounds__(MAXT, 1)
forest_compact_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const ForestLaunch fl, const KeyTiles kt) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ForestDesc& fd = m.forest; const CompactDesc& cd = fd.compact;
  const int E = blockDim.x, tid = threadIdx.x;
  const int depth = D ? D : fd.depth;
  const uint32_t TB = (uint32_t)cd.tree_bytes, nleaf = 1u << depth;
  const uint32_t LB = MODE == 4 ? (uint32_t)cd.leaf_bytes : 8u;   // bytes per leaf record
  const uint32_t off_leaf = 4u * nleaf - LB * nleaf;   // leaf h (h >= 2^depth) lives at 4*2^depth + LB*(h - 2^depth); wraps, 32-bit address arithmetic
  const uint32_t off_w = (4u + LB) * nleaf;
  double* dvotes = reinterpret_cast<double*>(smem + fl.off_votes);   // MODE 4: summed probabilities [label][event]
  const int G = fl.trees_per_stage;
  uint8_t* ring = smem + fl.off_ring;
  uint16_t* votes = reinterpret_cast<uint16_t*>(smem + fl.off_votes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + fl.off_bars);
  const int T = fd.n_trees, NG = (T + G - 1) / G;

  // bars: [0],[1] ring slot filled (bulk-copy bytes), [2] key tile filled, [3],[4] ring slot drained (one arrival per warp)
  if (tid == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1); mbar_init(&bars[3], E >> 5); mbar_init(&bars[4], E >> 5); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();

  uint32_t q_issue = 0, q_wait = 0, tile_parity = 0;
  auto issue = [&](int g) {  // thread 0: fill ring slot q_issue&1 with tree group g
    uint32_t slot = q_issue & 1u; int t0 = g * G; int nt = min(G, T - t0); uint32_t bytes = (uint32_t)nt * TB;
    mbar_expect_tx(&bars[slot], bytes);
    bulk_g2s(ring + (size_t)slot * G * TB, reinterpret_cast<const uint8_t*>(cd.blobs) + (size_t)t0 * TB, bytes, &bars[slot]);
    q_issue++;
  };

  for (long tile = blockIdx.x; tile < kt.ntiles; tile += gridDim.x) {
    if (tid == 0) {
      fence_proxy_async();
      mbar_expect_tx(&bars[2], kt.tile_bytes);
      const uint8_t* src = reinterpret_cast<const uint8_t*>(kt.keys) + (size_t)tile * kt.tile_bytes;
      for (uint32_t o = 0; o < kt.tile_bytes; o += 32768u) bulk_g2s(smem + o, src + o, min(32768u, kt.tile_bytes - o), &bars[2]);
      issue(0); if (NG > 1) issue(1);
    }
    const long row = kt.row_begin + tile * E + tid;
    const bool failed = kt.fail != nullptr && kt.fail[tile * E + tid] != 0;
    if (MODE == 2) for (int l = 0; l < m.n_labels; l++) { votes[l * E + tid] = 0; votes[(m.n_labels + l) * E + tid] = 0xFFFF; }
    if (MODE == 4) for (int l = 0; l < m.n_labels; l++) dvotes[l * E + tid] = 0.0;
    uint32_t seen_mask = 0; uint64_t order = 0;   // MODE 4: labels met so far, and the position (nibble) at which each was first met
    mbar_wait(&bars[2], tile_parity); tile_parity ^= 1u;

    const uint32_t kbase = smem_u32(smem) + 4u * tid;    // byte address of this event's column of key-pair words
    const uint32_t Eb = (uint32_t)E;                     // (node & 0x1FC) * E = (feat / 2) * 4E bytes; node & 2 picks the half
    double sum = 0.0, wsum = 0.0; uint32_t seen = 0;
    uint32_t cur_ch = 0;                                  // modelChain: channel of the running sum (node[0] of a tree = its channel)
    double* chv = reinterpret_cast<double*>(votes);      // modelChain: finished channel values [channel][event]
    for (int g = 0; g < NG; g++) {
      uint32_t slot = q_wait & 1u;
      mbar_wait(&bars[slot], (q_wait >> 1) & 1u); q_wait++;
      const uint32_t base = smem_u32(ring + (size_t)slot * G * TB);
      const int nt = min(G, T - g * G);
      for (int t = 0; t < nt; t += TI) {
        uint32_t idx[TI], bl[TI];
#pragma unroll
        for (int k = 0; k < TI; k++) { idx[k] = 1; bl[k] = base + (uint32_t)min(t + k, nt - 1) * TB; }
#pragma unroll
        for (int d = 0; d < (D ? D : 16); d++) {
          if (!D && d >= depth) break;
#pragma unroll
          for (int k = 0; k < TI; k++) {
            uint32_t nd, key;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(nd) : "r"(bl[k] + idx[k] * 4u));
            uint32_t ka, vv;
            asm("lop3.b32 %0, %1, %2, 2, 0xF8;" : "=r"(ka) : "r"(kbase + (nd & 0x1FCu) * Eb), "r"(nd));   // | (nd & 2): the half of the word
            asm volatile("ld.shared.u16 %0, [%1];" : "=r"(key) : "r"(ka));
            asm("lop3.b32 %0, %1, %2, 0x8000, 0xD0;" : "=r"(vv) : "r"(key), "r"(nd));   // a (node bit 15) keeps the missing bit, else key & 0x7FFF
            const uint32_t ny = (uint32_t)((int32_t)nd >> 16);
            // ny = ~K: vv + ny carries out exactly when vv > K ("go right"); idx = 2*idx + carry
            asm("{\n\t.reg .u32 t;\n\tadd.cc.u32 t, %2, %3;\n\taddc.u32 %0, %1, %1;\n\t}" : "=r"(idx[k]) : "r"(idx[k]), "r"(ny), "r"(vv));
          }
        }
#pragma unroll
        for (int k = 0; k < TI; k++) {
          if (t + k < nt) {
            if (MODE == 4) {
              const uint32_t la = bl[k] + off_leaf + idx[k] * LB; uint32_t hlo, hhi; double ww = 1.0;
              asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(hlo), "=r"(hhi) : "r"(la));
              uint64_t nib = ((uint64_t)hhi << 32) | hlo; const uint32_t cnt = hhi >> 28;
              if (fd.method == FM_WAVG_PROB || fd.method == FM_WVOTE) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(ww) : "r"(bl[k] + off_w));
              for (uint32_t i = 0; i < cnt; i++, nib >>= 4) {   // the leaf's distribution in document order
                const uint32_t lab = (uint32_t)nib & 15u; double p; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p) : "r"(la + 8u + 8u * i));
                if (!((seen_mask >> lab) & 1u)) { order |= (uint64_t)seen << (4u * lab); seen++; seen_mask |= 1u << lab; }
                dvotes[lab * E + tid] += ww * p;
              }
            } else if (MODE != 2) {
              double lv;
              asm volatile("ld.shared.f64 %0, [%1];" : "=d"(lv) : "r"(bl[k] + off_leaf + idx[k] * 8u));
              if (MODE == 1) { double w; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(w) : "r"(bl[k] + off_w)); sum += w * lv; wsum += w; }
              else if (MODE == 3) {
                uint32_t ch; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(ch) : "r"(bl[k]));
                if (ch != cur_ch) { chv[cur_ch * E + tid] = channel_value(fd.channels[cur_ch], sum); sum = 0.0; cur_ch = ch; }   // trees arrive channel by channel
                sum += lv;
              } else sum += lv;
            } else {
              uint32_t lab;
              asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lab) : "r"(bl[k] + off_leaf + idx[k] * 8u));
              uint16_t c = votes[lab * E + tid];
              if (c == 0) votes[(m.n_labels + lab) * E + tid] = (uint16_t)seen++;
              votes[lab * E + tid] = c + 1;
            }
          }
        }
      }
      // this warp is done with the slot; only the producer waits for the other warps before refilling it
      __syncwarp();
      if ((tid & 31) == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bars[3 + slot])) : "memory");
      if (tid == 0 && g + 2 < NG) { mbar_wait(&bars[3 + slot], ((q_wait - 1) >> 1) & 1u); fence_proxy_async(); issue(g + 2); }
    }
    if (row < a.n) {
      if (failed) write_out(a, row, 0, ST_PREP, -1);
      else if (MODE == 4) {   // largest summed probability; ties go to the label met first (dividing by the tree count would not move the winner)
        int best = -1; double bv = 0; uint32_t bo = 0;
        for (int l = 0; l < m.n_labels; l++) {
          if (!((seen_mask >> l) & 1u)) continue;
          const double v = dvotes[l * E + tid]; const uint32_t o = (uint32_t)(order >> (4 * l)) & 15u;
          if (best < 0) { best = l; bv = v; bo = o; continue; }
          const int c = java_cmp(v, bv);
          if (c > 0 || (c == 0 && o < bo)) { best = l; bv = v; bo = o; }
        }
        write_label(m, a, row, best, best);
      } else if (MODE == 3) {
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
    __syncthreads();  // the key tile and the votes are reused by the next tile
  }
}

// ---- sparse form: trees with explicit children (model.h CompactDesc, "sparse") --------------------------------------------
// Same tiles, keys and node word as forest_compact_kernel; a node also names its two successors, so trees keep their own
// shape (leaf-wise grown or heavily pruned trees, DAGs of compound predicates) instead of being padded to a common depth.
// Groups of up to CK_SG_MAX trees stream through the 2-slot ring; a thread walks TI trees of the group at a time until all
// of them have reached a leaf.  MODE: 0 sum / average, 1 weighted sum / average, 2 majority vote, 3 modelChain (the tree
// header's weight slot then holds the channel), 4 (weighted) sums of leaf class distributions.
template <int TI, int MODE, int MAXT>
__global__ void __launch_bounds__(MAXT, 1)
forest_sparse_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const ForestLaunch fl, const KeyTiles kt) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ForestDesc& fd = m.forest; const CompactDesc& cd = fd.compact;
  const int E = blockDim.x, tid = threadIdx.x;
  uint8_t* ring = smem + fl.off_ring;
  uint16_t* votes = reinterpret_cast<uint16_t*>(smem + fl.off_votes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + fl.off_bars);
  const int T = fd.n_trees, NG = cd.n_groups; const uint32_t slot_bytes = (uint32_t)cd.group_max_bytes;
  const uint32_t LB = MODE == 4 ? (uint32_t)cd.leaf_bytes : 8u;   // bytes per leaf record
  double* dvotes = reinterpret_cast<double*>(smem + fl.off_votes);   // MODE 4: summed probabilities [label][event]

  if (tid == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1); mbar_init(&bars[3], E >> 5); mbar_init(&bars[4], E >> 5); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }   // as in forest_compact_kernel
  __syncthreads();

  uint32_t q_issue = 0, q_wait = 0, tile_parity = 0;
  auto issue = [&](int g) {  // thread 0: fill ring slot q_issue&1 with tree group g
    const uint32_t slot = q_issue & 1u, o0 = cd.group_off[g], bytes = cd.group_off[g + 1] - o0;
    mbar_expect_tx(&bars[slot], bytes);
    bulk_g2s(ring + (size_t)slot * slot_bytes, reinterpret_cast<const uint8_t*>(cd.blobs) + o0, bytes, &bars[slot]);
    q_issue++;
  };

  for (long tile = blockIdx.x; tile < kt.ntiles; tile += gridDim.x) {
    if (tid == 0) {
      fence_proxy_async();
      mbar_expect_tx(&bars[2], kt.tile_bytes);
      const uint8_t* src = reinterpret_cast<const uint8_t*>(kt.keys) + (size_t)tile * kt.tile_bytes;
      for (uint32_t o = 0; o < kt.tile_bytes; o += 32768u) bulk_g2s(smem + o, src + o, min(32768u, kt.tile_bytes - o), &bars[2]);
      issue(0); if (NG > 1) issue(1);
    }
    const long row = kt.row_begin + tile * E + tid;
    const bool failed = kt.fail != nullptr && kt.fail[tile * E + tid] != 0;
    if (MODE == 2) for (int l = 0; l < m.n_labels; l++) { votes[l * E + tid] = 0; votes[(m.n_labels + l) * E + tid] = 0xFFFF; }
    if (MODE == 4) for (int l = 0; l < m.n_labels; l++) dvotes[l * E + tid] = 0.0;
    uint32_t seen_mask = 0; uint64_t order = 0;   // MODE 4: labels met so far, and the position (nibble) at which each was first met
    mbar_wait(&bars[2], tile_parity); tile_parity ^= 1u;

    const uint32_t kbase = smem_u32(smem) + 4u * tid, Eb = (uint32_t)E;
    double sum = 0.0, wsum = 0.0; uint32_t seen = 0;
    uint32_t cur_ch = 0; double* chv = reinterpret_cast<double*>(votes);   // modelChain: running channel, finished channel values [channel][event]
    for (int g = 0; g < NG; g++) {
      const uint32_t slot = q_wait & 1u;
      mbar_wait(&bars[slot], (q_wait >> 1) & 1u); q_wait++;
      const uint32_t base = smem_u32(ring + (size_t)slot * slot_bytes);
      uint32_t nt; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(nt) : "r"(base));
      for (uint32_t t = 0; t < nt; t += TI) {
        uint32_t tb[TI], code[TI], nn[TI];
#pragma unroll
        for (int k = 0; k < TI; k++) {
          uint32_t off; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(off) : "r"(base + 8u + 4u * min(t + k, nt - 1)));
          tb[k] = base + off;
          asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(nn[k]), "=r"(code[k]) : "r"(tb[k]));
        }
        for (;;) {   // one level of every unfinished tree per round
          bool busy = false;
#pragma unroll
          for (int k = 0; k < TI; k++) busy |= code[k] < 0x8000u;
          if (!busy) break;
#pragma unroll
          for (int k = 0; k < TI; k++) {   // straight-line: a finished chain re-reads node 0 and keeps its leaf code (no divergent branch)
            const bool done = code[k] >= 0x8000u;
            uint32_t nd, ch, ka, key, vv;
            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(nd), "=r"(ch) : "r"(tb[k] + 16u + (done ? 0u : code[k]) * 8u));
            asm("lop3.b32 %0, %1, %2, 2, 0xF8;" : "=r"(ka) : "r"(kbase + (nd & 0x1FCu) * Eb), "r"(nd));
            asm volatile("ld.shared.u16 %0, [%1];" : "=r"(key) : "r"(ka));
            asm("lop3.b32 %0, %1, %2, 0x8000, 0xD0;" : "=r"(vv) : "r"(key), "r"(nd));
            const uint32_t K = ~((uint32_t)((int32_t)nd >> 16));            // node holds ~K
            const uint32_t next = vv > K ? ch >> 16 : ch & 0xFFFFu;
            code[k] = done ? code[k] : next;
          }
        }
#pragma unroll
        for (int k = 0; k < TI; k++) {
          if (t + k < nt) {
            const uint32_t la = tb[k] + 16u + nn[k] * 8u + (code[k] & 0x7FFFu) * LB;
            if (MODE == 4) {
              uint32_t hlo, hhi; double ww = 1.0;
              asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(hlo), "=r"(hhi) : "r"(la));
              uint64_t nib = ((uint64_t)hhi << 32) | hlo; const uint32_t cnt = hhi >> 28;
              if (fd.method == FM_WAVG_PROB || fd.method == FM_WVOTE) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(ww) : "r"(tb[k] + 8u));
              for (uint32_t i = 0; i < cnt; i++, nib >>= 4) {   // the leaf's distribution in document order
                const uint32_t lab = (uint32_t)nib & 15u; double p; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p) : "r"(la + 8u + 8u * i));
                if (!((seen_mask >> lab) & 1u)) { order |= (uint64_t)seen << (4u * lab); seen++; seen_mask |= 1u << lab; }
                dvotes[lab * E + tid] += ww * p;
              }
            } else if (MODE != 2) {
              double lv; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(lv) : "r"(la));
              if (MODE == 1) { double w; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(w) : "r"(tb[k] + 8u)); sum += w * lv; wsum += w; }
              else if (MODE == 3) {
                uint32_t ch; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(ch) : "r"(tb[k] + 8u));
                if (ch != cur_ch) { chv[cur_ch * E + tid] = channel_value(fd.channels[cur_ch], sum); sum = 0.0; cur_ch = ch; }   // trees arrive channel by channel
                sum += lv;
              } else sum += lv;
            } else {
              uint32_t lab; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lab) : "r"(la));
              uint16_t c = votes[lab * E + tid];
              if (c == 0) votes[(m.n_labels + lab) * E + tid] = (uint16_t)seen++;
              votes[lab * E + tid] = c + 1;
            }
          }
        }
      }
      __syncwarp();     // this warp is done with the slot; only the producer waits for the others before refilling it
      if ((tid & 31) == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bars[3 + slot])) : "memory");
      if (tid == 0 && g + 2 < NG) { mbar_wait(&bars[3 + slot], ((q_wait - 1) >> 1) & 1u); fence_proxy_async(); issue(g + 2); }
    }
    if (row < a.n) {
      if (failed) write_out(a, row, 0, ST_PREP, -1);
      else if (MODE == 4) {
        int best = -1; double bv = 0; uint32_t bo = 0;
        for (int l = 0; l < m.n_labels; l++) {
          if (!((seen_mask >> l) & 1u)) continue;
          const double v = dvotes[l * E + tid]; const uint32_t o = (uint32_t)(order >> (4 * l)) & 15u;
          if (best < 0) { best = l; bv = v; bo = o; continue; }
          const int c = java_cmp(v, bv);
          if (c > 0 || (c == 0 && o < bo)) { best = l; bv = v; bo = o; }
        }
        write_label(m, a, row, best, best);
      } else if (MODE == 3) {
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
    __syncthreads();  // the key tile and the votes are reused by the next tile
  }
}
