// sliced_gemm.cuh -- dense layers (NeuralNetwork, J6) on the 5th-generation tensor cores: tcgen05.mma kind::i8 with the
// accumulators in tensor memory.  Compiled as sliced_gemm.cu (device helpers from device_common.cuh).  Replaces, for large
// batches, the FP64 CUDA-core tiles of neural_pair_kernel on the path the reference hands to JPMML's NeuralNetworkEvaluator
// (flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/PmmlModel.scala:159-160).
//
// Arithmetic (model.h "tensor-core form"): both operands of a layer are fixed-point images cut into S balanced base-256
// digits (S = 4: 30-bit operands); the digit products of level c = k + l are summed EXACTLY in int32 accumulators (one TMEM
// accumulator of N columns per level), and the epilogue recombines the S kept levels:
//     z_j = b_j + sigma tau 2^-(2Q) 256^(S-1) (sum_c 256^(S-1-c) L_c).
// Nothing in the tensor pipe rounds, so the only errors are the digit truncation of the operands and the dropped low-order
// pairs, bounded per layer and event by  K sigma tau SG_CS;  the bound is carried from layer to layer (activations with
// Lipschitz constant <= 1, |W| row sums) and the last layer's epilogue accepts a label only when the winner's margin exceeds
// twice the bound.  Every other event -- and every event with a missing / invalid / non-finite input -- goes on a list that
// neural_pair_kernel re-scores in FP64 in JPMML's order: labels are bit-exact by construction.
//
// Kernels per pass of <= 8 tiles per SM (a pass's digit planes go through L2 / HBM between the launches):
//   sg_slice_inputs_kernel  NeuralInputs (same FP64 code as the exact kernels) -> row scale sigma -> S digit planes
//   sg_layer_kernel<MODE>   one launch per layer; CTA = 128 events (UMMA M) x all output neurons in chunks of SG_NC (UMMA N):
//       warp 0    producer: the tile's input digit planes (K chunks of 64 inputs) and the layer's weight digit planes (one stage
//                 per (neuron chunk, K chunk)) arrive with cp.async.bulk (TMA unit) on mbarriers; both live in global memory in
//                 exactly the shared-memory image the MMA reads (K-major, no swizzle: 8 x 16 B core matrices)
//       warp 1    issues S (S + 1) / 2 digit pairs x 2 K steps of tcgen05.mma (M128 N<=64 K32) per stage from one elected lane;
//                 tcgen05.commit hands smem stages back to the producer and full accumulator sets (double-buffered in the 512
//                 TMEM columns) to the epilogue
//       warps 2-17 epilogue: tcgen05.ld S levels x 8 columns at a time, recombination, bias, activation (integer-only for
//                 rectifier / identity: one rounding shift per value), the byte transpose into the next layer's digit planes
//                 (coalesced 16-byte stores) or, for the last layer, arg-max + margin test
// Measured on the B200 (C5 MLP 64-256-256-8, profiles/): the 32-wide MMA was bound by the shared-memory reads of its 128 x 32 A
// operand (40 cycles for 16 of math), hence 64-neuron chunks; the epilogue is paced by tensor-memory reads (S x 4 B per output).
#pragma once
#include <type_traits>

namespace fjp {
namespace sg {

constexpr int THREADS = 576;                 // producer warp, MMA warp, 16 epilogue warps
constexpr int EPI_WARPS = 16;
constexpr int NSTAGE = 5;                    // weight stages in flight
constexpr int COLQ = SG_NC / 16;             // 16-column shares of a neuron chunk
constexpr int GROUPS = 16 / (4 * COLQ);      // epilogue warp groups: N = 32 -> one group of 8 warps per accumulator set; N = 64 -> all 16 warps on every chunk
constexpr uint32_t A_CHUNK = SG_S * SG_M * SG_KC;      // one K chunk of a tile, all digit planes (32 KB for S = 4)
constexpr uint32_t A_PLANE = SG_M * SG_KC;             // 8 KB
constexpr uint32_t B_STAGE = SG_S * SG_NC * SG_KC;     // 16 KB for S = 4, N = 64
constexpr uint32_t B_PLANE = SG_NC * SG_KC;            // 4 KB
constexpr int TMEM_COLS = SG_S * SG_NC * 2 <= 256 ? 256 : 512;   // two accumulator sets of SG_S x 32 columns
constexpr int MAX_KC = 4;                    // K <= 256
constexpr long PASS_TILES_PER_SM = 24;        // a pass = one launch per layer over this many tiles per SM: each launch pays its weight staging, tensor-memory
                                             // allocation and pipeline fill once (measured on C5-MLP: 8 -> 6.06e8, 16 -> 6.58e8, 24 -> 6.77e8, 32 -> 6.87e8 events/s;
                                             // the work area is 2.3 KB per row of a pass: 1.0 GB at 24)

struct Pass {
  const int8_t* a_in;    // [tile][Kp / 64][S][4][128][16]
  int8_t* a_out;         // next layer's input planes (NULL on the last layer)
  const SgRow* rows_in;  // row state of the pass's events as the previous kernel left it
  SgRow* rows_out;       // ... as this kernel leaves it (another array: the two column halves of an event read rows_in at their own pace)
  long row_begin, n_rows;
  int layer;
  int32_t* fb_count; int32_t* fb_rows;
};

// Balanced base-256 digits of four fixed-point values at once, written as byte lanes of the S plane words:
// q + 0x808080 has the bytes d_k + 128 (k < S - 1) and the top digit above them; XOR 0x80 turns a byte into the int8 digit.  The
// 4 x 4 byte transpose (value-major -> plane-major) is 8 PRMTs (S = 5: 3 more for the fifth digit).  w[0] = most significant plane.
using qint = std::conditional<SG_S == 5, long long, int>::type;   // a fixed-point value: 30 bits + sign for S = 4, 38 for S = 5
__device__ __forceinline__ qint to_q(double v) { if constexpr (SG_S == 5) return __double2ll_rn(v); else return __double2int_rn(v); }   // |v| <= 2^Q
__device__ __forceinline__ void pack4(qint q0, qint q1, qint q2, qint q3, uint32_t (&w)[SG_S][4], int wi) {
  if constexpr (SG_S == 5) {
    const long long B = 0x80808080LL;
    const long long p0 = (long long)q0 + B, p1 = q1 + B, p2 = q2 + B, p3 = q3 + B;
    const uint32_t l0 = (uint32_t)p0 ^ 0x80808080u, l1 = (uint32_t)p1 ^ 0x80808080u, l2 = (uint32_t)p2 ^ 0x80808080u, l3 = (uint32_t)p3 ^ 0x80808080u;
    const uint32_t t0 = __byte_perm(l0, l1, 0x5140), t1 = __byte_perm(l2, l3, 0x5140), t2 = __byte_perm(l0, l1, 0x7362), t3 = __byte_perm(l2, l3, 0x7362);
    w[4][wi] = __byte_perm(t0, t1, 0x5410); w[3][wi] = __byte_perm(t0, t1, 0x7632);
    w[2][wi] = __byte_perm(t2, t3, 0x5410); w[1][wi] = __byte_perm(t2, t3, 0x7632);
    const uint32_t u0 = __byte_perm((uint32_t)(p0 >> 32), (uint32_t)(p1 >> 32), 0x0040), u1 = __byte_perm((uint32_t)(p2 >> 32), (uint32_t)(p3 >> 32), 0x0040);
    w[0][wi] = __byte_perm(u0, u1, 0x5410);
  } else {   // S = 4: |q| <= 2^30, everything fits 32 bits; byte 3 of q + 0x808080 IS the (signed) top digit
    const uint32_t l0 = ((uint32_t)q0 + 0x00808080u) ^ 0x00808080u, l1 = ((uint32_t)q1 + 0x00808080u) ^ 0x00808080u;
    const uint32_t l2 = ((uint32_t)q2 + 0x00808080u) ^ 0x00808080u, l3 = ((uint32_t)q3 + 0x00808080u) ^ 0x00808080u;
    const uint32_t t0 = __byte_perm(l0, l1, 0x5140), t1 = __byte_perm(l2, l3, 0x5140), t2 = __byte_perm(l0, l1, 0x7362), t3 = __byte_perm(l2, l3, 0x7362);
    w[SG_S - 1][wi] = __byte_perm(t0, t1, 0x5410); w[SG_S - 2][wi] = __byte_perm(t0, t1, 0x7632);
    w[SG_S - 3][wi] = __byte_perm(t2, t3, 0x5410); w[0][wi] = __byte_perm(t2, t3, 0x7632);
  }
}
__device__ __forceinline__ double pow2(int e) { return __longlong_as_double((long long)(e + 1023) << 52); }   // -1022 <= e <= 1023
// smallest power-of-two exponent e with 2^e > v (v >= 0, finite); v == 0 -> 0
__device__ __forceinline__ int exp_above(double v) {
  if (!(v > 0.0)) return 0;
  const int be = (int)((__double_as_longlong(v) >> 52) & 0x7FF);
  return be == 0 ? -1021 : be - 1022;
}

// ---------------------------------------------------------------------------------------------- NeuralInputs -> digit planes
__device__ double input_value(const DeviceModel& m, const BatchArgs& a, long row, int i, int& fl) {
  const NeuralDesc& nd = m.nn;
  if (i >= nd.n_inputs) return 0.0;
  const NeuralInput in = nd.inputs[i];
  if (in.kind == NI_CONSTANT) return in.value;
  double v; const int st = in.slot < 0 ? 1 : load_field(m, a, row, in.slot, v);
  if (st == 2) { fl |= 1; return 0.0; }
  if (st == 1) { if (in.has_map_missing) return in.map_missing; fl |= 2; return 0.0; }
  if (in.kind == NI_FIELDREF) return v;
  if (in.kind == NI_NORMDISC) return java_eq(v, in.value) ? 1.0 : 0.0;
  bool miss = false; const double out = norm_cont(nd.pts + 2 * in.pts_off, in.n_pts, in.outliers, v, miss);
  if (miss) fl |= 2;
  return out;
}

// 4 lanes per event: lane q takes inputs 16 g + 4 q .. + 3 of every 16-input group g, so a quad reads 128 contiguous bytes
// of the row and writes one 16-byte unit of every plane (a warp: 8 rows = 128 contiguous bytes per plane and group).  Two
// sweeps over the inputs: row maximum, then digits.
// PLAIN: every NeuralInput is a FieldRef to the vector slot of the same index, the fields are plain doubles and the batch is
// dense (no present bits, no replaceNan) -- the rows are read straight from X with 16-byte loads.
template <bool PLAIN>
__global__ void __launch_bounds__(256) sg_slice_inputs_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const Pass ps) {
  const SgLayer& L0 = m.nn.sg.layers[0];
  const int groups = L0.Kp / 16, n_kc = L0.Kp / SG_KC, n_in = m.nn.n_inputs;
  const long t = blockIdx.x * (long)blockDim.x + threadIdx.x;
  const long r_pass = t >> 2; const int q4 = (int)(t & 3);
  const bool live = r_pass < ps.n_rows;
  const long row = ps.row_begin + (live ? r_pass : 0);
  const double* xrow = a.X + row * (long)m.F;
  const bool vec_ok = PLAIN && (m.F & 1) == 0;     // rows start 16-byte aligned
  auto load4 = [&](int g, double (&v)[4], int& fl) {
    const int i0 = g * 16 + q4 * 4;
    if (PLAIN) {
      if (vec_ok && i0 + 3 < n_in) {
        const double2 a0 = __ldg(reinterpret_cast<const double2*>(xrow + i0)), a1 = __ldg(reinterpret_cast<const double2*>(xrow + i0 + 2));
        v[0] = a0.x; v[1] = a0.y; v[2] = a1.x; v[3] = a1.y;
      } else {
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = i0 + k < n_in ? __ldg(xrow + i0 + k) : 0.0;
      }
    } else {
#pragma unroll
      for (int k = 0; k < 4; k++) v[k] = input_value(m, a, row, i0 + k, fl);
    }
  };
  int fl = 0; double amax = 0.0;
  for (int g = 0; g < groups; g++) {
    double v[4]; load4(g, v, fl);
#pragma unroll
    for (int k = 0; k < 4; k++) { amax = fmax(amax, fabs(v[k])); if (!(fabs(v[k]) <= 1e300)) fl |= 4; }
  }
  amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, 1)); amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
  fl |= __shfl_xor_sync(0xffffffffu, fl, 1); fl |= __shfl_xor_sync(0xffffffffu, fl, 2);
  if (!PLAIN) {
    if (q4 == 0 && live && !fl && row_fails_preparation(m, a, row)) fl |= 1;
    fl = __shfl_sync(0xffffffffu, fl, (threadIdx.x & 31) & ~3);
  }
  int sexp = exp_above(amax);
  if (sexp < -400 || sexp > 400) { fl |= 4; sexp = 0; }
  if (!live) return;
  const double qscale = pow2(SG_QBITS - sexp);
  const long tile = r_pass / SG_M; const int r = (int)(r_pass % SG_M);
  int8_t* tbase = ps.a_out + (size_t)tile * n_kc * A_CHUNK + (size_t)r * 16 + (size_t)q4 * 4;
  for (int g = 0; g < groups; g++) {
    double v[4]; int dummy = 0; load4(g, v, dummy);
    qint qv[4];
#pragma unroll
    for (int k = 0; k < 4; k++) qv[k] = fl ? 0 : to_q(v[k] * qscale);      // |v| < sigma: |q| <= 2^Q, off by at most 1/2
    uint32_t w[SG_S][4]; pack4(qv[0], qv[1], qv[2], qv[3], w, 0);
    int8_t* dst = tbase + (size_t)(g >> 2) * A_CHUNK + (size_t)(g & 3) * (SG_M * 16);
#pragma unroll
    for (int s = 0; s < SG_S; s++) *reinterpret_cast<uint32_t*>(dst + (size_t)s * A_PLANE) = w[s][0];
  }
  if (q4 == 0) {
    SgRow rs; rs.amax = (float)amax; rs.sexp = sexp;
    // the planes hold x to within sigma 2^-(Q+1) per input; anything irregular sends the event to the FP64 kernels
    rs.err = fl ? __longlong_as_double(0x7FF0000000000000LL) : pow2(sexp - SG_QBITS - 1);
    if ((double)rs.amax < amax) rs.amax = __float_as_uint(rs.amax) == 0x7f7fffffu ? rs.amax : __uint_as_float(__float_as_uint(rs.amax) + 1u);   // round up
    ps.rows_out[r_pass] = rs;
  }
}

// ---------------------------------------------------------------------------------------------- the layer kernel
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes) {
  // K-major, no swizzle: 8-row groups 128 B apart (SBO), the two 16-byte K halves of one MMA lbo_bytes apart (LBO); version 1
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | ((uint64_t)(128u >> 4) << 32) | ((uint64_t)1 << 46);
}
__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  const uint32_t z = 0;
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
               ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(z), "r"(z), "r"(z), "r"(z) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                 "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(taddr));
}

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
}

struct Smem { uint32_t off_b, off_bias, off_tau, off_hmax, off_bars, off_tmem, bytes; };
__host__ __device__ inline Smem smem_plan(int n_kc, int Np) {
  Smem p; uint32_t o = (uint32_t)n_kc * A_CHUNK;
  p.off_b = o; o += NSTAGE * B_STAGE;
  p.off_bias = o; o += (uint32_t)Np * 8;
  p.off_tau = o; o += (uint32_t)Np * 4; o = (o + 15u) & ~15u;
  p.off_hmax = o; o += 4u * SG_M * 8u;
  p.off_bars = o; o += 8u * (2 * MAX_KC + 2 * NSTAGE + 4);
  p.off_tmem = o; o += 16;
  p.bytes = o;
  return p;
}

// MODE: what the epilogue does with z.  Rectifier / identity layers re-slice in integer arithmetic; tanh / logistic go through
// double; the last layer keeps z for the arg-max.
enum : int { M_RELU = 0, M_IDENT = 1, M_TANH = 2, M_LOGISTIC = 3, M_LAST = 4 };
template <int MODE>
__global__ void __launch_bounds__(THREADS, 1) sg_layer_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, const Pass ps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const NeuralDesc& nd = m.nn; const SgLayer& L = nd.sg.layers[ps.layer];
  const int n_kc = L.Kp / SG_KC, n_nc = L.Np / SG_NC;
  const Smem sp = smem_plan(n_kc, L.Np);
  uint8_t* sA = smem; uint8_t* sB = smem + sp.off_b;
  double* s_bias = reinterpret_cast<double*>(smem + sp.off_bias); int* s_tau = reinterpret_cast<int*>(smem + sp.off_tau);   // (the fixed-point biases)
  double* s_hmax = reinterpret_cast<double*>(smem + sp.off_hmax);     // [2 groups x 2 column halves][128 events]: largest |h| of the tile (inf = digit overflow)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + sp.off_bars);
  uint64_t* a_full = bars; uint64_t* a_empty = bars + MAX_KC; uint64_t* b_full = bars + 2 * MAX_KC; uint64_t* b_empty = b_full + NSTAGE;
  uint64_t* t_full = b_empty + NSTAGE; uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + sp.off_tmem);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long ntiles = (ps.n_rows + SG_M - 1) / SG_M;

  // biases twice: as doubles (double-precision epilogues) and as 31-bit fixed point b_j 2^(30 - E_b), 2^E_b > max_j |b_j| (the
  // integer epilogue shifts them to each event's output scale)
  const int e_b = exp_above(L.bmax);
  for (int j = tid; j < L.Np; j += THREADS) {
    const double b = j < L.N ? nd.biases[L.b_off + j] : 0.0;
    s_bias[j] = b; s_tau[j] = __double2int_rn(b * pow2(30 - e_b));
  }
  if (tid == 0) {
    for (int i = 0; i < MAX_KC; i++) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < NSTAGE; i++) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < 2; i++) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], EPI_WARPS / GROUPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ===================================================================== producer
    if (lane == 0) {
      const int8_t* wsl = nd.sg.weights + L.w_off;
      uint32_t it_b = 0, it_t = 0;
      for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it_t++) {
        const int8_t* asrc = ps.a_in + (size_t)tile * n_kc * A_CHUNK;
        for (int nc = 0; nc < n_nc; nc++)
          for (int kc = 0; kc < n_kc; kc++, it_b++) {
            const uint32_t st = it_b % NSTAGE;
            mbar_wait(&b_empty[st], ((it_b / NSTAGE) & 1u) ^ 1u);
            mbar_expect_tx(&b_full[st], B_STAGE);
            bulk_g2s(sB + (size_t)st * B_STAGE, wsl + ((size_t)nc * n_kc + kc) * B_STAGE, B_STAGE, &b_full[st]);
            if (nc == 0) {   // the tile's own K chunk follows its first weight stage: the slot frees up later than the weight ring does
              mbar_wait(&a_empty[kc], (it_t & 1u) ^ 1u);
              mbar_expect_tx(&a_full[kc], A_CHUNK);
#pragma unroll
              for (int s = 0; s < SG_S; s++) bulk_g2s(sA + (size_t)kc * A_CHUNK + (size_t)s * A_PLANE, asrc + (size_t)kc * A_CHUNK + (size_t)s * A_PLANE, A_PLANE, &a_full[kc]);
            }
          }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    // The whole warp walks the loops (so every address below is warp-uniform and lives in uniform registers); one elected lane
    // issues the tcgen05 instructions.
    const bool leader = elect_one();
    const uint32_t idesc0 = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(SG_M >> 4) << 24);   // S32 += S8 x S8, K-major A and B, M = 128; N is set per chunk
    const uint64_t da0 = smem_desc(smem_u32(sA), SG_M * 16), db0 = smem_desc(smem_u32(sB), SG_NC * 16);
    uint32_t it_b = 0, it_t = 0, it_acc = 0;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it_t++) {
      for (int nc = 0; nc < n_nc; nc++, it_acc++) {
        const uint32_t buf = it_acc & 1u;
        mbar_wait(&t_empty[buf], ((it_acc >> 1) & 1u) ^ 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t d0 = tmem + buf * (SG_S * SG_NC);
        // a narrow last chunk (or a narrow layer) is multiplied at its own width: N = the chunk's neurons rounded up to 16
        const int n_left = L.N - nc * SG_NC;
        const uint32_t n_mma = n_left >= SG_NC ? (uint32_t)SG_NC : (uint32_t)((n_left + 15) & ~15);
        const uint32_t idesc = idesc0 | ((n_mma >> 3) << 17);
        for (int kc = 0; kc < n_kc; kc++, it_b++) {
          const uint32_t st = it_b % NSTAGE;
          if (nc == 0) mbar_wait(&a_full[kc], it_t & 1u);
          mbar_wait(&b_full[st], (it_b / NSTAGE) & 1u);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint64_t da_st = da0 + (uint64_t)(((uint32_t)kc * A_CHUNK) >> 4), db_st = db0 + (uint64_t)((st * B_STAGE) >> 4);
          const uint32_t first = kc == 0 ? 0u : 1u;
          if (leader) {
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {
#pragma unroll
              for (int k = 0; k < SG_S; k++) {
#pragma unroll
                for (int l = 0; l + k < SG_S; l++)
                  mma_i8(d0 + (uint32_t)(k + l) * SG_NC, da_st + (uint64_t)(((uint32_t)k * A_PLANE + (uint32_t)ks * 2u * (SG_M * 16)) >> 4),
                         db_st + (uint64_t)(((uint32_t)l * B_PLANE + (uint32_t)ks * 2u * (SG_NC * 16)) >> 4), idesc, (ks == 0 && k == 0) ? first : 1u);
              }
            }
            mma_commit(&b_empty[st]);                       // the weight stage can be refilled once these MMAs have read it
            if (nc == n_nc - 1) mma_commit(&a_empty[kc]);   // last use of this K chunk of the tile
          }
          __syncwarp();
        }
        if (leader) mma_commit(&t_full[buf]);
        __syncwarp();
      }
    }
  } else {
    // ===================================================================== epilogue: 16 warps, lane <-> event
    // A thread owns one event (TMEM lane) and 16 columns of a neuron chunk, taken in two rounds of 8.  N = 64: all 16 warps drain
    // every chunk (4 lane quarters x 4 column shares).  N = 32: two groups of 8 warps, one per accumulator set.
    const int ew = warp - 2, quarter = warp & 3;                                  // TMEM lanes 32 * (warp % 4) .. + 31 are this warp's
    const int group = GROUPS == 2 ? (ew >> 2) & 1 : 0, half = GROUPS == 2 ? ew >> 3 : ew >> 2;   // half: which 16 columns of the chunk
    const int r = quarter * 32 + lane;
    uint32_t it_acc = 0;
    const int next_kc = MODE == M_LAST ? 0 : nd.sg.layers[ps.layer + 1].Kp / SG_KC;
    constexpr bool INT_PATH = MODE == M_RELU || MODE == M_IDENT;
    constexpr int ACT = MODE == M_TANH ? ACT_TANH : ACT_LOGISTIC;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long r_pass = tile * SG_M + r; const bool live = r_pass < ps.n_rows;
      const SgRow rs = ps.rows_in[live ? r_pass : 0];
      const double sigma_in = pow2(rs.sexp);
      // scale of this layer's outputs, known before any of them: |h_j| <= |b_j| + sum_i |w_ji| |a_i| <= bmax + amax_W * max|a|
      const double out_bound = INT_PATH || MODE == M_LAST ? L.bmax + L.amax * (double)rs.amax : 1.0;
      int sexp_out = exp_above(out_bound * (1.0 + 1e-9));
      bool overflow = !(out_bound < 1e280) || sexp_out < -400 || sexp_out > 400;
      if (overflow) sexp_out = 0;
      const double qscale = pow2(SG_QBITS - sexp_out);          // value -> fixed point of the next layer's planes
      const double zscale = pow2(rs.sexp + L.tau_max_exp - 2 * SG_QBITS + 8 * (SG_S - 1));  // z - b = V sigma tau 2^-(2Q) 256^(S-1): the kept levels start at 256^(S-1)
      // V sigma tau 2^-(2Q) 256^(S-1) * 2^(Q - sexp_out) = V >> sh, the same shift for every neuron of the row
      int sh = sexp_out + (SG_QBITS - 8 * (SG_S - 1)) - rs.sexp - L.tau_max_exp;
      if (INT_PATH && (sh < 1 || sh > 62)) { overflow = true; sh = 1; }
      const long long rnd = 1LL << (sh - 1);
      int shb = sexp_out - e_b + 30 - SG_QBITS;                 // b_j 2^(30 - E_b) -> b_j 2^(Q - sexp_out)
      if (INT_PATH && (shb < 0 || shb > 31)) { if (shb < 0) overflow = true; shb = shb < 0 ? 0 : 31; }
      const int rndb = shb > 0 ? (1 << (shb - 1)) : 0;
      uint32_t ovf = 0; int qmax32 = 0;
      long long qmax = 0;                                        // largest |q| this thread produced (|h| = q sigma_out 2^-34)
      double zl[16];
      for (int nc = 0; nc < n_nc; nc++, it_acc++) {
        const uint32_t buf = it_acc & 1u;
        if (GROUPS == 2 && buf != (uint32_t)group) continue;     // the other group's chunk
        mbar_wait(&t_full[buf], (it_acc >> 1) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int j0 = nc * SG_NC + half * 16;
        const bool partial = j0 + 16 > L.N;                    // warp-uniform: this share reaches past the layer's last neuron
        uint32_t w[SG_S][4];
#pragma unroll
        for (int round = 0; round < 2; round++) {
          uint32_t lv[SG_S][8];
          const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + buf * (SG_S * SG_NC) + (uint32_t)half * 16u + (uint32_t)round * 8u;
#pragma unroll
          for (int c = 0; c < SG_S; c++) tmem_ld8(taddr + (uint32_t)c * SG_NC, lv[c]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          if (round == 1) {
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&t_empty[buf]);          // the accumulator set is in registers: the next chunk may overwrite it
          }
          qint qv[8];
#pragma unroll
          for (int c = 0; c < 8; c++) {
            const int cc = round * 8 + c, j = j0 + cc;
            // V = sum_c 256^(S-1-c) L_c   (|V| < 2^(8 S + 16))
            const int hi = (int)lv[0][c] * 256 + (int)lv[1][c];
            if (INT_PATH) {
              // all-integer: V is brought to the next planes' fixed point with one rounding shift (the same for the whole row), the
              // bias with another: the planes hold h to within 1.5 sigma_out 2^-Q
              long long V;
              if constexpr (SG_S == 4) V = (long long)hi * 65536 + ((long long)(int)lv[2][c] * 256 + (long long)(int)lv[3][c] + rnd);
              else { V = hi; for (int lc = 2; lc < SG_S; lc++) V = (V << 8) + (long long)(int)lv[lc][c]; V += rnd; }
              const long long vs = V >> sh;
              const int bq = (s_tau[j] + rndb) >> shb;
              if constexpr (SG_S == 4) {
                const int lo = (int)vs;
                if (!(partial && j >= L.N)) ovf |= (uint32_t)(vs >> 32) ^ (uint32_t)(lo >> 31);   // non-zero: the value left 32 bits (a violated bound)
                int q = lo + bq;
                if (MODE == M_RELU) q = max(q, 0);
                if (partial && j >= L.N) q = 0;                                 // columns the narrow MMA did not write
                qmax32 = max(qmax32, MODE == M_RELU ? q : abs(q));
                qv[c] = q;
              } else {
                long long q = vs + bq;
                if (MODE == M_RELU) q = q < 0 ? 0 : q;
                if (partial && j >= L.N) q = 0;
                const long long qa = q < 0 ? -q : q; qmax = qa > qmax ? qa : qmax;
                qv[c] = (qint)q;
              }
            } else {
              double V = (double)hi;
#pragma unroll
              for (int lc = 2; lc < SG_S; lc++) V = fma(V, 256.0, (double)(int)lv[lc][c]);
              const double z = fma(V, zscale, s_bias[j]);
              if (MODE == M_LAST) { zl[cc] = z; qv[c] = 0; }
              else {
                const long long q = (j < L.N) ? __double2ll_rn(activate(ACT, z, 0.0) * qscale) : 0;
                const long long qa = q < 0 ? -q : q; qmax = qa > qmax ? qa : qmax;
                qv[c] = (qint)q;
              }
            }
          }
          if (MODE != M_LAST) { pack4(qv[0], qv[1], qv[2], qv[3], w, round * 2); pack4(qv[4], qv[5], qv[6], qv[7], w, round * 2 + 1); }
        }
        if (MODE != M_LAST) {
          // output neuron j is input k = j of the next layer: plane s, K chunk j / 64, 16-byte column (j % 64) / 16, row r
          int8_t* dst = ps.a_out + (size_t)tile * next_kc * A_CHUNK + (size_t)(j0 >> 6) * A_CHUNK + (size_t)((j0 & 63) >> 4) * (SG_M * 16) + (size_t)r * 16;
#pragma unroll
          for (int s = 0; s < SG_S; s++) *reinterpret_cast<uint4*>(dst + (size_t)s * A_PLANE) = make_uint4(w[s][0], w[s][1], w[s][2], w[s][3]);
        }
      }
      // error bound of this layer's outputs for the event (model.h): propagated input error + slicing / dropped pairs + FP64 rounding
      // of both this evaluation and the reference's sequential sum
      const double mag = L.bmax + L.amax * (double)rs.amax;
      const double e_out = L.amax * rs.err + (double)L.K * sigma_in * pow2(L.tau_max_exp) * SG_CS + (double)(L.K + 8) * 2.220446049250313e-16 * mag;
      if (MODE != M_LAST) {
        // pad columns of the next layer's planes (Np of this layer < Kp of the next) are zero digits
        if (half == 0 && group == 0) {
          for (int j = L.Np; j < next_kc * SG_KC; j += 16) {
            int8_t* dst = ps.a_out + (size_t)tile * next_kc * A_CHUNK + (size_t)(j >> 6) * A_CHUNK + (size_t)((j & 63) >> 4) * (SG_M * 16) + (size_t)r * 16;
#pragma unroll
            for (int s = 0; s < SG_S; s++) *reinterpret_cast<uint4*>(dst + (size_t)s * A_PLANE) = make_uint4(0, 0, 0, 0);
          }
        }
        // the four threads of an event (2 groups x 2 column halves) exchange their largest |h|: the next layer scales ITS outputs
        // from the actual maximum of its inputs, so the a-priori bounds do not compound from layer to layer
        if (ovf) overflow = true;
        if ((long long)qmax32 > qmax) qmax = qmax32;
        if (qmax > (1LL << SG_QBITS)) overflow = true;
        s_hmax[(group * COLQ + half) * SG_M + r] = overflow ? __longlong_as_double(0x7FF0000000000000LL) : (double)qmax * pow2(sexp_out - SG_QBITS);
        asm volatile("bar.sync 1, %0;" ::"r"(EPI_WARPS * 32) : "memory");
        if (half == 0 && group == 0 && live) {
          const double hm = fmax(fmax(s_hmax[r], s_hmax[SG_M + r]), fmax(s_hmax[2 * SG_M + r], s_hmax[3 * SG_M + r]));
          SgRow o; o.sexp = sexp_out;
          o.amax = (float)fmin(hm * (1.0 + 1e-6), 3.0e38);      // largest |a_i| the planes hold
          // an activation is evaluated on z instead of the exact z*: at most e_out away (Lipschitz constant <= 1) plus its own
          // rounding; the planes then hold it to within 1.5 sigma_out 2^-Q
          o.err = !(hm <= out_bound * (1.0 + 1e-9) + pow2(sexp_out - SG_QBITS + 1)) ? __longlong_as_double(0x7FF0000000000000LL)
                                                                         : e_out + 4.0 * 2.220446049250313e-16 * fmax(1.0, out_bound) + 1.5 * pow2(sexp_out - SG_QBITS);
          ps.rows_out[r_pass] = o;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(EPI_WARPS * 32) : "memory");   // s_hmax is rewritten by the next tile
      }
      // the last layer has one neuron chunk per tile: the group whose accumulator set it landed in finishes the event
      if (MODE == M_LAST && half == 0 && group == (GROUPS == 2 ? (int)((it_acc - 1) & 1u) : 0)) {
        // first maximal NeuralOutput, accepted when it wins by more than twice the error bound (softmax is monotone; |z| < 600
        // keeps exp() finite and non-zero so the order of the probabilities is the order of the z)
        bool ok = live && rs.err < 1e300 && e_out < 1e300;
        int best = -1; double zb = 0, z2 = -1e308, zabs = 0;
        for (int o = 0; o < nd.n_outputs; o++) {
          const double z = zl[nd.out_neuron[o]]; zabs = fmax(zabs, fabs(z));
          if (best < 0) { best = o; zb = z; }
          else if (z > zb) { z2 = zb; zb = z; best = o; }
          else z2 = fmax(z2, z);
        }
        ok = ok && zabs < 600.0 && (nd.n_outputs == 1 || zb - z2 > 2.0 * e_out + 1e-12 * fmax(1.0, zabs));
        const long row = ps.row_begin + r_pass;
        if (ok) write_label(m, a, row, best, best);
        const unsigned need = __ballot_sync(0xffffffffu, live && !ok), alive = __ballot_sync(0xffffffffu, live);
        if (a.stats && lane == 0) { atomicAdd(a.stats, (unsigned long long)__popc(alive)); if (need) atomicAdd(a.stats + 1, (unsigned long long)__popc(need)); }
        if (need) {
          int base = 0;
          if (lane == 0) base = atomicAdd(ps.fb_count, __popc(need));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (live && !ok) ps.fb_rows[base + __popc(need & ((1u << lane) - 1u))] = (int32_t)row;
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
}

// ---------------------------------------------------------------------------------------------- fused single dense layer (C3)
// RegressionModel with softmax over plain dense rows: y_t = b_t + sum_f w_tf x_f for F up to thousands of fields, a handful
// of tables.  The contraction is tiny for the tensor cores; what has to be fast is getting 8 F bytes per event from HBM into
// int8 digit planes, so the slicing is fused into the kernel and the rows are read ONCE:
//   warps 0-15  producers: a quad per event reads 64 fields (one K chunk, 512 contiguous bytes) with 16-byte streaming loads,
//               software-pipelined one chunk ahead; every (event, chunk) gets its OWN power-of-two scale (largest |x| of those
//               64 values), so no pass over the whole row is needed; the four digit planes go to shared memory in the MMA's
//               layout (conflict-free 4-byte stores), then fence.proxy.async + one mbarrier arrive per warp;
//   warp 16     TMA: the weight digit planes of the chunk (cp.async.bulk);
//   warp 17     MMA issuer: S (S + 1) / 2 pairs x 2 K steps of tcgen05.mma.kind::i8 per chunk into a FRESH accumulator set (the chunks
//               have different scales, so they are not accumulated in tensor memory);
//   warps 18-21 epilogue: per chunk tcgen05.ld of the S levels, recombination to one int64 per table, scaled by the event's
//               chunk scale and added into double registers; after the last chunk: intercepts, arg-max, margin test.
// Error bound per event: sum over chunks of 64 sigma_c tau SG_CS (+ the FP64 rounding of both orders); labels inside it go to the
// FP64 tile kernel through the re-scoring list, like everywhere else.
namespace dense {
constexpr int PROD_WARPS = 16, TMA_WARP = 16, MMA_WARP = 17, EPI_WARP0 = 18, EPI_WARPS = 4;
constexpr int THREADS = 32 * (PROD_WARPS + 2 + EPI_WARPS);
constexpr int A_SLOTS = 3, B_SLOTS = 4, SEXP_SLOTS = 8;
constexpr int MAXT = 32;                    // MMA width (N = 32)
constexpr int MAXZ = 24;                    // tables the epilogue keeps running sums for
struct Smem { uint32_t off_b, off_sexp, off_bars, off_tmem, bytes; };
__host__ __device__ inline Smem smem_plan() {
  Smem p; uint32_t o = A_SLOTS * A_CHUNK;
  p.off_b = o; o += B_SLOTS * B_STAGE;
  p.off_sexp = o; o += SEXP_SLOTS * SG_M * 4;
  p.off_bars = o; o += 8u * (2 * A_SLOTS + 2 * B_SLOTS + 4);
  p.off_tmem = o; o += 16; p.bytes = o;
  return p;
}

__global__ void __launch_bounds__(THREADS, 1) sg_dense_kernel(const __grid_constant__ DeviceModel m, const BatchArgs a, int32_t* fb_count, int32_t* fb_rows) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const RegressionDesc& rd = m.reg; const SgLayer& L = rd.sg.layers[0];
  const int F = m.F, n_kc = L.Kp / SG_KC, C = rd.n_tables;
  const Smem sp = smem_plan();
  uint8_t* sA = smem; uint8_t* sB = smem + sp.off_b;
  int* s_sexp = reinterpret_cast<int*>(smem + sp.off_sexp);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + sp.off_bars);
  uint64_t* a_full = bars; uint64_t* a_empty = bars + A_SLOTS; uint64_t* b_full = bars + 2 * A_SLOTS; uint64_t* b_empty = b_full + B_SLOTS;
  uint64_t* t_full = b_empty + B_SLOTS; uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + sp.off_tmem);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long ntiles = (a.n + SG_M - 1) / SG_M;
  if (tid == 0) {
    for (int i = 0; i < A_SLOTS; i++) { mbar_init(&a_full[i], PROD_WARPS); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < B_SLOTS; i++) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < 2; i++) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == MMA_WARP) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  constexpr uint32_t ACC_COLS = SG_S * MAXT;          // columns of one accumulator set

  if (warp < PROD_WARPS) {
    // ===================================================================== producers
    const int q4 = lane & 3, r = warp * 8 + (lane >> 2);
    uint32_t it = 0;                                   // chunks produced so far (all tiles)
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long row = tile * SG_M + r; const bool live = row < a.n;
      const double* xrow = a.X + (live ? row : 0) * (long)F + q4 * 4;
      double cur[16], nxt[16];
      auto load = [&](int kc, double (&v)[16]) {
#pragma unroll
        for (int g = 0; g < 4; g++) {
          const int i0 = kc * SG_KC + g * 16 + q4 * 4;
          if (live && i0 + 3 < F) {
            const double2 u0 = __ldcs(reinterpret_cast<const double2*>(xrow + kc * SG_KC + g * 16)), u1 = __ldcs(reinterpret_cast<const double2*>(xrow + kc * SG_KC + g * 16 + 2));
            v[4 * g] = u0.x; v[4 * g + 1] = u0.y; v[4 * g + 2] = u1.x; v[4 * g + 3] = u1.y;
          } else {
#pragma unroll
            for (int k = 0; k < 4; k++) v[4 * g + k] = (live && i0 + k < F) ? __ldcs(xrow + kc * SG_KC + g * 16 + k) : 0.0;
          }
        }
      };
      load(0, cur);
      for (int kc = 0; kc < n_kc; kc++, it++) {
        if (kc + 1 < n_kc) load(kc + 1, nxt);
        // the chunk's scale for this event: largest |x| of its 64 values (high words compare like the doubles)
        int hm = 0;
#pragma unroll
        for (int k = 0; k < 16; k++) hm = max(hm, __double2hiint(cur[k]) & 0x7fffffff);
        hm = max(hm, __shfl_xor_sync(0xffffffffu, hm, 1)); hm = max(hm, __shfl_xor_sync(0xffffffffu, hm, 2));
        int sexp = (hm >> 20) - 1022;                  // 2^sexp > |x| (zero / denormal chunk: far below, harmless)
        const bool wild = hm >= 0x7ff00000 || sexp > 400;   // NaN / Inf / huge: the event is re-scored
        if (sexp < -400) sexp = -400;
        if (wild) sexp = 0;
        const double qscale = pow2(SG_QBITS - sexp);
        const uint32_t slot = it % A_SLOTS;
        mbar_wait(&a_empty[slot], ((it / A_SLOTS) & 1u) ^ 1u);
        uint8_t* dst = sA + (size_t)slot * A_CHUNK + (size_t)r * 16 + (size_t)q4 * 4;
#pragma unroll
        for (int g = 0; g < 4; g++) {
          qint qv[4];
#pragma unroll
          for (int k = 0; k < 4; k++) qv[k] = wild ? 0 : to_q(cur[4 * g + k] * qscale);
          uint32_t w[SG_S][4]; pack4(qv[0], qv[1], qv[2], qv[3], w, 0);
#pragma unroll
          for (int s = 0; s < SG_S; s++) *reinterpret_cast<uint32_t*>(dst + (size_t)s * A_PLANE + (size_t)g * (SG_M * 16)) = w[s][0];
        }
        if (q4 == 0) s_sexp[(it % SEXP_SLOTS) * SG_M + r] = wild ? 0x7fffffff : sexp;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full[slot]);
#pragma unroll
        for (int k = 0; k < 16; k++) cur[k] = nxt[k];
      }
    }
  } else if (warp == TMA_WARP) {
    // ===================================================================== weight planes
    if (lane == 0) {
      const int8_t* wsl = rd.sg.weights + L.w_off;
      uint32_t it = 0;
      for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
        for (int kc = 0; kc < n_kc; kc++, it++) {
          const uint32_t st = it % B_SLOTS;
          mbar_wait(&b_empty[st], ((it / B_SLOTS) & 1u) ^ 1u);
          mbar_expect_tx(&b_full[st], B_STAGE);
          bulk_g2s(sB + (size_t)st * B_STAGE, wsl + (size_t)kc * B_STAGE, B_STAGE, &b_full[st]);
        }
    }
  } else if (warp == MMA_WARP) {
    // ===================================================================== MMA issuer
    const bool leader = elect_one();
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(MAXT >> 3) << 17) | ((uint32_t)(SG_M >> 4) << 24);
    const uint64_t da0 = smem_desc(smem_u32(sA), SG_M * 16), db0 = smem_desc(smem_u32(sB), SG_NC * 16);
    uint32_t it = 0;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
      for (int kc = 0; kc < n_kc; kc++, it++) {
        const uint32_t sa = it % A_SLOTS, sb = it % B_SLOTS, buf = it & 1u;
        mbar_wait(&t_empty[buf], ((it >> 1) & 1u) ^ 1u);
        mbar_wait(&a_full[sa], (it / A_SLOTS) & 1u);
        mbar_wait(&b_full[sb], (it / B_SLOTS) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint64_t da_st = da0 + (uint64_t)((sa * A_CHUNK) >> 4), db_st = db0 + (uint64_t)((sb * B_STAGE) >> 4);
        const uint32_t d0 = tmem + buf * ACC_COLS;
        if (leader) {
#pragma unroll
          for (int ks = 0; ks < 2; ks++) {
#pragma unroll
            for (int k = 0; k < SG_S; k++) {
#pragma unroll
              for (int l = 0; l + k < SG_S; l++)
                mma_i8(d0 + (uint32_t)(k + l) * MAXT, da_st + (uint64_t)(((uint32_t)k * A_PLANE + (uint32_t)ks * 2u * (SG_M * 16)) >> 4),
                       db_st + (uint64_t)(((uint32_t)l * B_PLANE + (uint32_t)ks * 2u * (SG_NC * 16)) >> 4), idesc, (ks == 0 && k == 0) ? 0u : 1u);
            }
          }
          mma_commit(&a_empty[sa]); mma_commit(&b_empty[sb]); mma_commit(&t_full[buf]);
        }
        __syncwarp();
      }
  } else {
    // ===================================================================== epilogue: 4 warps, lane <-> event
    const int quarter = warp & 3, r = quarter * 32 + lane;
    uint32_t it = 0;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long row = tile * SG_M + r; const bool live = row < a.n;
      double z[MAXZ];
#pragma unroll
      for (int j = 0; j < MAXZ; j++) z[j] = 0.0;
      double sig_sum = 0.0, sig_max = 0.0; bool wild = false;
      for (int kc = 0; kc < n_kc; kc++, it++) {
        const uint32_t buf = it & 1u;
        mbar_wait(&t_full[buf], (it >> 1) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int se = s_sexp[(it % SEXP_SLOTS) * SG_M + r];
        if (se == 0x7fffffff) wild = true;
        const double sigma = pow2(se == 0x7fffffff ? 0 : se);
        const double scale = sigma * pow2(L.tau_max_exp - 2 * SG_QBITS + 8 * (SG_S - 1));
        sig_sum += sigma; sig_max = fmax(sig_max, sigma);
#pragma unroll
        for (int round = 0; round < MAXZ / 8; round++) {
          uint32_t lv[SG_S][8];
          const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + buf * ACC_COLS + (uint32_t)round * 8u;
#pragma unroll
          for (int c = 0; c < SG_S; c++) tmem_ld8(taddr + (uint32_t)c * MAXT, lv[c]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          if (round == MAXZ / 8 - 1) {
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&t_empty[buf]);
          }
#pragma unroll
          for (int c = 0; c < 8; c++) {
            long long V = (long long)((int)lv[0][c] * 256 + (int)lv[1][c]);
#pragma unroll
            for (int lc = 2; lc < SG_S; lc++) V = (V << 8) + (long long)(int)lv[lc][c];
            z[round * 8 + c] = fma((double)V, scale, z[round * 8 + c]);      // exact product, one rounding per chunk
          }
        }
      }
      // intercepts, first maximal table, margin against the bound
      const double e = (double)SG_KC * sig_sum * pow2(L.tau_max_exp) * SG_CS + (double)(F + 64) * 2.220446049250313e-16 * (L.bmax + L.amax * sig_max);
      int best = 0; double zb = 0, z2 = -1e308, zabs = 0;
#pragma unroll
      for (int j = 0; j < MAXZ; j++) {
        if (j < C) {
          const double zz = z[j] + rd.intercept[j]; zabs = fmax(zabs, fabs(zz));
          if (j == 0) { zb = zz; } else if (zz > zb) { z2 = zb; zb = zz; best = j; } else z2 = fmax(z2, zz);
        }
      }
      const bool ok = live && !wild && zabs < 600.0 && (C == 1 || zb - z2 > 2.0 * e + 1e-12 * fmax(1.0, zabs));
      if (ok) write_label(m, a, row, best, best);
      const unsigned need = __ballot_sync(0xffffffffu, live && !ok), alive = __ballot_sync(0xffffffffu, live);
      if (a.stats && lane == 0) { atomicAdd(a.stats, (unsigned long long)__popc(alive)); if (need) atomicAdd(a.stats + 1, (unsigned long long)__popc(need)); }
      if (need) {
        int base = 0;
        if (lane == 0) base = atomicAdd(fb_count, __popc(need));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (live && !ok) fb_rows[base + __popc(need & ((1u << lane) - 1u))] = (int32_t)row;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256) : "memory");
}
}  // namespace dense

}  // namespace sg
}  // namespace fjp
