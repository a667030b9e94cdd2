// model.h -- compiled (flattened) model representation shared by the PMML compiler, the engine and the kernels.
//
// What the reference keeps as a JPMML object graph behind `Evaluator` (api/Evaluator.scala:49-91) becomes a set of
// flat arrays that live in HBM for as long as the model is in the engine's table.  Device-visible structs are PODs.
#pragma once
#include <climits>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

namespace fjp {

// ---- per-event status byte (SURVEY.md §8b "Error convention"); 0 <=> Prediction(Score(d)), !=0 <=> EmptyScore
enum : uint8_t {
  ST_OK = 0,        // Score
  ST_SIZE = 1,      // InputValidationException  (api/PmmlModel.scala:127-134)
  ST_PREP = 2,      // InputPreparationException (api/PmmlModel.scala:143-152, api/pipeline/Pipeline.scala:49-54)
  ST_EVAL = 3,      // JPMML EvaluationException / null prediction
  ST_EXTRACT = 4,   // JPMMLExtractionException / non-Double target (api/pipeline/Pipeline.scala:94-98)
  ST_NOMODEL = 5,   // EmptyEvaluatorException (api/Evaluator.scala:69-77)
};

enum : uint8_t { DT_STRING = 0, DT_INTEGER = 1, DT_FLOAT = 2, DT_DOUBLE = 3, DT_BOOLEAN = 4, DT_OTHER = 5 };
enum : uint8_t { IVT_RETURN_INVALID = 0, IVT_AS_IS = 1, IVT_AS_MISSING = 2 };
enum : uint8_t { OT_AS_IS = 0, OT_AS_MISSING = 1, OT_AS_EXTREME = 2 };

// One entry per active MiningField, in vector-slot order (api/converter/VectorConverter.scala:68-73,97-98).
// Encodes JPMML's FieldValueUtil.prepareInputValue for a java.lang.Double input.
struct PrepField {
  double repl, low, high;
  uint32_t pool_off;  // doubles in prep_pool: n_intervals*(lo,hi,flags) then valid, invalid, missing value lists
  uint16_t n_intervals, n_valid, n_invalid, n_missing;
  uint8_t dtype, continuous, ivt, outlier;
  uint8_t has_repl;   // 1: missingValueReplacement present; 2: present but unusable -> preparation failure
  uint8_t has_low, has_high;
  uint8_t has_valid_space;  // DataField lists property="valid" values: anything else is invalid
};

struct TargetXform {  // <Targets><Target rescaleFactor rescaleConstant min max castInteger>
  double factor, constant, mn, mx;
  int has_min, has_max, cast;  // cast: 0 none, 1 round, 2 ceiling, 3 floor
};

enum ModelKind : int { MK_CLUSTERING = 0, MK_FOREST = 1, MK_REGRESSION = 2, MK_NEURAL = 3, MK_GENERAL_REGRESSION = 4 };

// ---- tree ensembles -----------------------------------------------------------------------------
enum : int { FOREST_COMPLETE = 0, FOREST_GENERAL = 1 };
enum : int { ENC_FLOAT = 0, ENC_RANK = 1 };
enum : int { FM_SINGLE = 0, FM_SUM = 1, FM_AVERAGE = 2, FM_WSUM = 3, FM_WAVERAGE = 4, FM_VOTE = 5, FM_WVOTE = 6,
             FM_AVG_PROB = 7, FM_WAVG_PROB = 8,
             FM_CHAIN = 9,    // modelChain: per-channel tree sums feed RegressionTables (DeviceModel::reg, slot = channel)
             FM_SELECT_FIRST = 10, FM_MEDIAN = 11, FM_MAX = 12 };  // general form only (forest_general_kernel)
static constexpr int MAX_MEDIAN_SEGMENTS = 64;
static constexpr int MAX_CHANNELS = 16;
struct ChannelSpec { TargetXform tx; int32_t n_trees; int32_t average; };  // one regression segment of a modelChain  // classification average / weightedAverage of the leaves' distributions

// General decision node: key = key(x[feat]); key == 0 (missing) -> on_missing, key <= k -> on_true, else on_false.
// Successors >= 0 are node indices, < 0 are ~leaf index, LEAF_NULL is "null prediction", LEAF_SKIP is "the Segment's
// predicate is not TRUE: the segment takes no part in the aggregation" (general form only).
struct GNode {
  uint32_t feat;
  uint32_t k;
  int32_t on_true, on_false, on_missing;
};
static constexpr int32_t LEAF_NULL = INT32_MIN;  // successor meaning "null prediction"
static constexpr int32_t LEAF_SKIP = INT32_MIN + 1;

#ifdef __CUDACC__
#define FJP_HD __host__ __device__
#else
#define FJP_HD
#endif

// Compact form of a complete forest for the throughput path (large batches): 16-bit rank keys and 4-byte nodes halve the
// shared-memory footprint of a key tile (twice the resident warps) and the bank conflicts of the node gather.
//   key(x[f])   0x8000 = missing, else 2*lo + 1 (+1 when x equals threshold lo), lo = number of distinct thresholds of f below
//               x (the scheme of ENC_RANK, < 0x7FFF); computed by forest_rank_keys_kernel from the float image of x: a bucket
//               table over the value range gives the few candidate thresholds, a short scan among them the exact rank.
//               A key tile in shared memory is word[f / 2][event]: the keys of features 2p and 2p + 1 share a 32-bit word.
//   tree blob   u32 node[2^D] in heap order (node[0] unused), then 2^D leaves (8 B), then the segment weight (double).
//               node = feat << 1 | a << 15 | (~K & 0xFFFF) << 16;  right iff (a ? key : key & 0x7FFF) > K, evaluated as the
//               carry of that value + sign_extend16(node >> 16).  a = 1 sends a missing value right.  Without a threshold:
//               ~K = 0 (never right) or K = 0x7FFF with a = 1 (right only when missing).
constexpr int CK_FB = 8;             // features per block of the key pre-pass
struct CkFeat {   // byte offsets inside the feature block's table area
  uint32_t thr_off, n_thr, tab_off, nb; float lo, scale;
  // direct table (float thresholds only): nb_e entries {u32 image of the bucket's first threshold or all-ones, u16 thresholds in lower
  // buckets, u16 thresholds in this bucket}; with about four buckets per threshold one 8-byte load decides almost every value
  uint32_t ent_off, nb_e; float scale_e; uint32_t pad;
};
//   sparse      trees that do not fit one common depth (leaf-wise grown, heavily pruned, DAGs of compound predicates) keep
//               explicit children: tree = {u32 n_nodes, u32 root, double weight, u64 node[n_nodes], 8-byte leaf[...]},
//               node = the 32-bit word above | (left | right << 16) << 32; a child / root code >= 0x8000 is leaf (code & 0x7FFF).
//               Trees are streamed in groups: {u32 n_trees, u32 0, u32 byte offset of tree k [CK_SG_MAX]} followed by the trees.
constexpr int CK_SG_MAX = 16;        // trees per group of the sparse form
constexpr int CK_SG_HDR = 8 + 4 * CK_SG_MAX;   // 72 -> trees start at 80
struct CompactDesc {
  int enabled, wide, tree_bytes, n_blocks, block_smem;   // wide: 64-bit threshold images (thresholds that are not floats); block_smem: largest table area of one feature block
  int sparse, n_groups, group_max_bytes;                 // sparse form: blobs = the groups back to back
  int uniform_mode;            // the rounding mode (ForestDesc::feat_mode) all features share, or -1
  int leaf_bytes;              // complete form: 8, or 8 + 8 * (most entries of a leaf's class distribution) for ensembles that average them:
                               // {u64: label nibbles in document order, entry count in the top nibble; double probability[...]}
  const uint32_t* group_off;   // n_groups + 1 byte offsets into blobs (multiples of 16)
  const uint32_t* blobs;       // n_trees x tree_bytes (complete form) / the groups (sparse form)
  const CkFeat* feat;          // F
  const uint8_t* tables;       // per feature block: sorted distinct threshold images (u32, or u64 when wide; all-ones-terminated) and bucket tables (u16 x nb)
  const uint32_t* block_off;   // n_blocks + 1 byte offsets into tables (multiples of 16)
};
FJP_HD inline int compact_tree_bytes(int depth) { return (int)(((12u << depth) + 8u + 15u) & ~15u); }
// bucket of a (finite or infinite, not NaN) float value; monotone non-decreasing in x, the same arithmetic on host and device
FJP_HD inline uint32_t ck_bucket(float x, float lo, float scale, uint32_t nb) {
  float t = (x - lo) * scale;
  if (!(t > 0.0f)) return 0u;
  if (t >= (float)nb) return nb - 1u;
  return (uint32_t)t;
}

struct ForestDesc {
  int n_trees, depth, format, enc, method, classify;
  CompactDesc compact;
  int n_channels; const ChannelSpec* channels; const int32_t* tree_channel;   // FM_CHAIN only
  // FOREST_COMPLETE: n_trees blobs of 2^(depth+1) 8-byte words: [0] = segment weight (double), [1 .. 2^d-1] = nodes
  // in heap order {lo: feat | a<<31, hi: ~ka}, [2^d .. 2^(d+1)-1] = leaves (double value, or label index as u64);
  // the walk goes right iff key(x[feat]) - a > ka (unsigned; key 0 = missing, a = 1 sends missing right)
  const uint64_t* blobs;
  const int32_t* leaf_node_id;  // complete: n_trees * 2^depth PMML node ids (aux output)
  // FOREST_GENERAL
  const GNode* gnodes; const int32_t* troot; const double* gleaf_value; const int32_t* gleaf_label; const double* tweight;
  const int32_t* gleaf_node_id;
  // classification leaves: ScoreDistribution as CSR (label index, probability) in document order; a leaf without one
  // holds the single entry (its label, 1.0)
  const int32_t* gdist_off; const int32_t* gdist_label; const double* gdist_prob;
  // ENC_FLOAT: per feature 0 = value is a float already, 1 = round up to float, 2 = round down
  const uint8_t* feat_mode;
  // ENC_RANK: per active feature sorted distinct thresholds
  const int32_t* rank_off; const double* rank_thr;
};

// ---- clustering ----------------------------------------------------------------------------------
enum : int { CM_EUCLIDEAN = 0, CM_SQEUCLIDEAN = 1, CM_CHEBYCHEV = 2, CM_CITYBLOCK = 3, CM_MINKOWSKI = 4 };
enum : int { CF_ABSDIFF = 0, CF_GAUSSSIM = 1, CF_DELTA = 2, CF_EQUAL = 3 };
struct ClusterDesc {
  int k, nf, measure, fast;  // fast: absDiff everywhere, weights 1, q all 1, (squared)Euclidean
  int slots_identity;        // ClusteringField i reads vector slot i and there are no other slots
  double p;
  const int32_t* slot;      // nf: active-field index of each ClusteringField
  const double* weight;     // nf
  const double* q;          // nf missing-value weights
  const int32_t* compare;   // nf
  const double* sim_scale;  // nf
  const double* centers;    // k x nf
  // FMA form (fast models): |x - c_t|^2 - |x|^2 = fast_icpt[t] + sum_f fast_wt[f][t] x_f, fast_wt = -2 c transposed [nf][KP] (KP as for
  // the regression tables: k rounded up to 8 / 10 / 12 per thread), fast_icpt = |c_t|^2; cmax = largest |centre coordinate|
  const double* fast_wt; const double* fast_icpt; double cmax;
  const double* fast_csq;     // c^2 transposed [nf + 1][KP], the last row zero: what a missing field takes back out of |c_t|^2
};

static constexpr int MAX_LAYERS = 8;
// ---- tensor-core form of a dense layer (tcgen05.mma kind::i8; sliced_gemm.cuh) -----------------------------------------
// y_j = b_j + sum_i w_ji a_i is evaluated in EXACT integer arithmetic on fixed-point images of the operands:
//   a_i = sigma * qa_i * 2^-Q, sigma = a power of two > max_i |a_i| of the event's row      (|qa_i| <= 2^Q, Q = 8 S - 2)
//   w_ji = tau * qw_ji * 2^-Q,  tau = a power of two > max_ji |w_ji| of the layer
//   q = sum_{k<S} d_k 256^k with balanced digits d_k in [-128, 127] (the top one in [-65, 65]): S int8 "slices" per operand --
//   the bytes of q + 0x808080, the low S - 1 of them XOR 0x80.
// The digit products of level c = k + l accumulate in their own int32 TMEM accumulator (|.| < 2^25 for K <= 256); the pairs of
// the S most significant levels are kept (S (S + 1) / 2 MMAs per K step) and the epilogue recombines them.  What is lost is
// bounded per element by SG_CS * sigma * tau (digit truncation of both operands: 2 x 2^-(Q+1); dropped pairs: 2 (S - 1) x
// 2^-(Q+1)), so every value carries a rigorous error bound; labels whose top-2 margin is inside it are re-scored by the FP64
// kernels in JPMML's order.  S = 4 (30-bit operands, 10 MMAs per K step) keeps the re-scored share of the C5 MLP near 1e-3;
// S = 5 brings it to 1e-5 for 50 % more tensor work and 25 % more plane traffic.
constexpr int SG_S = 4;             // slices (base-256 digits) per operand: 4 or 5
constexpr int SG_M = 128;           // events per tile (UMMA M)
constexpr int SG_NC = SG_S == 4 ? 64 : 32;   // output neurons per accumulator set (UMMA N): two sets of S x N columns fill the 512 TMEM columns.
                                    // (N = 32 leaves the MMA bound by the shared-memory reads of its 128 x 32 A operand: 40 cycles for 16 of math)
constexpr int SG_KC = 64;           // input neurons per pipeline stage (two K = 32 MMA steps)
constexpr int SG_QBITS = 8 * SG_S - 2;        // |q| <= 2^Q
constexpr double SG_CS = (2.0 * SG_S + 0.5) / (double)(1ULL << (8 * SG_S - 1));   // (2 + 2 (S - 1) + slack) * 2^-(Q+1): see DESIGN.md
struct SgLayer {
  int32_t K, Kp, N, Np, act, last;     // Kp = K rounded up to SG_KC (zero digits), Np = N rounded up to SG_NC
  int64_t w_off;       // bytes into SgDesc::weights: [Np/NC][Kp/64][S][4][NC][16] int8 -- one (nc, kc) stage = S * NC * 64 contiguous bytes
  int64_t tau_off;     // index into SgDesc::tau_exp (Np entries, all equal to tau_max_exp: one scale per layer)
  int64_t b_off;       // index into NeuralDesc::biases (N entries) / RegressionDesc::intercept
  double amax, bmax;   // max_j sum_i |w_ji| (rounded up), max_j |b_j|
  int32_t tau_max_exp, norm;
};
struct SgDesc {
  int enabled, n_layers;
  int inputs_identity;    // NeuralInput i is a plain FieldRef to vector slot i (no mapMissingTo): rows can be read straight from X
  int pad;
  SgLayer layers[MAX_LAYERS]; const int8_t* weights; const int32_t* tau_exp;
};
struct SgRow { double err; float amax; int32_t sexp; };   // per event between layers: error bound so far (inf = not on this path), max |a_i|, exponent of sigma

// ---- regression / neural network ---------------------------------------------------------------------
enum : int { NORM_NONE = 0, NORM_SIMPLEMAX = 1, NORM_SOFTMAX = 2, NORM_LOGIT = 3, NORM_PROBIT = 4, NORM_CLOGLOG = 5,
             NORM_EXP = 6, NORM_LOGLOG = 7, NORM_CAUCHIT = 8 };
struct RegTerm { int32_t slot; int32_t exponent; double coef; };           // NumericPredictor
struct RegCat { int32_t slot; int32_t pad; double value; double coef; };   // CategoricalPredictor on a numeric field
struct RegressionDesc {
  int n_tables, classify, norm, dense;  // dense: every table lists all F active fields in slot order, exponent 1
  const double* intercept;    // n_tables
  const int32_t* term_off;    // n_tables+1
  const RegTerm* terms;
  const int32_t* cat_off;     // n_tables+1
  const RegCat* cats;
  const double* dense_wt;     // dense only: F x KP transposed, KP = n_tables rounded up to a multiple of 8, 10 or 12 (least padding), zero padded
  double dense_amax, dense_bmax;   // dense only: max_t sum_f |coef|, max_t |intercept| (error bound of the FMA path)
  SgDesc sg;                       // dense softmax classification with <= 24 tables: weight digit planes of the fused tensor-core layer
};

// ---- GeneralRegressionModel (PPMatrix x ParamMatrix) -------------------------------------------------------
enum : int { GR_REGRESSION = 0, GR_GENERAL_LINEAR = 1, GR_MULTINOMIAL = 2 };
enum : int { GL_IDENTITY = 0, GL_LOG, GL_LOGIT, GL_PROBIT, GL_CLOGLOG, GL_LOGLOG, GL_LOGC, GL_CAUCHIT, GL_POWER, GL_NEGBIN };
struct GCell { int32_t slot; int32_t is_factor; double value; };  // factor: indicator [x == value]; covariate: x ^ value
static constexpr int GR_MAX_PARAMS = 64, GR_MAX_CATS = 32;
struct GeneralRegDesc {
  int n_params, n_rows, n_cats, classify, model_type, link;
  double link_param, offset;
  const int32_t* cell_off;    // n_params + 1
  const GCell* cells;
  const int32_t* beta_off;    // n_rows + 1
  const int32_t* beta_param;  // parameter index per PCell
  const double* beta;
  const int32_t* row_cat;     // n_rows: index of the row's targetCategory among the labels (-1: regression)
};

enum : int { ACT_THRESHOLD = 0, ACT_LOGISTIC, ACT_TANH, ACT_IDENTITY, ACT_EXPONENTIAL, ACT_RECIPROCAL, ACT_SQUARE,
             ACT_GAUSS, ACT_SINE, ACT_COSINE, ACT_ELLIOTT, ACT_ARCTAN, ACT_RECTIFIER,
             ACT_RADIAL };   // radialBasis: exp(fanIn ln(altitude) - sum (x - w)^2 / (2 width^2)); biases hold [bias | 2 width^2 | fanIn ln(altitude)] x n_out
enum : int { NI_FIELDREF = 0, NI_NORMCONT = 1, NI_NORMDISC = 2, NI_CONSTANT = 3 };
struct NeuralInput {  // DerivedField of a NeuralInput
  int32_t kind, slot, pts_off, n_pts, outliers, has_map_missing;  // outliers: 0 asIs, 1 asMissingValues, 2 asExtremeValues
  double value, map_missing;
};
struct NeuralLayer { int32_t n_in, n_out, act, norm; double threshold; int64_t w_off, b_off; int64_t wt_off; int32_t n_out_pad, pad; };

struct NeuralDesc {
  int n_inputs, n_layers, classify, n_outputs, max_width;
  NeuralLayer layers[MAX_LAYERS];
  const NeuralInput* inputs;
  const double* pts;          // LinearNorm (orig, norm) pairs
  const double* weights;      // per layer row-major [n_out][n_in]
  const double* weights_t;    // per layer transposed [n_in][n_out_pad] (n_out rounded up to 8, zero padded): streamed in row chunks
  const double* biases;
  const int32_t* out_neuron;  // n_outputs: index into the last layer
  int out_denorm_off, out_denorm_n;  // regression: NormContinuous of the output to invert (n = 0: none)
  SgDesc sg;                  // tensor-core form (classification MLPs with 1-Lipschitz activations, layers up to 256 inputs)
};

// ---- derived fields (TransformationDictionary / LocalTransformations) ---------------------------------------------
// Every DerivedField is a postfix program over a small value stack (value, present).  A pre-pass kernel prepares the raw
// slots, runs the programs and writes an augmented row [F_in raw + n_derived] that the family kernel then reads.
enum : int { DX_CONST = 0, DX_FIELD, DX_NORMDISC, DX_NORMCONT, DX_DISCRETIZE, DX_MAPVALUES, DX_APPLY };
enum : int { DF_ADD = 0, DF_SUB, DF_MUL, DF_DIV, DF_MIN, DF_MAX, DF_SUM, DF_AVG, DF_PRODUCT, DF_POW, DF_EXP, DF_LN, DF_LOG10, DF_SQRT,
             DF_ABS, DF_FLOOR, DF_CEIL, DF_ROUND, DF_THRESHOLD, DF_EQ, DF_NE, DF_LT, DF_LE, DF_GT, DF_GE, DF_AND, DF_OR, DF_NOT,
             DF_ISIN, DF_ISNOTIN, DF_IF, DF_ISMISSING, DF_ISNOTMISSING,
             DF_COALESCE,    // FieldRef to a DefineFunction parameter carrying mapMissingTo: the argument's value, or x when it is missing
             DF_USERCALL };  // DefineFunction call: [arguments when mapMissingTo needs them] body; outliers = 1: float result
struct DxOp {
  int32_t op, a, b, c, flags;   // a: field slot / function id; b, c: pool offset + count, or argument count; flags: 1 mapMissingTo, 2 defaultValue
  int32_t outliers;             // NORMCONT: 0 asIs, 1 asMissingValues, 2 asExtremeValues
  double x, y, z;               // x: constant / match / mapMissingTo value, y: defaultValue, z: spare
};
static constexpr int DX_STACK = 16;
struct DerivedDesc {
  int n_derived;
  const int32_t* prog_off;   // n_derived + 1
  const DxOp* ops;
  const double* pool;        // LinearNorm pairs, Discretize bins (l, r, flags, value), MapValues pairs
  const uint8_t* dtype;      // n_derived: dataType of the derived field (double / float / integer)
};

struct DeviceModel {
  int kind, F, prep_simple, n_labels;
  int F_in;                  // size of the caller's vectors (= F without derived fields; F = F_in + derived.n_derived with them)
  const PrepField* prep_in; const double* prep_in_pool;   // preparation of the F_in raw slots (used by the pre-pass when deriving)
  int prep_in_simple;
  DerivedDesc derived;
  int prep_identity;  // every active field is a plain double: a present value is used as is
  int target_status;  // 0, or ST_EXTRACT when no result can ever be extracted (no named target / non-double target)
  const PrepField* prep; const double* prep_pool;
  const double* label_value; const uint8_t* label_status;  // class label / cluster id -> Double (String.toDouble)
  TargetXform tx;
  ForestDesc forest; ClusterDesc cluster; RegressionDesc reg; NeuralDesc nn; GeneralRegDesc greg;
};

// ---- host side ---------------------------------------------------------------------------------------
struct CompileError : std::runtime_error { using std::runtime_error::runtime_error; };

// Host copy of everything; pointers inside `dev` are rebased onto the device blob by the engine.
struct CompiledModel {
  std::string model_name, family;
  std::vector<std::string> active_names;
  std::vector<std::string> labels;
  DeviceModel dev{};            // scalar fields valid; pointer fields filled at upload
  // one byte blob + (offset-from-blob, address-of-pointer-field) fixups
  std::vector<uint8_t> blob;
  struct Fixup { size_t off; size_t field_off; };  // field_off: byte offset of the pointer inside DeviceModel
  std::vector<Fixup> fixups;
  size_t n_nodes = 0;           // statistics
  // per vector slot, when the model can be scored bit-identically from FLOAT rows (fjp_score_batch_f32): 0 = the field is a float
  // field (plain cast), 1 = round the double up to float, 2 = round it down.  Empty: double rows only.
  std::vector<uint8_t> float_row_modes;
};

CompiledModel compile_pmml(const char* data, size_t len);

}  // namespace fjp
