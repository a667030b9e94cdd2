// engine.cu -- engine, device model table and the extern "C" boundary declared in include/fjpmml.h.
//
// Stands where the reference keeps per-operator JVM state:
//   the lazily loaded evaluator     S/api/functions/EvaluationFunction.scala:43-48
//   servingModels (model cache)     S/api/functions/EvaluationCoFunction.scala:66-67, 98-131
//   PmmlModel.predict per event     S/api/PmmlModel.scala:109-119     (here: per micro-batch, on the GPU)
// (S/ = flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/)
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/fjpmml.h"
#include "kernels.h"
#include "model.h"

using namespace fjp;

static thread_local std::string g_last_error;
static int fail(const std::string& msg, char* err = nullptr, size_t cap = 0) {
  g_last_error = msg;
  if (err && cap) { strncpy(err, msg.c_str(), cap - 1); err[cap - 1] = 0; }
  return -1;
}
#define CK(call)                                                                                      \
  do {                                                                                                \
    cudaError_t e_ = (call);                                                                          \
    if (e_ != cudaSuccess) return fail(std::string(#call) + ": " + cudaGetErrorString(e_));           \
  } while (0)

struct fjp_model {
  CompiledModel cm;
  std::atomic<int> refs{1};
  int device = -1;
  void* d_blob = nullptr;
  DeviceModel dev{};
  std::string describe;
};

static void rebase(fjp_model* m) {
  m->dev = m->cm.dev;
  for (auto& fx : m->cm.fixups) {
    const void* p = static_cast<const uint8_t*>(m->d_blob) + fx.off;
    memcpy(reinterpret_cast<uint8_t*>(&m->dev) + fx.field_off, &p, sizeof(p));
  }
}

static void make_describe(fjp_model* m) {
  const DeviceModel& d = m->cm.dev; std::string s = "family=" + m->cm.family;
  if (d.kind == MK_FOREST) {
    s += " trees=" + std::to_string(d.forest.n_trees) + " format=" + (d.forest.format == FOREST_COMPLETE ? "complete" : "general") +
         " depth=" + std::to_string(d.forest.depth) + " enc=" + (d.forest.enc == ENC_FLOAT ? "float" : "rank") +
         " method=" + std::to_string(d.forest.method) + " nodes=" + std::to_string(m->cm.n_nodes) + " compact=" + std::to_string(d.forest.compact.enabled + d.forest.compact.wide) +   // 1: float thresholds, 2: 64-bit threshold images
         (d.forest.compact.sparse ? " sparse=1" : "");
    ForestLaunch fl;
    if (m->d_blob && d.forest.method != FM_SINGLE && plan_forest_complete(d, &fl))
      s += " tile_events=" + std::to_string(fl.events_per_cta) + " trees_per_stage=" + std::to_string(fl.trees_per_stage) + " interleave=" + std::to_string(fl.trees_interleaved) +
           " smem=" + std::to_string(fl.smem_bytes);
  } else if (d.kind == MK_CLUSTERING) s += " k=" + std::to_string(d.cluster.k) + " fields=" + std::to_string(d.cluster.nf) + " fast=" + std::to_string(d.cluster.fast);
  else if (d.kind == MK_REGRESSION) s += " tables=" + std::to_string(d.reg.n_tables) + " dense=" + std::to_string(d.reg.dense) + " norm=" + std::to_string(d.reg.norm);
  else if (d.kind == MK_GENERAL_REGRESSION) s += " params=" + std::to_string(d.greg.n_params) + " rows=" + std::to_string(d.greg.n_rows) + " type=" + std::to_string(d.greg.model_type) + " link=" + std::to_string(d.greg.link);
  else if (d.kind == MK_NEURAL) s += " inputs=" + std::to_string(d.nn.n_inputs) + " layers=" + std::to_string(d.nn.n_layers) + " max_width=" + std::to_string(d.nn.max_width);
  if (d.derived.n_derived) s += " derived=" + std::to_string(d.derived.n_derived);
  s += " F=" + std::to_string(d.F_in) + " prep_simple=" + std::to_string(d.prep_simple) + " blob_bytes=" + std::to_string(m->cm.blob.size());
  m->describe = s;
}

struct Table {
  std::map<std::pair<std::string, int64_t>, fjp_model*> models;
  uint64_t epoch = 0;
  Table() = default;
  Table(const Table& o) : models(o.models), epoch(o.epoch) { for (auto& kv : models) fjp_model_retain(kv.second); }
  ~Table() { for (auto& kv : models) fjp_model_release(kv.second); }
};

static constexpr int NSLOT = 8;        // stream slots shared by every host thread that uses the engine
static constexpr int MAX_INFLIGHT = 3; // open tickets one caller may hold (fjp_max_inflight)
struct Slot {
  cudaStream_t s = nullptr; cudaEvent_t done = nullptr;
  cudaEvent_t copied = nullptr;   // the copy-in of the chunk this slot carries has finished (the next chunk's copy-in waits for it)
  uint8_t* buf = nullptr; size_t cap = 0;   // one device arena per slot, carved per call
  bool busy = false; fjp_model* held = nullptr;
  const uint8_t* pending_status = nullptr; long pending_n = 0;   // submitted batch whose outcomes are counted at fjp_wait
  std::vector<fjp_model*> held_more; std::vector<std::pair<const uint8_t*, long>> pending_more;   // fjp_submit_many: the other batches of the ticket
};

struct fjp_engine {
  int device = 0;
  Slot slots[NSLOT];
  std::mutex slot_mu; std::condition_variable slot_cv;
  std::mutex table_mu; std::shared_ptr<const Table> table{std::make_shared<Table>()};
  std::atomic<long> launches{0};
  std::atomic<uint64_t> status_counts[6];   // per-batch aggregation of the outcomes the reference logs one WARN line for
  unsigned long long* d_stats = nullptr;    // device counters of the tensor-core path (BatchArgs::stats)
  fjp_engine() { for (auto& c : status_counts) c = 0; }
};

static int acquire_slot(fjp_engine* e) {
  std::unique_lock<std::mutex> lk(e->slot_mu);
  for (;;) {
    for (int i = 0; i < NSLOT; i++) if (!e->slots[i].busy) { e->slots[i].busy = true; return i; }
    e->slot_cv.wait(lk);
  }
}
static int try_acquire_slot(fjp_engine* e) {
  std::lock_guard<std::mutex> lk(e->slot_mu);
  for (int i = 0; i < NSLOT; i++) if (!e->slots[i].busy) { e->slots[i].busy = true; return i; }
  return -1;
}
static void release_slot(fjp_engine* e, int i) {
  { std::lock_guard<std::mutex> lk(e->slot_mu); e->slots[i].busy = false; }
  e->slot_cv.notify_one();
}
static int slot_reserve(Slot& s, size_t bytes) {
  if (bytes <= s.cap) return 0;
  if (s.buf) { CK(cudaStreamSynchronize(s.s)); CK(cudaFree(s.buf)); s.buf = nullptr; s.cap = 0; }
  size_t cap = (bytes + (size_t(1) << 20)) & ~((size_t(1) << 20) - 1);
  CK(cudaMalloc(&s.buf, cap));
  s.cap = cap;
  return 0;
}

static size_t score_scratch_bytes(const fjp_model* m, long n, bool want_probs = false, bool want_node_ids = false);
static bool env_flag(const char* name) { const char* v = getenv(name); return v && *v && *v != '0'; }

struct Carve {  // sub-allocation inside a slot arena
  size_t off = 0;
  size_t take(size_t bytes) { size_t o = off; off = (off + bytes + 255) & ~size_t(255); return o; }
};
struct HostBatch {
  const double* X; const uint32_t* present; const double* replace; long n; int vec_size;
  double* score; uint8_t* status; int32_t* aux; int32_t* node_ids; int node_stride; double* probs; int prob_stride;
  const float* Xf;   // float rows (fjp_score_batch_f32): X is NULL then
};

// FJP_TRACE_CHUNKS=1 (dev aid): device timeline of the chunk pipeline of one host-buffer call, printed to stderr
struct ChunkTrace { std::vector<cudaEvent_t> ev; bool on = false; };
static thread_local ChunkTrace g_trace;
static void trace_mark(cudaStream_t st) {
  if (!g_trace.on) return;
  cudaEvent_t ev; if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, st); g_trace.ev.push_back(ev);
}

// enqueue H2D + kernel + D2H for rows [lo, hi) of hb on slot s
// copy_after: the previous chunk's `copied` event.  Two copy-ins issued on two streams would otherwise run side by side at half
// the PCIe rate each, finish together, and leave the link idle while both chunks' kernels run; chained, every copy-in has the
// link to itself and the kernels of chunk i run under the copy-in of chunk i + 1.
// shared: the arena is already reserved and carved from *shared on (several batches behind one ticket, fjp_submit_many);
// dry_run: only advance *shared (sizing pass).
static int enqueue_chunk(fjp_engine* e, fjp_model* m, Slot& s, const HostBatch& hb, long lo, long hi, cudaEvent_t copy_after = nullptr,
                         Carve* shared = nullptr, bool dry_run = false) {
  const long n = hi - lo; const int F = hb.vec_size, W = (F + 31) / 32;
  Carve own; Carve& c = shared ? *shared : own;
  size_t oX = c.take((size_t)n * F * 8), oP = hb.present ? c.take((size_t)n * W * 4) : 0, oR = hb.replace ? c.take((size_t)n * 8) : 0;
  size_t oF = hb.Xf ? c.take((size_t)n * F * 4) : 0;
  size_t oS = c.take((size_t)n * 8), oT = c.take((size_t)n), oA = hb.aux ? c.take((size_t)n * 4) : 0;
  size_t oN = hb.node_ids ? c.take((size_t)n * hb.node_stride * 4) : 0, oQ = hb.probs ? c.take((size_t)n * hb.prob_stride * 8) : 0;
  const size_t dsz = score_scratch_bytes(m, n, hb.probs != nullptr, hb.node_ids != nullptr); size_t oD = dsz ? c.take(dsz) : 0;   // augmented rows (derived fields) / leaf values (small-batch forest path) / digit planes
  if (dry_run) return 0;
  if (!shared && slot_reserve(s, c.off)) return -1;
  BatchArgs a{}; a.stats = e->d_stats;
  a.X = reinterpret_cast<double*>(s.buf + oX); a.n = n; a.present_stride = W;
  int launches = 0;
  if (copy_after) CK(cudaStreamWaitEvent(s.s, copy_after, 0));
  trace_mark(s.s);
  if (hb.Xf) CK(cudaMemcpyAsync(s.buf + oF, hb.Xf + lo * (long)F, (size_t)n * F * 4, cudaMemcpyHostToDevice, s.s));   // half the bytes over PCIe
  else CK(cudaMemcpyAsync(s.buf + oX, hb.X + lo * (long)F, (size_t)n * F * 8, cudaMemcpyHostToDevice, s.s));
  if (hb.present) { a.present = reinterpret_cast<uint32_t*>(s.buf + oP); CK(cudaMemcpyAsync(s.buf + oP, hb.present + lo * (long)W, (size_t)n * W * 4, cudaMemcpyHostToDevice, s.s)); }
  if (hb.replace) { a.replace = reinterpret_cast<double*>(s.buf + oR); CK(cudaMemcpyAsync(s.buf + oR, hb.replace + lo, (size_t)n * 8, cudaMemcpyHostToDevice, s.s)); }
  CK(cudaEventRecord(s.copied, s.s));
  if (hb.Xf) CK(launch_widen_f32(reinterpret_cast<const float*>(s.buf + oF), reinterpret_cast<double*>(s.buf + oX), n * (long)F, s.s, &launches));   // widened on the device (exact)
  a.score = reinterpret_cast<double*>(s.buf + oS); a.status = s.buf + oT;
  a.aux = hb.aux ? reinterpret_cast<int32_t*>(s.buf + oA) : nullptr;
  a.node_ids = hb.node_ids ? reinterpret_cast<int32_t*>(s.buf + oN) : nullptr; a.node_stride = hb.node_stride;
  a.probs = hb.probs ? reinterpret_cast<double*>(s.buf + oQ) : nullptr; a.prob_stride = hb.prob_stride;
  if (hb.node_ids) CK(cudaMemsetAsync(s.buf + oN, 0xFF, (size_t)n * hb.node_stride * 4, s.s));
  if (hb.probs) CK(cudaMemsetAsync(s.buf + oQ, 0xFF, (size_t)n * hb.prob_stride * 8, s.s));
  trace_mark(s.s);
  CK(launch_score_any(m->dev, a, dsz ? s.buf + oD : nullptr, s.s, &launches));
  trace_mark(s.s);
  e->launches += launches;
  CK(cudaMemcpyAsync(hb.score + lo, s.buf + oS, (size_t)n * 8, cudaMemcpyDeviceToHost, s.s));
  CK(cudaMemcpyAsync(hb.status + lo, s.buf + oT, (size_t)n, cudaMemcpyDeviceToHost, s.s));
  if (hb.aux) CK(cudaMemcpyAsync(hb.aux + lo, s.buf + oA, (size_t)n * 4, cudaMemcpyDeviceToHost, s.s));
  if (hb.node_ids) CK(cudaMemcpyAsync(hb.node_ids + lo * (long)hb.node_stride, s.buf + oN, (size_t)n * hb.node_stride * 4, cudaMemcpyDeviceToHost, s.s));
  if (hb.probs) CK(cudaMemcpyAsync(hb.probs + lo * (long)hb.prob_stride, s.buf + oQ, (size_t)n * hb.prob_stride * 8, cudaMemcpyDeviceToHost, s.s));
  trace_mark(s.s);
  return 0;
}

// work area a batch of n events wants besides its inputs and outputs: augmented rows (derived fields) or the per-tree leaf
// values of the small-batch forest path
static size_t score_scratch_bytes(const fjp_model* m, long n, bool want_probs, bool want_node_ids) {
  size_t d = derived_scratch_bytes(m->dev, n);
  if (d) return d;
  // the fast paths look at the same fields of the batch the launcher will see (probabilities / node ids asked for or not)
  BatchArgs t{}; t.n = n; t.probs = want_probs ? reinterpret_cast<double*>(8) : nullptr; t.node_ids = want_node_ids ? reinterpret_cast<int32_t*>(8) : nullptr;
  if (size_t sl = sliced_scratch_bytes(m->dev, t)) return sl;       // digit planes of the tensor-core path (dense layers)
  if (size_t df = dense_fma_scratch_bytes(m->dev, t)) return df;    // re-scoring list of the FMA paths (dense tile kernel)
  return forest_slice_scratch_bytes(m->dev, n);
}

// status histogram of a finished batch (host side; rows are mostly 0, so whole zero words are skipped)
static void count_statuses(fjp_engine* e, const uint8_t* st, long n) {
  uint64_t c[6] = {0, 0, 0, 0, 0, 0}; long i = 0;
  for (; i < n && (reinterpret_cast<uintptr_t>(st + i) & 7); i++) c[st[i] < 6 ? st[i] : 3]++;
  for (; i + 8 <= n; i += 8) {
    uint64_t w; memcpy(&w, st + i, 8);
    if (w == 0) c[0] += 8; else for (int k = 0; k < 8; k++) c[st[i + k] < 6 ? st[i + k] : 3]++;
  }
  for (; i < n; i++) c[st[i] < 6 ? st[i] : 3]++;
  for (int k = 0; k < 6; k++) if (c[k]) e->status_counts[k] += c[k];
}

// Small micro-batches whose buffers are all pinned are scored in place: the kernel reads the rows from host memory over
// PCIe and writes score / status straight back (zero-copy), so the batch costs one launch and one synchronisation instead
// of three to six copy calls around the launch.  Returns 1 when taken, 0 when the copy path should run, -1 on error.
static constexpr long kZeroCopyMaxRows = 8;   // the tree-sliced latency path re-reads the rows once per slice: keep that off PCIe
static bool pinned_dev_ptr(const void* host, const void** dev) {
  if (!host) { *dev = nullptr; return true; }
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); return false; }
  if (at.type != cudaMemoryTypeHost || !at.devicePointer) return false;
  *dev = at.devicePointer; return true;
}
static int try_zero_copy(fjp_engine* e, fjp_model* m, Slot& s, const HostBatch& hb) {
  if (hb.n > kZeroCopyMaxRows || hb.node_ids || hb.probs) return 0;
  const void *dX, *dP, *dR, *dS, *dT, *dA;
  if (!pinned_dev_ptr(hb.X, &dX) || !pinned_dev_ptr(hb.present, &dP) || !pinned_dev_ptr(hb.replace, &dR) || !pinned_dev_ptr(hb.score, &dS) ||
      !pinned_dev_ptr(hb.status, &dT) || !pinned_dev_ptr(hb.aux, &dA)) return 0;
  const size_t dsz = score_scratch_bytes(m, hb.n);
  if (dsz && slot_reserve(s, dsz)) return -1;
  BatchArgs a{}; a.X = static_cast<const double*>(dX); a.present = static_cast<const uint32_t*>(dP); a.replace = static_cast<const double*>(dR);
  a.n = hb.n; a.present_stride = (hb.vec_size + 31) / 32;
  a.score = static_cast<double*>(const_cast<void*>(dS)); a.status = static_cast<uint8_t*>(const_cast<void*>(dT)); a.aux = static_cast<int32_t*>(const_cast<void*>(dA));
  int launches = 0;
  cudaError_t ce = launch_score_any(m->dev, a, dsz ? s.buf : nullptr, s.s, &launches);
  e->launches += launches;
  if (ce != cudaSuccess) { fail(std::string("scoring kernel: ") + cudaGetErrorString(ce)); return -1; }
  return 1;
}

// whole-batch outcomes that need no device work: empty model (5) and size mismatch (1)
static bool host_short_circuit(fjp_model* m, const HostBatch& hb) {
  int st = -1;
  if (!m) st = ST_NOMODEL;
  else if (hb.vec_size != m->cm.dev.F_in) st = ST_SIZE;
  if (st < 0) return false;
  uint64_t nanbits = 0x7FF8000000000000ULL; double nanv; memcpy(&nanv, &nanbits, 8);
  for (long i = 0; i < hb.n; i++) { hb.score[i] = nanv; hb.status[i] = (uint8_t)st; if (hb.aux) hb.aux[i] = -1; }
  return true;
}

extern "C" {

const char* fjp_version(void) { return "flink-jpmml-b200 0.1 (sm_100a)"; }
const char* fjp_last_error(void) { return g_last_error.c_str(); }
int fjp_max_inflight(void) { return MAX_INFLIGHT; }

int fjp_engine_create(int device, fjp_engine** out, char* err, size_t errcap) {
  int count = 0;
  cudaError_t ce = cudaGetDeviceCount(&count);
  if (ce != cudaSuccess || count == 0)
    return fail(std::string("no CUDA device available (") + cudaGetErrorString(ce) + "): this engine has no CPU scoring path", err, errcap);
  if (device < 0 || device >= count) return fail("device ordinal out of range", err, errcap);
  if (cudaSetDevice(device) != cudaSuccess) return fail("cudaSetDevice failed", err, errcap);
  auto* e = new fjp_engine(); e->device = device;
  {  // fjp_score_device takes its work areas from the stream-ordered allocator: keep freed blocks in the pool between calls
    cudaMemPool_t pool; uint64_t keep = UINT64_MAX;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    cudaGetLastError();
  }
  if (cudaMalloc(&e->d_stats, 64) != cudaSuccess || cudaMemset(e->d_stats, 0, 64) != cudaSuccess) { delete e; return fail("engine counters: cudaMalloc failed", err, errcap); }
  for (int i = 0; i < NSLOT; i++) {
    if (cudaStreamCreateWithFlags(&e->slots[i].s, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->slots[i].done, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->slots[i].copied, cudaEventDisableTiming) != cudaSuccess) {
      delete e; return fail("stream/event creation failed", err, errcap);
    }
  }
  *out = e; return 0;
}

void fjp_engine_destroy(fjp_engine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  for (auto& s : e->slots) { if (s.s) cudaStreamSynchronize(s.s); if (s.held) fjp_model_release(s.held); if (s.buf) cudaFree(s.buf); if (s.done) cudaEventDestroy(s.done); if (s.copied) cudaEventDestroy(s.copied); if (s.s) cudaStreamDestroy(s.s); }
  { std::lock_guard<std::mutex> lk(e->table_mu); e->table.reset(); }
  if (e->d_stats) cudaFree(e->d_stats);
  delete e;
}
int fjp_engine_device(const fjp_engine* e) { return e->device; }
int fjp_engine_sync(fjp_engine* e) { CK(cudaSetDevice(e->device)); for (auto& s : e->slots) CK(cudaStreamSynchronize(s.s)); return 0; }
long fjp_engine_launch_count(const fjp_engine* e) { return e->launches.load(); }
void fjp_engine_status_counts(const fjp_engine* e, uint64_t out[6]) { for (int k = 0; k < 6; k++) out[k] = e->status_counts[k].load(); }
int fjp_engine_fast_path_counts(fjp_engine* e, uint64_t out[4]) {
  if (fjp_engine_sync(e) != 0) return -1;
  unsigned long long h[4] = {0, 0, 0, 0};
  CK(cudaMemcpy(h, e->d_stats, sizeof h, cudaMemcpyDeviceToHost));
  for (int k = 0; k < 4; k++) out[k] = h[k];
  return 0;
}

int fjp_model_compile(fjp_engine* e, const char* pmml, size_t len, fjp_model** out, char* err, size_t errcap) {
  std::unique_ptr<fjp_model> m(new fjp_model());
  try { m->cm = compile_pmml(pmml, len); }
  catch (std::exception& ex) { return fail(ex.what(), err, errcap); }
  if (e) {
    if (cudaSetDevice(e->device) != cudaSuccess) return fail("cudaSetDevice failed", err, errcap);
    m->device = e->device;
    size_t bytes = std::max<size_t>(m->cm.blob.size(), 256);
    cudaError_t ce = cudaMalloc(&m->d_blob, bytes);
    if (ce != cudaSuccess) return fail(std::string("cudaMalloc(model): ") + cudaGetErrorString(ce), err, errcap);
    // upload on a slot stream so that scoring on the other streams is not stalled
    cudaStream_t st = e->slots[NSLOT - 1].s;
    ce = cudaMemcpyAsync(m->d_blob, m->cm.blob.data(), m->cm.blob.size(), cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
    if (ce != cudaSuccess) { cudaFree(m->d_blob); return fail(std::string("model upload: ") + cudaGetErrorString(ce), err, errcap); }
    rebase(m.get());
  }
  make_describe(m.get());
  *out = m.release(); return 0;
}

void fjp_model_retain(fjp_model* m) { if (m) m->refs.fetch_add(1); }
void fjp_model_release(fjp_model* m) {
  if (!m) return;
  if (m->refs.fetch_sub(1) == 1) {
    if (m->d_blob) { int cur = 0; cudaGetDevice(&cur); cudaSetDevice(m->device); cudaFree(m->d_blob); cudaSetDevice(cur); }
    delete m;
  }
}

int fjp_model_replicate(fjp_engine* dst, const fjp_model* src, fjp_model** out, char* err, size_t errcap) {
  if (!src || !src->d_blob) return fail("source model is not device-resident", err, errcap);
  std::unique_ptr<fjp_model> m(new fjp_model());
  m->cm = src->cm; m->device = dst->device;
  if (cudaSetDevice(dst->device) != cudaSuccess) return fail("cudaSetDevice failed", err, errcap);
  size_t bytes = std::max<size_t>(m->cm.blob.size(), 256);
  cudaError_t ce = cudaMalloc(&m->d_blob, bytes);
  if (ce != cudaSuccess) return fail(std::string("cudaMalloc(model): ") + cudaGetErrorString(ce), err, errcap);
  cudaStream_t st = dst->slots[NSLOT - 1].s;
  ce = cudaMemcpyPeerAsync(m->d_blob, dst->device, src->d_blob, src->device, m->cm.blob.size(), st);
  if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
  if (ce != cudaSuccess) { cudaFree(m->d_blob); return fail(std::string("peer copy: ") + cudaGetErrorString(ce), err, errcap); }
  rebase(m.get()); make_describe(m.get());
  *out = m.release(); return 0;
}

int fjp_model_num_fields(const fjp_model* m) { return m->cm.dev.F_in; }
const char* fjp_model_field_name(const fjp_model* m, int i) { return (i >= 0 && i < (int)m->cm.active_names.size()) ? m->cm.active_names[i].c_str() : nullptr; }
const char* fjp_model_name(const fjp_model* m) { return m->cm.model_name.c_str(); }
const char* fjp_model_family(const fjp_model* m) { return m->cm.family.c_str(); }
int fjp_model_num_labels(const fjp_model* m) { return (int)m->cm.labels.size(); }
const char* fjp_model_label(const fjp_model* m, int i) { return (i >= 0 && i < (int)m->cm.labels.size()) ? m->cm.labels[i].c_str() : nullptr; }
int fjp_model_num_segments(const fjp_model* m) { return m->cm.dev.kind == MK_FOREST ? m->cm.dev.forest.n_trees : 0; }
const char* fjp_model_describe(const fjp_model* m) { return m->describe.c_str(); }
size_t fjp_model_device_bytes(const fjp_model* m) { return m->d_blob ? m->cm.blob.size() : 0; }

int fjp_table_add(fjp_engine* e, const char* name, int64_t version, fjp_model* m) {
  if (!m) return fail("fjp_table_add: null model");
  std::lock_guard<std::mutex> lk(e->table_mu);
  auto cur = std::atomic_load(&e->table);
  auto key = std::make_pair(std::string(name), version);
  if (cur->models.count(key)) return 1;
  auto next = std::make_shared<Table>(*cur);
  fjp_model_retain(m); next->models[key] = m; next->epoch = cur->epoch + 1;
  std::atomic_store(&e->table, std::shared_ptr<const Table>(next));
  return 0;
}
int fjp_table_del(fjp_engine* e, const char* name, int64_t version) {
  std::lock_guard<std::mutex> lk(e->table_mu);
  auto cur = std::atomic_load(&e->table);
  auto key = std::make_pair(std::string(name), version);
  if (!cur->models.count(key)) return 1;
  auto next = std::make_shared<Table>(*cur);
  fjp_model_release(next->models[key]); next->models.erase(key); next->epoch = cur->epoch + 1;
  std::atomic_store(&e->table, std::shared_ptr<const Table>(next));
  return 0;
}
fjp_model* fjp_table_get(fjp_engine* e, const char* name, int64_t version) {
  auto cur = std::atomic_load(&e->table);  // the snapshot keeps its models alive while we look
  auto it = cur->models.find(std::make_pair(std::string(name), version));
  if (it == cur->models.end()) return nullptr;
  fjp_model_retain(it->second); return it->second;
}
int fjp_table_size(const fjp_engine* e) { return (int)std::atomic_load(&e->table)->models.size(); }
uint64_t fjp_table_epoch(const fjp_engine* e) { return std::atomic_load(&e->table)->epoch; }

void* fjp_host_alloc(size_t bytes) { void* p = nullptr; if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) return nullptr; return p; }
void fjp_host_free(void* p) { if (p) cudaFreeHost(p); }

static int score_host_batch(fjp_engine* e, fjp_model* m, const HostBatch& hb) {
  const long n = hb.n; const int vec_size = hb.vec_size;
  if (n <= 0) return 0;
  if (host_short_circuit(m, hb)) { count_statuses(e, hb.status, n); return 0; }
  if (!m->d_blob) return fail("model was compiled without an engine: it cannot score");
  CK(cudaSetDevice(e->device));
  if (m->device != e->device) return fail("model lives on another device; use fjp_model_replicate");
  // chunked 2-slot pipeline: copy-in of chunk i+1 overlaps the kernel and copy-out of chunk i
  long rows = (long)((size_t(24) << 20) / ((size_t)vec_size * (hb.Xf ? 4 : 8) + 16));
  if (rows < 4096) rows = 4096; if (rows > (1 << 18)) rows = 1 << 18;
  // large batches: eight chunks (up to 2^18 rows each) -- a chunk then fills the chip on the throughput kernels, and the
  // un-overlapped first copy / last kernel stay an eighth of the batch (measured on C4: tools/e2e_chunks.py)
  if (n / 8 > rows) rows = std::min<long>(n / 8, 1 << 18);
  if (const char* v = getenv("FJP_CHUNK_ROWS")) { long r = atol(v); if (r >= 256) rows = r; }   // tuning aid
  int s0 = acquire_slot(e), s1 = n > rows ? try_acquire_slot(e) : -1;
  int used[2] = {s0, s1}; int ns = s1 >= 0 ? 2 : 1; int rc = 0; long k = 0;
  g_trace.on = env_flag("FJP_TRACE_CHUNKS");
  // the last chunk's kernels and copy-out run under no copy-in: taper the end (half, quarter, quarter of a chunk)
  for (long lo = 0; lo < n && rc == 0; k++) {
    long take = std::min(rows, n - lo);
    if (ns > 1 && n - lo <= rows && n - lo > 65536) take = (n - lo + 1) / 2;
    rc = enqueue_chunk(e, m, e->slots[used[k % ns]], hb, lo, lo + take, (k > 0 && ns > 1) ? e->slots[used[(k - 1) % ns]].copied : nullptr);
    lo += take;
  }
  for (int i = 0; i < ns; i++) { cudaError_t ce = cudaStreamSynchronize(e->slots[used[i]].s); if (ce != cudaSuccess && rc == 0) rc = fail(std::string("scoring failed: ") + cudaGetErrorString(ce)); release_slot(e, used[i]); }
  if (g_trace.on) {   // per chunk: copy-in start, copy-in end = kernels start, kernels end, copy-out end (ms from the first mark)
    for (size_t i = 0; i + 3 < g_trace.ev.size(); i += 4) {
      float t[4]; for (int q = 0; q < 4; q++) cudaEventElapsedTime(&t[q], g_trace.ev[0], g_trace.ev[i + q]);
      fprintf(stderr, "[fjp trace] chunk %zu: h2d %.3f..%.3f  kernels ..%.3f  d2h ..%.3f ms\n", i / 4, t[0], t[1], t[2], t[3]);
    }
    for (cudaEvent_t ev : g_trace.ev) cudaEventDestroy(ev);
    g_trace.ev.clear(); g_trace.on = false;
  }
  if (rc == 0) count_statuses(e, hb.status, n);
  return rc;
}

int fjp_score_batch_ex(fjp_engine* e, fjp_model* m, const double* X, const uint32_t* present, const double* replace,
                       long n, int vec_size, double* score, uint8_t* status, int32_t* aux,
                       int32_t* node_ids, int node_stride, double* probs, int prob_stride) {
  HostBatch hb{X, present, replace, n, vec_size, score, status, aux, node_ids, node_stride, probs, prob_stride, nullptr};
  return score_host_batch(e, m, hb);
}

int fjp_model_float_row_modes(const fjp_model* m, uint8_t* modes, int cap) {
  if (!m || m->cm.float_row_modes.empty()) return 0;
  const int F = (int)m->cm.float_row_modes.size();
  if (modes) for (int i = 0; i < F && i < cap; i++) modes[i] = m->cm.float_row_modes[i];
  return F;
}

int fjp_score_batch_f32(fjp_engine* e, fjp_model* m, const float* X, const uint32_t* present, long n, int vec_size,
                        double* score, uint8_t* status, int32_t* aux) {
  if (m && m->cm.float_row_modes.empty()) return fail("fjp_score_batch_f32: this model's tests need the double values (fjp_model_float_row_modes says 0); use fjp_score_batch");
  HostBatch hb{nullptr, present, nullptr, n, vec_size, score, status, aux, nullptr, 0, nullptr, 0, X};
  return score_host_batch(e, m, hb);
}

int fjp_score_sparse(fjp_engine* e, fjp_model* m, const int64_t* row_ptr, const int32_t* indices, const double* values,
                     const double* replace, long n, int vec_size, double* score, uint8_t* status, int32_t* aux) {
  if (n <= 0) return 0;
  HostBatch hb{nullptr, nullptr, replace, n, vec_size, score, status, aux, nullptr, 0, nullptr, 0, nullptr};
  if (host_short_circuit(m, hb)) { count_statuses(e, status, n); return 0; }
  if (!m->d_blob) return fail("model was compiled without an engine: it cannot score");
  CK(cudaSetDevice(e->device));
  if (m->device != e->device) return fail("model lives on another device; use fjp_model_replicate");
  const int F = vec_size, W = (F + 31) / 32;
  long rows = (long)((size_t(24) << 20) / ((size_t)F * 8 + 16)); if (rows < 4096) rows = 4096; if (rows > (1 << 18)) rows = 1 << 18;
  int s0 = acquire_slot(e), s1 = n > rows ? try_acquire_slot(e) : -1;
  int used[2] = {s0, s1}; int ns = s1 >= 0 ? 2 : 1; int rc = 0; long k = 0;
  for (long lo = 0; lo < n && rc == 0; lo += rows, k++) {
    Slot& s = e->slots[used[k % ns]];
    const long hi = std::min(n, lo + rows), cn = hi - lo; const int64_t nz0 = row_ptr[lo], nnz = row_ptr[hi] - nz0;
    Carve c;
    size_t oRp = c.take((size_t)(cn + 1) * 8), oIx = c.take((size_t)std::max<int64_t>(nnz, 1) * 4), oVl = c.take((size_t)std::max<int64_t>(nnz, 1) * 8);
    size_t oX = c.take((size_t)cn * F * 8), oP = c.take((size_t)cn * W * 4), oR = replace ? c.take((size_t)cn * 8) : 0;   // oX, oP: scratch for the expanded rows
    size_t oS = c.take((size_t)cn * 8), oT = c.take((size_t)cn), oA = aux ? c.take((size_t)cn * 4) : 0;
    const size_t dsz = score_scratch_bytes(m, cn); size_t oD = dsz ? c.take(dsz) : 0;
    if (slot_reserve(s, c.off)) { rc = -1; break; }
    // row_ptr is re-based on the device side by passing shifted index / value pointers
    rc = [&]() -> int {
      CK(cudaMemcpyAsync(s.buf + oRp, row_ptr + lo, (size_t)(cn + 1) * 8, cudaMemcpyHostToDevice, s.s));
      if (nnz > 0) {
        CK(cudaMemcpyAsync(s.buf + oIx, indices + nz0, (size_t)nnz * 4, cudaMemcpyHostToDevice, s.s));
        CK(cudaMemcpyAsync(s.buf + oVl, values + nz0, (size_t)nnz * 8, cudaMemcpyHostToDevice, s.s));
      }
      if (replace) CK(cudaMemcpyAsync(s.buf + oR, replace + lo, (size_t)cn * 8, cudaMemcpyHostToDevice, s.s));
      int launches = 0;
      CK(launch_expand_sparse(reinterpret_cast<const long long*>(s.buf + oRp), reinterpret_cast<const int32_t*>(s.buf + oIx) - nz0,
                              reinterpret_cast<const double*>(s.buf + oVl) - nz0, cn, F, reinterpret_cast<double*>(s.buf + oX),
                              reinterpret_cast<uint32_t*>(s.buf + oP), s.s, &launches));
      BatchArgs a{}; a.stats = e->d_stats; a.X = reinterpret_cast<double*>(s.buf + oX); a.present = reinterpret_cast<uint32_t*>(s.buf + oP); a.present_stride = W;
      a.replace = replace ? reinterpret_cast<double*>(s.buf + oR) : nullptr; a.n = cn;
      a.score = reinterpret_cast<double*>(s.buf + oS); a.status = s.buf + oT; a.aux = aux ? reinterpret_cast<int32_t*>(s.buf + oA) : nullptr;
      CK(launch_score_any(m->dev, a, dsz ? s.buf + oD : nullptr, s.s, &launches));
      e->launches += launches;
      CK(cudaMemcpyAsync(score + lo, s.buf + oS, (size_t)cn * 8, cudaMemcpyDeviceToHost, s.s));
      CK(cudaMemcpyAsync(status + lo, s.buf + oT, (size_t)cn, cudaMemcpyDeviceToHost, s.s));
      if (aux) CK(cudaMemcpyAsync(aux + lo, s.buf + oA, (size_t)cn * 4, cudaMemcpyDeviceToHost, s.s));
      return 0;
    }();
  }
  for (int i = 0; i < ns; i++) { cudaError_t ce = cudaStreamSynchronize(e->slots[used[i]].s); if (ce != cudaSuccess && rc == 0) rc = fail(std::string("scoring failed: ") + cudaGetErrorString(ce)); release_slot(e, used[i]); }
  if (rc == 0) count_statuses(e, status, n);
  return rc;
}

int fjp_score_batch(fjp_engine* e, fjp_model* m, const double* X, const uint32_t* present, const double* replace,
                    long n, int vec_size, double* score, uint8_t* status, int32_t* aux) {
  return fjp_score_batch_ex(e, m, X, present, replace, n, vec_size, score, status, aux, nullptr, 0, nullptr, 0);
}

int fjp_score_device(fjp_engine* e, fjp_model* m, const double* dX, const uint32_t* dpresent, const double* dreplace,
                     long n, int vec_size, double* dscore, uint8_t* dstatus, int32_t* daux, void* cuda_stream) {
  if (n <= 0) return 0;
  CK(cudaSetDevice(e->device));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : e->slots[0].s;
  BatchArgs a{}; a.stats = e->d_stats; a.X = dX; a.present = dpresent; a.replace = dreplace; a.n = n; a.present_stride = (vec_size + 31) / 32;
  a.score = dscore; a.status = dstatus; a.aux = daux;
  int launches = 0;
  if (!m) { CK(launch_fill_status(a, ST_NOMODEL, st, &launches)); }
  else if (vec_size != m->cm.dev.F_in) { CK(launch_fill_status(a, ST_SIZE, st, &launches)); }
  else {
    if (!m->d_blob) return fail("model was compiled without an engine: it cannot score");
    if (m->device != e->device) return fail("model lives on another device; use fjp_model_replicate");
    size_t dsz = derived_scratch_bytes(m->dev, n);
    if (!dsz) dsz = sliced_scratch_bytes(m->dev, a);
    if (!dsz) dsz = dense_fma_scratch_bytes(m->dev, a);
    if (!dsz) dsz = forest_compact_scratch_bytes(m->dev, n);
    if (dsz) {   // augmented rows (derived fields) / rank-key tiles (large forest batches): stream-ordered allocation on the caller's stream
      void* scratch = nullptr; CK(cudaMallocAsync(&scratch, dsz, st));
      cudaError_t ce = launch_score_any(m->dev, a, scratch, st, &launches);
      cudaFreeAsync(scratch, st);
      if (ce != cudaSuccess) return fail(std::string("scoring kernel: ") + cudaGetErrorString(ce));
    } else CK(launch_score(m->dev, a, st, &launches));
  }
  e->launches += launches;
  return 0;
}

int fjp_submit(fjp_engine* e, fjp_model* m, const double* X, const uint32_t* present, const double* replace,
               long n, int vec_size, double* score, uint8_t* status, int32_t* aux, int* ticket) {
  HostBatch hb{X, present, replace, n, vec_size, score, status, aux, nullptr, 0, nullptr, 0, nullptr};
  *ticket = -1;
  CK(cudaSetDevice(e->device));
  // never blocks: slots are shared by all host threads, so a thread that waited here while holding tickets of its own
  // could wait for ever on a thread doing the same.  FJP_BUSY tells the caller to fjp_wait its oldest ticket and retry.
  const int si = try_acquire_slot(e);
  if (si < 0) return FJP_BUSY;
  Slot& s = e->slots[si];
  auto give_up = [&](int rc) { if (s.held) { fjp_model_release(s.held); s.held = nullptr; } s.pending_status = nullptr; s.pending_n = 0; release_slot(e, si); return rc; };
  s.pending_status = status; s.pending_n = n > 0 ? n : 0;
  if (n <= 0 || host_short_circuit(m, hb)) { *ticket = si; return 0; }
  if (!m->d_blob || m->device != e->device) return give_up(fail("model is not resident on this engine's device"));
  fjp_model_retain(m); s.held = m;  // the batch keeps its model alive across a concurrent fjp_table_del
  int zc = try_zero_copy(e, m, s, hb);
  if (zc < 0 || (zc == 0 && enqueue_chunk(e, m, s, hb, 0, n))) { cudaStreamSynchronize(s.s); return give_up(-1); }
  cudaError_t ce = cudaEventRecord(s.done, s.s);
  if (ce != cudaSuccess) { cudaStreamSynchronize(s.s); return give_up(fail(std::string("cudaEventRecord: ") + cudaGetErrorString(ce))); }
  *ticket = si;
  return 0;
}

// Several micro-batches, each for its own model, behind ONE ticket: one stream slot, one arena carved for all of them, copies and
// kernels enqueued back to back without a host-side wait in between.  This is what the dynamic operator hands over per flush
// (one bucket per model), so that the task thread can pack the next generation of events while the GPU scores this one.
int fjp_submit_many(fjp_engine* e, const fjp_batch_desc* b, int count, int* ticket) {
  *ticket = -1;
  if (count <= 0) return fail("fjp_submit_many: no batches");
  CK(cudaSetDevice(e->device));
  for (int i = 0; i < count; i++) {
    if (!b[i].model) return fail("fjp_submit_many: batch without a model");
    if (!b[i].model->d_blob || b[i].model->device != e->device) return fail("model is not resident on this engine's device");
    if (b[i].vec_size != b[i].model->cm.dev.F_in) return fail("fjp_submit_many: vector size differs from the model's field count");
  }
  const int si = try_acquire_slot(e);
  if (si < 0) return FJP_BUSY;
  Slot& s = e->slots[si];
  auto give_up = [&](int rc) {
    for (fjp_model* hm : s.held_more) fjp_model_release(hm);
    s.held_more.clear(); s.pending_more.clear(); release_slot(e, si); return rc;
  };
  auto host_batch = [&](int i) { return HostBatch{b[i].X, b[i].present, b[i].replace, b[i].n, b[i].vec_size, b[i].score, b[i].status, nullptr, nullptr, 0, nullptr, 0, nullptr}; };
  Carve sizing;
  for (int i = 0; i < count; i++) if (b[i].n > 0) enqueue_chunk(e, b[i].model, s, host_batch(i), 0, b[i].n, nullptr, &sizing, true);
  if (slot_reserve(s, sizing.off)) return give_up(-1);
  Carve carve;
  for (int i = 0; i < count; i++) {
    if (b[i].n <= 0) continue;
    fjp_model_retain(b[i].model); s.held_more.push_back(b[i].model);   // a batch keeps its model alive across a concurrent fjp_table_del
    s.pending_more.emplace_back(b[i].status, b[i].n);
    if (enqueue_chunk(e, b[i].model, s, host_batch(i), 0, b[i].n, nullptr, &carve, false)) { cudaStreamSynchronize(s.s); return give_up(-1); }
  }
  cudaError_t ce = cudaEventRecord(s.done, s.s);
  if (ce != cudaSuccess) { cudaStreamSynchronize(s.s); return give_up(fail(std::string("cudaEventRecord: ") + cudaGetErrorString(ce))); }
  *ticket = si;
  return 0;
}

int fjp_wait(fjp_engine* e, int ticket) {
  if (ticket < 0 || ticket >= NSLOT) return fail("bad ticket");
  Slot& s = e->slots[ticket];
  CK(cudaSetDevice(e->device));
  cudaError_t ce = cudaStreamSynchronize(s.s);
  if (s.held) { fjp_model_release(s.held); s.held = nullptr; }
  for (fjp_model* hm : s.held_more) fjp_model_release(hm);
  s.held_more.clear();
  if (ce == cudaSuccess && s.pending_status) count_statuses(e, s.pending_status, s.pending_n);
  if (ce == cudaSuccess) for (auto& ps : s.pending_more) count_statuses(e, ps.first, ps.second);
  s.pending_status = nullptr; s.pending_n = 0; s.pending_more.clear();
  release_slot(e, ticket);
  if (ce != cudaSuccess) return fail(std::string("scoring failed: ") + cudaGetErrorString(ce));
  return 0;
}

}  // extern "C"

// ---- internal accessors for translation units that do not see the handle layouts (declared in engine_internal.h)
namespace fjp {
const DeviceModel* internal_device_model(const fjp_model* m) { return (m && m->d_blob) ? &m->dev : nullptr; }
int internal_model_device(const fjp_model* m) { return m ? m->device : -1; }
int internal_model_fields(const fjp_model* m) { return m ? m->cm.dev.F_in : -1; }
int internal_model_derived(const fjp_model* m) { return m ? m->cm.dev.derived.n_derived : 0; }
void internal_count_launches(fjp_engine* e, int launches) { if (e) e->launches += launches; }
void internal_set_error(const char* msg) { g_last_error = msg; }
}  // namespace fjp
