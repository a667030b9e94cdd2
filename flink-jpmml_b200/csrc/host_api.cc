// host_api.cc -- implementation of host/flinkjpmml.hpp and the extern "C" operator-emulator entry points declared in
// include/fjpmml_host.h (what pytest drives in place of a Flink mini-cluster; T/utils/FlinkTestKits.scala:32-67).
#include <algorithm>
#include <chrono>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <thread>

#include "../../include/fjpmml_host.h"
#include "host/flinkjpmml.hpp"

namespace flinkjpmml {

// ---------------------------------------------------------------------------------------- ModelId
int32_t javaStringHashCode(const std::string& s) {
  uint32_t h = 0;  // java.lang.String.hashCode over UTF-16 units; ids are ASCII
  for (unsigned char c : s) h = 31u * h + c;
  return (int32_t)h;
}
int32_t ModelId::hashCode() const { return javaStringHashCode(identifier()); }

// RegexParsers semantics of S/models/core/ModelId.scala:32-51: whitespace is skipped before every token; the name is
// a lowercase-hex UUID; the version token must be all digits up to the end of the input.
ModelId ModelId::fromIdentifier(const std::string& id) {
  size_t p = 0;
  auto skip_ws = [&] { while (p < id.size() && (id[p] == ' ' || id[p] == '\t' || id[p] == '\n' || id[p] == '\r' || id[p] == '\f' || id[p] == '\v')) p++; };
  auto hex = [&](size_t n) { for (size_t i = 0; i < n; i++, p++) { if (p >= id.size()) return false; char c = id[p]; if (!((c >= '0' && c <= '9') || (c >= 'a' && c <= 'f'))) return false; } return true; };
  auto dash = [&] { if (p < id.size() && id[p] == '-') { p++; return true; } return false; };
  skip_ws();
  size_t start = p;
  bool ok = hex(8) && dash() && hex(4) && dash() && hex(4) && dash() && hex(4) && dash() && hex(12);
  if (!ok) throw WrongModelIdFormat("string matching regex `[0-9a-f]{8}-[0-9a-f]{4}-[0-9a-f]{4}-[0-9a-f]{4}-[0-9a-f]{12}' expected");
  std::string name = id.substr(start, p - start);
  skip_ws();
  if (p >= id.size() || id[p] != '_') throw WrongModelIdFormat("`_' expected");
  p++;
  skip_ws();
  if (p >= id.size()) throw WrongModelIdFormat("string matching regex `^\\d+$' expected");
  int64_t v = 0;
  for (size_t q = p; q < id.size(); q++) {
    if (id[q] < '0' || id[q] > '9') throw WrongModelIdFormat("string matching regex `^\\d+$' expected");
    if (v > (std::numeric_limits<int64_t>::max() - (id[q] - '0')) / 10) throw WrongModelIdFormat("version does not fit a Long");
    v = v * 10 + (id[q] - '0');
  }
  return ModelId{name, v};
}

Metadata MetadataManager::apply(const ServingMessage& m, const Metadata& md) {
  Metadata out = md;
  if (m.kind == ServingMessage::Add) { if (!out.count(m.modelId())) out[m.modelId()] = m.modelInfo(); }  // existing id: ignored (+WARN in the reference)
  else out.erase(m.modelId());
  return out;
}

// ---------------------------------------------------------------------------------------- reader + model
std::string ModelReader::buildDistributedPath() const {
  std::ifstream f(sourcePath, std::ios::binary);
  if (!f) throw std::runtime_error("cannot open " + sourcePath);
  std::ostringstream ss; ss << f.rdbuf();
  return ss.str();
}

PmmlModel PmmlModel::fromReader(fjp_engine* e, const ModelReader& reader) {
  std::string text;
  try { text = reader.buildDistributedPath(); } catch (std::exception& ex) { throw ModelLoadingException(ex.what()); }
  fjp_model* m = nullptr; char err[512] = {0};
  if (fjp_model_compile(e, text.data(), text.size(), &m, err, sizeof err) != 0) throw ModelLoadingException(err);
  return PmmlModel(e, m);
}
std::string PmmlModel::modelName() const {
  if (!model_) throw EmptyEvaluatorException("EmptyEvaluator.None");
  return fjp_model_name(model_.get());
}
Prediction PmmlModel::predict(const Vector& v, std::optional<double> replaceNan) const {
  Prediction p;
  if (!model_) { p.status = FJP_ST_NOMODEL; return p; }                 // EmptyEvaluatorException -> EmptyScore
  if (v.size != activeFields()) { p.status = FJP_ST_SIZE; return p; }     // InputValidationException -> EmptyScore
  MicroBatch mb(v.size, 1);
  mb.append(v, replaceNan);
  if (mb.score(engine_, model_.get()) != 0) throw std::runtime_error(fjp_last_error());
  p.status = mb.statuses()[0];
  if (p.status == FJP_ST_SCORE) p.value = Target::score(mb.scores()[0]);
  return p;
}

// ---------------------------------------------------------------------------------------- micro-batch
MicroBatch::MicroBatch(int n_fields, int capacity) : F_(n_fields), W_((n_fields + 31) / 32), cap_(capacity) {
  size_t bx = (size_t)cap_ * F_ * 8, bp = (size_t)cap_ * W_ * 4, b8 = (size_t)cap_ * 8;
  X_ = static_cast<double*>(fjp_host_alloc(bx ? bx : 8)); present_ = static_cast<uint32_t*>(fjp_host_alloc(bp ? bp : 8));
  replace_ = static_cast<double*>(fjp_host_alloc(b8)); score_ = static_cast<double*>(fjp_host_alloc(b8));
  status_ = static_cast<uint8_t*>(fjp_host_alloc(cap_));
  if (!X_ || !present_ || !replace_ || !score_ || !status_) throw std::runtime_error("pinned host allocation failed");
}
MicroBatch::~MicroBatch() { fjp_host_free(X_); fjp_host_free(present_); fjp_host_free(replace_); fjp_host_free(score_); fjp_host_free(status_); }

void MicroBatch::need_present() {   // first sparse row of the batch: the dense rows before it become "all present"
  if (any_sparse_) return;
  any_sparse_ = true; memset(present_, 0xFF, (size_t)n_ * W_ * 4);
}
void MicroBatch::need_replace() {
  if (any_replace_) return;
  any_replace_ = true; for (int r = 0; r < n_; r++) replace_[r] = std::numeric_limits<double>::quiet_NaN();
}
int MicroBatch::append(const Vector& v, std::optional<double> replaceNan) {
  if (v.sparse) need_present();
  if (replaceNan) need_replace();
  int r = n_++;
  double* x = X_ + (size_t)r * F_; uint32_t* p = present_ + (size_t)r * W_;
  if (!v.sparse) {  // DenseVector: every slot present (VectorConverter.scala:68-73)
    memcpy(x, v.data.data(), (size_t)F_ * 8);
    if (any_sparse_) for (int w = 0; w < W_; w++) p[w] = 0xFFFFFFFFu;
  } else {          // SparseVector: only listed indices are present (VectorConverter.scala:80-86, 100-103)
    memset(x, 0, (size_t)F_ * 8); memset(p, 0, (size_t)W_ * 4);
    for (size_t k = 0; k < v.indices.size(); k++) { int i = v.indices[k]; if (i >= 0 && i < F_) { x[i] = v.data[k]; p[i >> 5] |= 1u << (i & 31); } }
  }
  if (any_replace_) replace_[r] = replaceNan ? *replaceNan : std::numeric_limits<double>::quiet_NaN();
  return r;
}
int MicroBatch::append_dense(const double* x) { return append_dense_block(x, 1); }
int MicroBatch::append_dense_block(const double* X, int rows) {
  const int r = n_; n_ += rows;
  memcpy(X_ + (size_t)r * F_, X, (size_t)rows * F_ * 8);
  if (any_sparse_) memset(present_ + (size_t)r * W_, 0xFF, (size_t)rows * W_ * 4);
  if (any_replace_) for (int k = 0; k < rows; k++) replace_[r + k] = std::numeric_limits<double>::quiet_NaN();
  return r;
}

int MicroBatch::commit_rows(int rows) {
  const int r = n_; n_ += rows;
  if (any_sparse_) memset(present_ + (size_t)r * W_, 0xFF, (size_t)rows * W_ * 4);
  if (any_replace_) for (int k = 0; k < rows; k++) replace_[r + k] = std::numeric_limits<double>::quiet_NaN();
  return r;
}

size_t PredictionQueue::ready_end() const {
  if (!open_) return status_.size();
  const void* f = memchr(status_.data() + head_, kOpen, status_.size() - head_);
  return f ? (size_t)(static_cast<const uint8_t*>(f) - status_.data()) : status_.size();
}
long PredictionQueue::pop(double* scores, uint8_t* status, long cap) {
  const long n = std::min<long>(cap, (long)size());
  for (long i = 0; i < n; i++) { const uint8_t st = status_[head_ + i]; status[i] = st; scores[i] = st == FJP_ST_SCORE ? score_[head_ + i] : std::numeric_limits<double>::quiet_NaN(); }
  head_ += n;
  if (head_ == status_.size()) { status_.clear(); score_.clear(); head_ = 0; }
  return n;
}
std::vector<Prediction> PredictionQueue::drain() {
  const size_t end = ready_end();
  std::vector<Prediction> r; r.reserve(end - head_);
  for (size_t i = head_; i < end; i++) { Prediction p; p.status = status_[i]; if (p.status == FJP_ST_SCORE) p.value = Target::score(score_[i]); r.push_back(p); }
  head_ = end;
  if (head_ == status_.size()) { status_.clear(); score_.clear(); head_ = 0; }
  return r;
}
int MicroBatch::submit(fjp_engine* e, fjp_model* m, int* ticket) {
  return fjp_submit(e, m, X_, any_sparse_ ? present_ : nullptr, any_replace_ ? replace_ : nullptr, n_, F_, score_, status_, nullptr, ticket);
}
// fjp_submit never blocks; FJP_BUSY = every stream slot of the engine is taken (other subtasks share it).  The caller gives
// back its own oldest ticket through `wait_oldest` (returns false when it holds none: then only other threads can free a
// slot, and they always can, so yielding is enough).
template <class WaitOldest>
static void submit_or_throw(MicroBatch& mb, fjp_engine* e, fjp_model* m, int* ticket, WaitOldest&& wait_oldest) {
  for (;;) {
    int rc = mb.submit(e, m, ticket);
    if (rc == 0) return;
    if (rc != FJP_BUSY) throw std::runtime_error(fjp_last_error());
    if (!wait_oldest()) std::this_thread::yield();
  }
}
int MicroBatch::score(fjp_engine* e, fjp_model* m) {
  if (n_ == 0) return 0;
  return fjp_score_batch(e, m, X_, any_sparse_ ? present_ : nullptr, any_replace_ ? replace_ : nullptr, n_, F_, score_, status_, nullptr);
}

// ---------------------------------------------------------------------------------------- EvaluationFunction
void EvaluationFunction::open() {
  model_ = PmmlModel::fromReader(engine_, reader_);   // ModelLoadingException propagates: the job fails (EvaluationFunction.scala:45-48)
  for (auto& b : batch_) b = std::make_unique<MicroBatch>(model_.activeFields(), max_batch_);
}
void EvaluationFunction::processElement(const Vector& v, std::optional<double> replaceNan) {
  if (model_.isEmpty()) open();   // `lazy val evaluator`
  if (v.size != model_.activeFields()) pending_[cur_].push_back(-1);
  else pending_[cur_].push_back(batch_[cur_]->append(v, replaceNan));
  if (batch_[cur_]->full() || (int)pending_[cur_].size() >= max_batch_) submit_current();
}
void EvaluationFunction::processDense(const double* x, int size) {
  if (model_.isEmpty()) open();
  if (size != model_.activeFields()) pending_[cur_].push_back(-1);
  else pending_[cur_].push_back(batch_[cur_]->append_dense(x));
  if (batch_[cur_]->full() || (int)pending_[cur_].size() >= max_batch_) submit_current();
}
void EvaluationFunction::processDenseBlock(const double* X, long n, int size) {
  if (model_.isEmpty()) open();
  if (size != model_.activeFields()) { for (long i = 0; i < n; i++) processDense(X + i * (long)size, size); return; }
  for (long i = 0; i < n;) {   // whole runs of rows go into the pinned batch with one copy
    const int take = (int)std::min<long>(n - i, std::min(batch_[cur_]->room(), max_batch_ - (int)pending_[cur_].size()));
    const int r0 = batch_[cur_]->append_dense_block(X + i * (long)size, take);
    const size_t p0 = pending_[cur_].size(); pending_[cur_].resize(p0 + take);
    for (int k = 0; k < take; k++) pending_[cur_][p0 + k] = r0 + k;
    i += take;
    if (batch_[cur_]->full() || (int)pending_[cur_].size() >= max_batch_) submit_current();
  }
}
double* EvaluationFunction::reserve(int size, int want_rows, int* granted) {
  if (model_.isEmpty()) open();
  *granted = 0;
  if (size != model_.activeFields() || want_rows <= 0) return nullptr;
  *granted = std::min(want_rows, std::min(batch_[cur_]->room(), max_batch_ - (int)pending_[cur_].size()));
  return batch_[cur_]->tail();
}
void EvaluationFunction::commit(int rows) {
  if (rows <= 0) return;
  const int r0 = batch_[cur_]->commit_rows(rows);
  const size_t p0 = pending_[cur_].size(); pending_[cur_].resize(p0 + rows);
  for (int k = 0; k < rows; k++) pending_[cur_][p0 + k] = r0 + k;
  if (batch_[cur_]->full() || (int)pending_[cur_].size() >= max_batch_) submit_current();
}
void EvaluationFunction::collect(int b) {
  if (ticket_[b] >= 0) { if (fjp_wait(engine_, ticket_[b]) != 0) throw std::runtime_error(fjp_last_error()); ticket_[b] = -1; }
  const double* sc = batch_[b]->scores(); const uint8_t* st = batch_[b]->statuses();
  for (int r : pending_[b]) { if (r < 0) out_.push(FJP_ST_SIZE, 0); else out_.push(st[r], sc[r]); }
  pending_[b].clear(); batch_[b]->clear();
}
void EvaluationFunction::submit_current() {
  const int b = cur_;
  if (batch_[b]->size() > 0)
    submit_or_throw(*batch_[b], engine_, model_.handle(), &ticket_[b], [&] {   // the older batch is the only other ticket this operator can hold
      if (ticket_[b ^ 1] < 0) return false;
      if (fjp_wait(engine_, ticket_[b ^ 1]) != 0) throw std::runtime_error(fjp_last_error());
      ticket_[b ^ 1] = -1; return true;
    });
  collect(b ^ 1);      // the older batch: its results come first, and its buffer becomes the one being packed
  cur_ = b ^ 1;
}
void EvaluationFunction::flush() {
  if (model_.isEmpty()) return;
  if (!pending_[cur_].empty()) submit_current();
  collect(cur_ ^ 1);   // whatever is still in flight (submit_current already collected the batch before it)
}

// ---------------------------------------------------------------------------------------- EvaluationCoFunction
void EvaluationCoFunction::initializeState(const std::vector<Metadata>& restored) {
  metadata_.clear();
  for (auto& m : restored) for (auto& kv : m) metadata_.insert(kv);   // union list state: every subtask's map
}
Metadata EvaluationCoFunction::snapshotState() { flush(); return metadata_; }

PmmlModel EvaluationCoFunction::loadModel(const std::string& path) { return PmmlModel::fromReader(engine_, ModelReader(path)); }

PmmlModel EvaluationCoFunction::fromMetadata(const std::string& modelId) {
  ModelId id = ModelId::fromIdentifier(modelId);                       // WrongModelIdFormat is not caught: job-fatal
  auto it = metadata_.find(id);
  if (it == metadata_.end()) return PmmlModel::empty();
  for (auto& p : parked_) if (!p->detached && p->id == id) {   // the bulk path parked events on this load: settle them, the model is cached then
    flush_buckets(true);
    if (fjp_model* cached = fjp_table_get(engine_, id.name.c_str(), id.version)) return PmmlModel(engine_, cached);
    break;
  }
  PmmlModel loaded;
  auto lf = loading_.find(id);
  if (lf != loading_.end()) { auto fut = lf->second; loading_.erase(lf); loaded = fut.get(); }   // background load started by the Add (rethrows its failure)
  else loaded = loadModel(it->second.path);                            // e.g. after a checkpoint restore: load on the first event
  fjp_table_add(engine_, id.name.c_str(), id.version, loaded.handle()); // servingModels += (id -> model), keyed by the full id
  return loaded;
}

std::unique_ptr<MicroBatch> EvaluationCoFunction::take_batch(int fields) {
  for (size_t i = 0; i < pool_.size(); i++)
    if (pool_[i]->fields() == fields && pool_[i]->capacity() == bucket_rows()) { auto mb = std::move(pool_[i]); pool_.erase(pool_.begin() + i); return mb; }
  return std::make_unique<MicroBatch>(fields, bucket_rows());
}
int EvaluationCoFunction::bucket_for(const PmmlModel& m) {
  for (size_t i = 0; i < buckets_.size(); i++) if (buckets_[i].model.handle() == m.handle()) return (int)i;
  buckets_.push_back(Bucket{m, take_batch(m.activeFields())});
  return (int)buckets_.size() - 1;
}

// Where the events of `modelId` go while its model is still loading in the background (bulk path only): the index of its Parked
// record, created on the first such event; -1 = not parked (bound already, unknown id, load finished, or no background load).
int EvaluationCoFunction::park_index(const std::string& modelId, int size) {
  if (parked_.empty() && loading_.empty()) return -1;
  if (fjp_table_epoch(engine_) == resolved_epoch_ && resolved_.count(modelId)) return -1;
  const ModelId id = ModelId::fromIdentifier(modelId);
  auto ready = [](const std::shared_future<PmmlModel>& f) { return f.wait_for(std::chrono::seconds(0)) == std::future_status::ready; };
  for (size_t i = 0; i < parked_.size(); i++) {
    Parked& p = *parked_[i];
    if (p.detached || !(p.id == id)) continue;
    const bool over = (int)p.batches.size() >= kMaxParkedBatches && p.batches.back()->full();
    if (ready(p.fut) || p.size != size || over) { flush_buckets(p.size != size || over); return -1; }   // settles it (waiting only when it has to)
    return (int)i;
  }
  auto lf = loading_.find(id);
  if (lf == loading_.end() || !metadata_.count(id) || ready(lf->second)) return -1;   // ready: the usual path picks it up without waiting
  auto p = std::make_unique<Parked>(); p->id = id; p->fut = lf->second; p->size = size;
  loading_.erase(lf);
  parked_.push_back(std::move(p));
  return (int)parked_.size() - 1;
}

void EvaluationCoFunction::settle(Parked& p) {
  PmmlModel loaded = p.fut.get();   // rethrows the load's ModelLoadingException: job-fatal, as on the first event in the reference
  if (!p.detached && metadata_.count(p.id)) { fjp_table_add(engine_, p.id.name.c_str(), p.id.version, loaded.handle()); }
  const bool size_ok = p.size == loaded.activeFields();
  size_t s = 0;
  for (auto& mb : p.batches) {
    if (size_ok && mb->score(engine_, loaded.handle()) != 0) throw std::runtime_error(fjp_last_error());
    const int n = mb->size();
    for (int r = 0; r < n; r++, s++) { if (size_ok) out_.fill(p.slots[s], mb->statuses()[r], mb->scores()[r]); else out_.fill(p.slots[s], FJP_ST_SIZE, 0); }
    mb->clear(); if (pool_.size() < 64) pool_.push_back(std::move(mb));
  }
  p.batches.clear(); p.slots.clear();
}

PmmlModel EvaluationCoFunction::resolve(const std::string& modelId) {
  if (!engine_) throw std::runtime_error("no engine: scoring needs a CUDA device (there is no CPU path)");
  const uint64_t epoch = fjp_table_epoch(engine_);
  if (epoch != resolved_epoch_) { resolved_.clear(); resolved_epoch_ = epoch; }
  auto hit = resolved_.find(modelId);
  if (hit != resolved_.end()) return hit->second;
  // the reference looks the cache up by the raw id's hashCode first; an id that was never parsed cannot be cached,
  // so parsing first is equivalent except that a malformed id fails here instead of in fromMetadata -- same outcome
  ModelId id = ModelId::fromIdentifier(modelId);
  PmmlModel model;
  if (fjp_model* cached = fjp_table_get(engine_, id.name.c_str(), id.version)) model = PmmlModel(engine_, cached);
  else model = fromMetadata(modelId);
  if (!model.isEmpty()) { resolved_epoch_ = fjp_table_epoch(engine_); resolved_[modelId] = model; }   // empty bindings are not remembered
  return model;
}

void EvaluationCoFunction::enqueue(const PmmlModel& model, const Vector* v, const double* dense, int size, std::optional<double> replaceNan) {
  if (model.isEmpty()) pending_.push_back({-1, FJP_ST_NOMODEL});
  else if (size != model.activeFields()) pending_.push_back({-1, FJP_ST_SIZE});
  else {
    int b = bucket_for(model);
    int row = v ? buckets_[b].batch->append(*v, replaceNan) : buckets_[b].batch->append_dense(dense);
    pending_.push_back({b, row});
    if (buckets_[b].batch->full()) flush_buckets(false);
  }
  if ((int)pending_.size() >= max_batch_) flush_buckets(false);
}

void EvaluationCoFunction::processElement1(const std::string& modelId, const Vector& v, std::optional<double> replaceNan) {
  // servingModels.getOrElse(event.modelId.hashCode, fromMetadata(event.modelId))   (S/package.scala:112)
  enqueue(resolve(modelId), &v, nullptr, v.size, replaceNan);
}
void EvaluationCoFunction::processDense1(const std::string& modelId, const double* x, int size) {
  enqueue(resolve(modelId), nullptr, x, size, std::nullopt);
}

void EvaluationCoFunction::processDenseBatch(const char* const* ids, const int32_t* id_index, const double* X, long n, int size) {
  int n_ids = 0; for (long i = 0; i < n; i++) n_ids = std::max(n_ids, id_index[i] + 1);
  // per id of this call: bucket (>= 0), -1 = empty model, -2 = size mismatch, -3 = not looked at yet
  std::vector<int> where(n_ids, -3); uint64_t seen_flush = flushes_;
  for (long i = 0; i < n; i++) {
    const int k = id_index[i];
    if (seen_flush != flushes_) { std::fill(where.begin(), where.end(), -3); seen_flush = flushes_; }   // a flush dropped the buckets
    int w = where[k];
    if (w == -3) {
      const uint64_t before = flushes_;
      const int parked = park_index(ids[k], size);
      if (before != flushes_) { std::fill(where.begin(), where.end(), -3); seen_flush = flushes_; }   // park_index settled a finished load
      if (parked >= 0) w = -4 - parked;
      else {
        PmmlModel model = resolve(ids[k]);
        w = model.isEmpty() ? -1 : (size != model.activeFields() ? -2 : bucket_for(model));
      }
      where[k] = w;
    }
    if (w <= -4) {   // the model is still loading: park the row, keep its place in the output order
      Parked& p = *parked_[-4 - w];
      if (p.batches.empty() || p.batches.back()->full()) {
        if ((int)p.batches.size() >= kMaxParkedBatches) { where[k] = -3; i--; continue; }   // too much parked: look again (park_index waits for the load)
        p.batches.push_back(take_batch(size));
      }
      p.batches.back()->append_dense(X + i * (long)size); parked_events_++;
      pending_.push_back({-2 - (-4 - w), 0});
    }
    else if (w < 0) pending_.push_back({-1, w == -1 ? FJP_ST_NOMODEL : FJP_ST_SIZE});
    else {
      pending_.push_back({w, buckets_[w].batch->append_dense(X + i * (long)size)});
      if (buckets_[w].batch->full()) flush_buckets(false);
    }
    if ((int)pending_.size() >= max_batch_) flush_buckets(false);
  }
}

void EvaluationCoFunction::processElement2(const ServingMessage& control) {
  // manageModels first, then manageMetadata (EvaluationCoFunction.scala:69-74, Appendix B of SURVEY.md)
  if (control.kind == ServingMessage::Del && engine_) fjp_table_del(engine_, control.name.c_str(), control.version);
  const bool known = metadata_.count(control.modelId()) != 0;
  metadata_ = MetadataManager::apply(control, metadata_);
  resolved_.clear();   // bindings are looked up again after any control message
  if (control.kind == ServingMessage::Del) {
    loading_.erase(control.modelId());
    for (auto& p : parked_) if (p->id == control.modelId()) p->detached = true;   // its parked events are still scored; the model is not cached
  }
  else if (!known && engine_) {   // a new id: parse + compile + upload off the task thread
    fjp_engine* eng = engine_; std::string path = control.path;
    loading_[control.modelId()] = std::async(std::launch::async, [eng, path] { return PmmlModel::fromReader(eng, ModelReader(path)); }).share();
  }
  // buffered events keep the model they were bound to (each bucket holds a reference), so nothing is flushed here
}

void EvaluationCoFunction::flush() { flush_buckets(true); }

EvaluationCoFunction::~EvaluationCoFunction() {
  if (inflight_ && inflight_->ticket >= 0 && engine_) fjp_wait(engine_, inflight_->ticket);   // nothing may outlive the pinned buffers
}

void EvaluationCoFunction::collect(Generation& g) {
  if (g.ticket >= 0) { const int t = g.ticket; g.ticket = -1; if (fjp_wait(engine_, t) != 0) throw std::runtime_error(fjp_last_error()); }
  for (size_t i = 0; i < g.pending.size(); i++) {
    const Pending& p = g.pending[i];
    if (p.bucket >= 0) out_.fill(g.slot0 + i, g.buckets[p.bucket].batch->statuses()[p.row], g.buckets[p.bucket].batch->scores()[p.row]);
  }
  for (auto& b : g.buckets) { b.batch->clear(); if (pool_.size() < 96) pool_.push_back(std::move(b.batch)); }
  g.buckets.clear(); g.pending.clear();   // dropping the buckets releases models deleted in the meantime
}

void EvaluationCoFunction::flush_buckets(bool wait_for_parked) {
  struct Timer { double& acc; std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
                 ~Timer() { acc += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); } } timer{flush_seconds_};
  auto settle_parked = [&] {
    for (size_t i = 0; i < parked_.size();) {
      Parked& p = *parked_[i];
      if (!wait_for_parked && p.fut.wait_for(std::chrono::seconds(0)) != std::future_status::ready) { i++; continue; }
      std::unique_ptr<Parked> done = std::move(parked_[i]); parked_.erase(parked_.begin() + i);   // out of the list first: settle may throw
      settle(*done);
    }
  };
  if (!pending_.empty()) {
    auto gen = std::make_unique<Generation>();
    gen->buckets = std::move(buckets_); gen->pending = std::move(pending_); buckets_.clear(); pending_.clear();
    gen->slot0 = out_.reserve_n(gen->pending.size());
    for (size_t i = 0; i < gen->pending.size(); i++) {   // what needs no GPU is known now; parked events learn their output position
      const Pending& p = gen->pending[i];
      if (p.bucket <= -2) parked_[-2 - p.bucket]->slots.push_back(gen->slot0 + i);
      else if (p.bucket < 0) out_.fill(gen->slot0 + i, (uint8_t)p.row, 0);
    }
    std::vector<fjp_batch_desc> descs;
    for (auto& b : gen->buckets) if (b.batch->size()) descs.push_back(b.batch->desc(b.model.handle()));
    while (!descs.empty()) {
      const int rc = fjp_submit_many(engine_, descs.data(), (int)descs.size(), &gen->ticket);
      if (rc == 0) break;
      if (rc != FJP_BUSY) throw std::runtime_error(fjp_last_error());
      if (inflight_ && inflight_->ticket >= 0) collect(*inflight_);   // every stream slot of the engine is taken: give back our own first
      else std::this_thread::yield();                                  // the others are held by other subtasks, which always finish
    }
    if (inflight_) collect(*inflight_);   // the older generation: its predictions come first, its buffers go back to the pool
    inflight_ = std::move(gen);
    flushes_++;
  }
  if (wait_for_parked && inflight_) { collect(*inflight_); inflight_.reset(); }
  if (!parked_.empty()) { settle_parked(); flushes_++; }   // nothing refers to parked_ by index any more
}

}  // namespace flinkjpmml

// ============================================================================================ C wrappers
using namespace flinkjpmml;

struct fjh_eval_function { std::unique_ptr<EvaluationFunction> op; };
struct fjh_cofunction { std::unique_ptr<EvaluationCoFunction> op; };

static void set_err(char* err, size_t cap, const char* msg) { if (err && cap) { strncpy(err, msg, cap - 1); err[cap - 1] = 0; } }
static Vector make_vector(const double* values, int size, const int* indices, int nnz) {
  if (!indices) return Vector::dense(std::vector<double>(values, values + size));
  return Vector::sparseOf(size, std::vector<int>(indices, indices + nnz), std::vector<double>(values, values + nnz));
}
static std::optional<double> opt(double r) { return std::isnan(r) ? std::nullopt : std::optional<double>(r); }
template <class F> static int guarded(char* err, size_t cap, F&& f) {
  try { f(); return FJH_OK; }
  catch (WrongModelIdFormat& e) { set_err(err, cap, e.what()); return FJH_WRONG_MODEL_ID_FORMAT; }
  catch (ModelLoadingException& e) { set_err(err, cap, e.what()); return FJH_MODEL_LOADING; }
  catch (std::exception& e) { set_err(err, cap, e.what()); return FJH_ENGINE_ERROR; }
}

extern "C" {

int fjh_model_id_parse(const char* id, char* name_out, size_t name_cap, int64_t* version_out) {
  try { ModelId m = ModelId::fromIdentifier(id); if (name_out && name_cap) { strncpy(name_out, m.name.c_str(), name_cap - 1); name_out[name_cap - 1] = 0; } if (version_out) *version_out = m.version; return FJH_OK; }
  catch (WrongModelIdFormat&) { return FJH_WRONG_MODEL_ID_FORMAT; }
}
int32_t fjh_model_id_hash(const char* name, int64_t version) { return ModelId{name, version}.hashCode(); }

int fjh_predict(fjp_engine* e, fjp_model* m, const double* values, int size, const int* indices, int nnz, double replace_nan,
                double* score, uint8_t* status) {
  try {
    PmmlModel model; if (m) { fjp_model_retain(m); model = PmmlModel(e, m); }
    Prediction p = model.predict(make_vector(values, size, indices, nnz), opt(replace_nan));
    *score = p.value.empty ? std::numeric_limits<double>::quiet_NaN() : p.value.value; *status = p.status; return FJH_OK;
  } catch (std::exception&) { return FJH_ENGINE_ERROR; }
}

fjh_eval_function* fjh_eval_function_open(fjp_engine* e, const char* path, int max_batch, char* err, size_t errcap) {
  auto h = std::make_unique<fjh_eval_function>();
  h->op = std::make_unique<EvaluationFunction>(e, ModelReader(path), max_batch);
  if (guarded(err, errcap, [&] { h->op->open(); }) != FJH_OK) return nullptr;
  return h.release();
}
const char* fjh_eval_function_model_name(fjh_eval_function* h) { static thread_local std::string s; s = h->op->evaluator().modelName(); return s.c_str(); }
int fjh_eval_function_process(fjh_eval_function* h, const double* values, int size, const int* indices, int nnz, double replace_nan) {
  return guarded(nullptr, 0, [&] { h->op->processElement(make_vector(values, size, indices, nnz), opt(replace_nan)); });
}
int fjh_eval_function_process_dense(fjh_eval_function* h, const double* X, long n, int size) {
  return guarded(nullptr, 0, [&] { h->op->processDenseBlock(X, n, size); });
}
double* fjh_eval_function_reserve(fjh_eval_function* h, int size, int want_rows, int* granted_rows) {
  double* p = nullptr; int g = 0;
  if (guarded(nullptr, 0, [&] { p = h->op->reserve(size, want_rows, &g); }) != FJH_OK) { p = nullptr; g = 0; }
  if (granted_rows) *granted_rows = g;
  return p;
}
int fjh_eval_function_commit(fjh_eval_function* h, int rows) { return guarded(nullptr, 0, [&] { h->op->commit(rows); }); }
int fjh_eval_function_flush(fjh_eval_function* h) { return guarded(nullptr, 0, [&] { h->op->flush(); }); }
long fjh_eval_function_poll(fjh_eval_function* h, double* scores, uint8_t* status, long cap) {
  return h->op->poll(scores, status, cap);
}
void fjh_eval_function_close(fjh_eval_function* h) { if (h) { try { h->op->close(); } catch (...) {} delete h; } }

fjh_cofunction* fjh_cofunction_create(fjp_engine* e, int max_batch) {
  auto h = new fjh_cofunction(); h->op = std::make_unique<EvaluationCoFunction>(e, max_batch); h->op->initializeState({}); return h;
}
int fjh_cofunction_process_event(fjh_cofunction* h, const char* model_id, const double* values, int size, const int* indices, int nnz,
                                 double replace_nan, char* err, size_t errcap) {
  return guarded(err, errcap, [&] { h->op->processElement1(model_id, make_vector(values, size, indices, nnz), opt(replace_nan)); });
}
int fjh_cofunction_process_dense(fjh_cofunction* h, const char* const* ids, const int32_t* id_index, const double* X, long n, int size,
                                 char* err, size_t errcap) {
  return guarded(err, errcap, [&] { h->op->processDenseBatch(ids, id_index, X, n, size); });
}
int fjh_cofunction_process_control(fjh_cofunction* h, int is_add, const char* name, int64_t version, const char* path, int64_t occurred_on) {
  return guarded(nullptr, 0, [&] { h->op->processElement2(is_add ? AddMessage(name, version, path ? path : "", occurred_on) : DelMessage(name, version, occurred_on)); });
}
int fjh_cofunction_flush(fjh_cofunction* h, char* err, size_t errcap) { return guarded(err, errcap, [&] { h->op->flush(); }); }
long fjh_cofunction_poll(fjh_cofunction* h, double* scores, uint8_t* status, long cap) {
  return h->op->poll(scores, status, cap);
}
// snapshot format: one "name_version<TAB>path<LF>" line per entry (the checkpointed HashMap[ModelId, ModelInfo])
long fjh_cofunction_snapshot(fjh_cofunction* h, char* buf, long cap) {
  std::string s;
  try { for (auto& kv : h->op->snapshotState()) s += kv.first.identifier() + "\t" + kv.second.path + "\n"; } catch (std::exception&) { return -1; }
  if ((long)s.size() + 1 > cap) return -(long)s.size() - 1;
  memcpy(buf, s.c_str(), s.size() + 1);
  return (long)s.size();
}
int fjh_cofunction_restore(fjh_cofunction* h, const char* const* snapshots, int n_snapshots) {
  return guarded(nullptr, 0, [&] {
    std::vector<Metadata> all;
    for (int k = 0; k < n_snapshots; k++) {
      Metadata md; std::istringstream in(snapshots[k]); std::string line;
      while (std::getline(in, line)) { size_t tab = line.find('\t'); if (tab == std::string::npos) continue; md[ModelId::fromIdentifier(line.substr(0, tab))] = ModelInfo{line.substr(tab + 1)}; }
      all.push_back(std::move(md));
    }
    h->op->initializeState(all);
  });
}
int fjh_cofunction_metadata_size(fjh_cofunction* h) { return (int)h->op->servingMetadata().size(); }
int fjh_cofunction_cached_models(fjh_cofunction* h) { return h->op->servingModels(); }
long fjh_cofunction_parked_events(fjh_cofunction* h) { return h->op->parkedEvents(); }
double fjh_cofunction_flush_seconds(fjh_cofunction* h) { return h->op->flushSeconds(); }
void fjh_cofunction_destroy(fjh_cofunction* h) { delete h; }

}  // extern "C"
