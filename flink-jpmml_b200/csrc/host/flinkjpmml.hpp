// flinkjpmml.hpp -- host-side mirror (C++17) of the reference's user-facing types and its two Flink operators,
// written above the C ABI (include/fjpmml.h).  The reference is Scala (no JVM toolchain in this image), so this is
// the compiled-language restatement a Scala/JNI shim would follow; names, argument meaning and error behaviour are
// the reference's.  S/ = flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/
//
//   Vector / DenseVector / SparseVector      org.apache.flink.ml.math (input type of S/api/PmmlModel.scala:109)
//   Target / Score / EmptyScore / Prediction S/models/prediction/Target.scala:44-84, Prediction.scala:30-72
//   exceptions                               S/api/exceptions/package.scala:27-53
//   ModelId / ModelInfo                      S/models/core/ModelId.scala:32-64, ModelInfo.scala:26
//   AddMessage / DelMessage                  S/models/control/ServingMessage.scala:27-60
//   MetadataManager / ModelsManager          S/api/managers/MetadataManager.scala:39-88, ModelsManager.scala:39-59
//   ModelReader                              S/api/reader/ModelReader.scala:27, FsReader.scala:41-49 (local files only)
//   PmmlModel                                S/api/PmmlModel.scala:43-176
//   EvaluationFunction                       S/api/functions/EvaluationFunction.scala:39-50 + S/package.scala:76-82, 138-142
//   EvaluationCoFunction                     S/api/functions/EvaluationCoFunction.scala:55-133 + S/package.scala:63-65, 107-119
//
// What is different on purpose: the operators buffer events into pinned-host micro-batches and emit predictions in
// arrival order when the batch is flushed (size reached, flush(), snapshotState(), close()).  Every buffered event
// holds the model it was bound to when it arrived, so control messages keep their "takes effect between events"
// meaning (T/RichConnectedStreamSpec.scala:83-95, 187-204) without flushing.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <future>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/fjpmml.h"

namespace flinkjpmml {

// ---------------------------------------------------------------------------------------- exceptions
struct InputValidationException : std::runtime_error { using std::runtime_error::runtime_error; };
struct InputPreparationException : std::runtime_error { using std::runtime_error::runtime_error; };
struct JPMMLExtractionException : std::runtime_error { using std::runtime_error::runtime_error; };
struct EmptyEvaluatorException : std::runtime_error { using std::runtime_error::runtime_error; };
struct ModelLoadingException : std::runtime_error { using std::runtime_error::runtime_error; };   // job-fatal
struct WrongModelIdFormat : std::runtime_error { using std::runtime_error::runtime_error; };       // job-fatal

// ---------------------------------------------------------------------------------------- vectors
struct Vector {
  int size = 0;
  bool sparse = false;
  std::vector<double> data;     // dense: size values; sparse: one value per index
  std::vector<int> indices;     // sparse only
  static Vector dense(std::vector<double> v) { Vector r; r.size = (int)v.size(); r.data = std::move(v); return r; }
  static Vector sparseOf(int size, std::vector<int> idx, std::vector<double> val) {
    Vector r; r.size = size; r.sparse = true; r.indices = std::move(idx); r.data = std::move(val); return r;
  }
};

// ---------------------------------------------------------------------------------------- prediction ADT
struct Target {
  bool empty = true; double value = 0;
  static Target score(double v) { Target t; t.empty = false; t.value = v; return t; }
  double get() const { if (empty) throw std::out_of_range("EmptyScore.nan"); return value; }   // Target.scala:83
  double getOrElse(double dflt) const { return empty ? dflt : value; }                           // Target.scala:51-55
  bool operator==(const Target& o) const { return empty == o.empty && (empty || value == o.value); }
};
struct Prediction {
  Target value; uint8_t status = FJP_ST_NOMODEL;  // status: why it is empty (diagnostics; the reference logs instead)
  bool operator==(const Prediction& o) const { return value == o.value; }
};

// ---------------------------------------------------------------------------------------- ids and control messages
struct ModelId {
  std::string name; int64_t version = 0;
  static constexpr const char* separatorSymbol = "_";
  static ModelId fromIdentifier(const std::string& id);   // throws WrongModelIdFormat (ModelId.scala:36-40, 47-51)
  std::string identifier() const { return name + separatorSymbol + std::to_string(version); }
  int32_t hashCode() const;                               // (name + "_" + version).hashCode (ModelId.scala:62)
  bool operator<(const ModelId& o) const { return name != o.name ? name < o.name : version < o.version; }
  bool operator==(const ModelId& o) const { return name == o.name && version == o.version; }
};
int32_t javaStringHashCode(const std::string& s);
struct ModelInfo { std::string path; };

struct ServingMessage {
  enum Kind { Add, Del } kind; std::string name; int64_t version; std::string path; int64_t occurredOn;
  ModelId modelId() const { return ModelId{name, version}; }
  ModelInfo modelInfo() const { return ModelInfo{path}; }
};
inline ServingMessage AddMessage(std::string name, int64_t version, std::string path, int64_t occurredOn) { return {ServingMessage::Add, std::move(name), version, std::move(path), occurredOn}; }
inline ServingMessage DelMessage(std::string name, int64_t version, int64_t occurredOn) { return {ServingMessage::Del, std::move(name), version, "", occurredOn}; }

using Metadata = std::map<ModelId, ModelInfo>;   // immutable.Map[ModelId, ModelInfo] / CheckpointType.scala:29-32
struct MetadataManager {  // Add on an existing id is ignored; Del removes (MetadataManager.scala:59-69, 80-86)
  static Metadata apply(const ServingMessage& m, const Metadata& md);
};

// ---------------------------------------------------------------------------------------- reader + model
struct ModelReader {
  std::string sourcePath;
  explicit ModelReader(std::string p) : sourcePath(std::move(p)) {}
  std::string buildDistributedPath() const;   // whole file as a string; throws std::runtime_error
};

class Engine;  // thin RAII over fjp_engine

class PmmlModel {
 public:
  PmmlModel() = default;                                   // PmmlModel.empty (EmptyEvaluator)
  PmmlModel(fjp_engine* e, fjp_model* m) : engine_(e), model_(m, &fjp_model_release) {}
  static PmmlModel fromReader(fjp_engine* e, const ModelReader& reader);   // throws ModelLoadingException
  static PmmlModel empty() { return PmmlModel(); }
  bool isEmpty() const { return !model_; }
  std::string modelName() const;                           // throws EmptyEvaluatorException on the empty model
  int activeFields() const { return model_ ? fjp_model_num_fields(model_.get()) : 0; }
  // One event, synchronously (a micro-batch of one): the literal drop-in for S/api/PmmlModel.scala:109-119.
  Prediction predict(const Vector& v, std::optional<double> replaceNan = std::nullopt) const;
  fjp_model* handle() const { return model_.get(); }
  fjp_engine* engine() const { return engine_; }
 private:
  fjp_engine* engine_ = nullptr;
  std::shared_ptr<fjp_model> model_;
};

// Predictions waiting to be emitted, in arrival order: two flat arrays and a read cursor (an operator emits hundreds of millions
// of them; one heap object or one vector shuffle per event would cost more than scoring it).
class PredictionQueue {
 public:
  static constexpr uint8_t kOpen = 0xFF;   // placeholder: the prediction is still being computed; nothing behind it is handed out
  void push(uint8_t status, double score) { status_.push_back(status); score_.push_back(score); }
  size_t reserve() { status_.push_back(kOpen); score_.push_back(0); open_++; return status_.size() - 1; }   // stays valid until filled
  size_t reserve_n(size_t n) { const size_t s0 = status_.size(); status_.insert(status_.end(), n, kOpen); score_.insert(score_.end(), n, 0.0); open_ += n; return s0; }
  void fill(size_t slot, uint8_t status, double score) { status_[slot] = status; score_[slot] = score; open_--; }
  // predictions that can be emitted now: everything up to the first placeholder
  size_t size() const { return ready_end() - head_; }
  // copies up to `cap` predictions out (score NaN for an empty one); returns how many
  long pop(double* scores, uint8_t* status, long cap);
  std::vector<Prediction> drain();
 private:
  size_t ready_end() const;
  std::vector<double> score_; std::vector<uint8_t> status_; size_t head_ = 0, open_ = 0;
};

// ---------------------------------------------------------------------------------------- micro-batch buffer
// Rows of one model packed for fjp_score_batch / fjp_submit: pinned X (dense slots), present bits, replace values.
class MicroBatch {
 public:
  MicroBatch(int n_fields, int capacity);
  ~MicroBatch();
  MicroBatch(const MicroBatch&) = delete;
  bool full() const { return n_ >= cap_; }
  int fields() const { return F_; }
  int capacity() const { return cap_; }
  int size() const { return n_; }
  int append(const Vector& v, std::optional<double> replaceNan);   // returns the row index
  int append_dense(const double* x);                                // DenseVector payload, no replaceNan
  int append_dense_block(const double* X, int rows);                // `rows` DenseVector payloads, contiguous: one copy; returns the first row index
  double* tail() { return X_ + (size_t)n_ * F_; }                   // where the next row goes: a producer may write room() rows there itself ...
  int commit_rows(int rows);                                        // ... and then declare them (dense, no replaceNan); returns the first row index
  void clear() { n_ = 0; any_sparse_ = false; any_replace_ = false; }
  int room() const { return cap_ - n_; }
  int score(fjp_engine* e, fjp_model* m);                         // fills score()/status()
  int submit(fjp_engine* e, fjp_model* m, int* ticket);           // asynchronous: fjp_submit; results valid after fjp_wait(ticket)
  fjp_batch_desc desc(fjp_model* m) const { return fjp_batch_desc{m, X_, any_sparse_ ? present_ : nullptr, any_replace_ ? replace_ : nullptr, n_, F_, score_, status_}; }
  const double* scores() const { return score_; }
  const uint8_t* statuses() const { return status_; }
 private:
  // present words / replace values are written lazily: dense rows leave them alone until the first sparse row (or replaceNan)
  // of the batch makes the kernels look at them, at which point the rows before it are back-filled
  void need_present(); void need_replace();
  int F_, W_, cap_, n_ = 0; bool any_sparse_ = false, any_replace_ = false;
  double* X_; uint32_t* present_; double* replace_; double* score_; uint8_t* status_;
};

// ---------------------------------------------------------------------------------------- fixed-model operator
// stream.evaluate(reader)(f) / quickEvaluate(reader): S/package.scala:76-82, 138-142.
class EvaluationFunction {
 public:
  EvaluationFunction(fjp_engine* e, ModelReader reader, int maxBatch) : engine_(e), reader_(std::move(reader)), max_batch_(maxBatch) {}
  void open();                                             // loads the model; failure -> ModelLoadingException
  const PmmlModel& evaluator() const { return model_; }
  // buffer one event; predictions come back through drain() in arrival order
  void processElement(const Vector& v, std::optional<double> replaceNan = std::nullopt);
  void processDense(const double* x, int size);           // same for a DenseVector payload given in place
  void flush();
  std::vector<Prediction> drain() { return out_.drain(); }
  long poll(double* scores, uint8_t* status, long cap) { return out_.pop(scores, status, cap); }
  void processDenseBlock(const double* X, long n, int size);   // n DenseVector payloads in arrival order
  // The same without the operator's copy: the producer (a JNI shim holding a Java double[]) writes up to *granted DenseVector
  // payloads straight into the pinned micro-batch and commits them; nothing else may be called on the operator in between.
  double* reserve(int size, int want_rows, int* granted);      // nullptr: `size` is not the model's field count (use processDense)
  void commit(int rows);
  void close() { flush(); }
 private:
  // two pinned micro-batches: while the GPU scores one, the task thread packs the other (fjp_submit / fjp_wait)
  fjp_engine* engine_; ModelReader reader_; int max_batch_; PmmlModel model_;
  std::unique_ptr<MicroBatch> batch_[2]; int cur_ = 0; int ticket_[2] = {-1, -1}; PredictionQueue out_;
  std::vector<int> pending_[2];   // per buffered event: row in its batch, or -1 = size mismatch (EmptyScore)
  void submit_current();          // hand the current batch to the GPU, then collect the older one
  void collect(int b);            // wait for batch b and emit its predictions in arrival order
};

// ---------------------------------------------------------------------------------------- dynamic operator
// events.withSupportStream(ctrl).evaluate(f): S/package.scala:63-65, 107-119, S/api/functions/EvaluationCoFunction.scala:55-133.
class EvaluationCoFunction {
 public:
  EvaluationCoFunction(fjp_engine* e, int maxBatch) : engine_(e), max_batch_(maxBatch) {}
  ~EvaluationCoFunction();
  void initializeState(const std::vector<Metadata>& restored);   // union of the restored maps (:81-96)
  Metadata snapshotState();                                      // flushes, then returns the metadata (:76-79)
  void processElement1(const std::string& modelId, const Vector& v, std::optional<double> replaceNan = std::nullopt);
  void processDense1(const std::string& modelId, const double* x, int size);
  // n dense events in arrival order, event i addressed to ids[id_index[i]]; ids are resolved once per call and flush
  void processDenseBatch(const char* const* ids, const int32_t* id_index, const double* X, long n, int size);
  void processElement2(const ServingMessage& control);           // manageModels then manageMetadata (:69-74)
  void flush();                                                   // every buffered event scored and ready to poll (waits for parked ones)
  std::vector<Prediction> drain() { return out_.drain(); }
  long poll(double* scores, uint8_t* status, long cap) { return out_.pop(scores, status, cap); }
  PmmlModel fromMetadata(const std::string& modelId);            // (:106-117)
  PmmlModel loadModel(const std::string& path);                  // (:98-104)
  const Metadata& servingMetadata() const { return metadata_; }
  int servingModels() const { return engine_ ? fjp_table_size(engine_) : 0; }
  long parkedEvents() const { return parked_events_; }
  double flushSeconds() const { return flush_seconds_; }   // time the task thread spent handing batches to the GPU and waiting for them
 private:
  long parked_events_ = 0; double flush_seconds_ = 0;
  struct Pending { int bucket; int row; };
  struct Bucket { PmmlModel model; std::unique_ptr<MicroBatch> batch; };
  fjp_engine* engine_; int max_batch_; Metadata metadata_;
  std::vector<Bucket> buckets_; std::vector<Pending> pending_; PredictionQueue out_; uint64_t flushes_ = 0;
  // id string -> model bound to it right now; dropped on every control message and whenever the table changes
  std::unordered_map<std::string, PmmlModel> resolved_; uint64_t resolved_epoch_ = ~0ull;
  std::vector<std::unique_ptr<MicroBatch>> pool_;   // pinned micro-batches are recycled: cudaHostAlloc is far too slow for the event path
  // Add starts the model load in the background; its outcome (model or ModelLoadingException) is consumed by the first
  // event for that id, which is where the reference loads -- and fails -- too (T/RichConnectedStreamSpec.scala:288-300)
  std::map<ModelId, std::shared_future<PmmlModel>> loading_;
  // Bulk ingest (processDenseBatch) does not stop the stream for a model that is still loading: its events are PARKED -- rows in
  // pinned micro-batches of their own, a placeholder keeping their place in the output order -- and scored as soon as the load
  // has finished, while events of the other models keep flowing.  The single-event calls keep the reference's behaviour to the
  // letter (the first event waits for the load and is where a ModelLoadingException surfaces); for parked events the exception
  // surfaces from the next call that settles them (a later processDenseBatch, flush or snapshotState): job-fatal either way.
  struct Parked {
    ModelId id; std::shared_future<PmmlModel> fut; int size; bool detached = false;   // detached: the id was deleted meanwhile -- score, but do not cache
    std::vector<std::unique_ptr<MicroBatch>> batches; std::vector<size_t> slots;       // slots[i]: output position of the i-th parked row
  };
  std::vector<std::unique_ptr<Parked>> parked_;
  static constexpr int kMaxParkedBatches = 8;
  int park_index(const std::string& modelId, int size);   // >= 0: the events of this id are parked there right now
  void settle(Parked& p);                                 // load finished: score the parked rows, fill their placeholders
  void flush_buckets(bool wait_for_parked);
  // A flush hands all buckets to the GPU behind one ticket (fjp_submit_many) and returns to packing; the generation before it is
  // collected then (its predictions come first).  Placeholders keep the output order while a generation is in flight.
  struct Generation { std::vector<Bucket> buckets; std::vector<Pending> pending; size_t slot0 = 0; int ticket = -1; };
  std::unique_ptr<Generation> inflight_;
  void collect(Generation& g);
  std::unique_ptr<MicroBatch> take_batch(int fields);
  // a generation holds at most max_batch events over all its buckets; one model's bucket gets an eighth of that (a full bucket
  // flushes the generation), so two generations of 32 models stay at a few hundred MB of pinned memory
  int bucket_rows() const { return std::min(max_batch_, std::max(1024, max_batch_ / 8)); }
  int bucket_for(const PmmlModel& m);
  PmmlModel resolve(const std::string& modelId);   // servingModels.getOrElse(id.hashCode, fromMetadata(id))
  void enqueue(const PmmlModel& model, const Vector* v, const double* dense, int size, std::optional<double> replaceNan);
};

}  // namespace flinkjpmml
