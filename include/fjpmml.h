/* fjpmml.h -- C ABI of the B200 batched PMML scoring engine (libfjpmml.so).
 *
 * This is the boundary a JNI / Panama binding of flink-jpmml would call in place of JPMML.  Every entry point cites
 * the reference interface it stands in for; paths are relative to
 *   flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/   (abbreviated S/ below).
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ or torch types.
 *   - return value 0 = success, negative = engine-fatal (CUDA error, out of memory, bad argument); message in `err`.
 *   - per-event outcomes never fail a call: they are reported in the `status` byte, mirroring the reference's rule
 *     that PmmlModel.predict never throws (S/models/prediction/Prediction.scala:37-62):
 *         0  Score(value)                                   any non-zero value <=> EmptyScore
 *         1  vector size != number of active fields         InputValidationException   S/api/PmmlModel.scala:127-134
 *         2  input preparation failed                       InputPreparationException  S/api/pipeline/Pipeline.scala:49-54
 *         3  evaluation failed / null prediction            JPMML EvaluationException  S/models/prediction/Prediction.scala:55
 *         4  first target not extractable as a Double       JPMMLExtractionException   S/api/pipeline/Pipeline.scala:94-98
 *         5  no model (unknown / deleted id)                EmptyEvaluatorException    S/api/Evaluator.scala:69-77
 *   - there is no CPU scoring path: scoring entry points need a CUDA device and fail loudly without one.
 */
#ifndef FJPMML_H
#define FJPMML_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fjp_engine fjp_engine;
typedef struct fjp_model fjp_model;

#define FJP_ST_SCORE 0
#define FJP_ST_SIZE 1
#define FJP_ST_PREPARE 2
#define FJP_ST_EVALUATE 3
#define FJP_ST_EXTRACT 4
#define FJP_ST_NOMODEL 5

/* Library version string. */
const char* fjp_version(void);

/* ---- engine: one per Flink subtask / GPU --------------------------------------------------------------------
 * Replaces the per-operator state that S/api/functions/EvaluationFunction.scala:43 (lazy evaluator) and
 * S/api/functions/EvaluationCoFunction.scala:60-67 (servingModels cache) keep on the JVM heap.
 * Usable from several host threads; each call picks its own stream slot. */
int fjp_engine_create(int device, fjp_engine** out, char* err, size_t errcap);
void fjp_engine_destroy(fjp_engine* e);
int fjp_engine_device(const fjp_engine* e);
int fjp_engine_sync(fjp_engine* e);
/* Number of scoring kernels launched by this engine so far (diagnostics / bench evidence). */
long fjp_engine_launch_count(const fjp_engine* e);
/* Events scored through the host entry points so far, by status (index = status byte).  Replaces the one WARN log line
 * per failed prediction of S/models/prediction/Prediction.scala:47-62 with per-batch counters an operator can export
 * as Flink metrics. */
void fjp_engine_status_counts(const fjp_engine* e, uint64_t out[6]);
/* Diagnostics of the two fast paths that decide a class label only outside a proven error bound and hand every other event
 * to the kernels that follow JPMML's arithmetic order:
 *   out[0], out[1]  tensor-core path of dense layers (NeuralNetwork models, large batches): events taken / handed back
 *   out[2], out[3]  fused-multiply-add path of the dense tile kernel (softmax RegressionModel, k-means ClusteringModel)
 * Waits for the engine's streams.  The paths stand in for JPMML's evaluators behind S/api/PmmlModel.scala:159-160. */
int fjp_engine_fast_path_counts(fjp_engine* e, uint64_t out[4]);

/* ---- model load: PMML bytes in, device-resident flattened model out -----------------------------------------
 * Replaces PmmlModel.fromReader (S/api/PmmlModel.scala:53-61): JAXB unmarshal + ModelEvaluatorFactory.  The bytes
 * come from the caller (Flink's FileSystem stays in Flink, S/api/reader/FsReader.scala:41-49).  A non-zero return
 * carries the message for ModelLoadingException (S/api/exceptions/package.scala:42-43).
 * `e` may be NULL: the model is then compiled but not uploaded (inspection only; it cannot score). */
int fjp_model_compile(fjp_engine* e, const char* pmml, size_t len, fjp_model** out, char* err, size_t errcap);
void fjp_model_retain(fjp_model* m);
void fjp_model_release(fjp_model* m);
/* Copies a model that lives on another engine's GPU onto `dst`'s GPU with a peer copy (NVLink), without
 * re-compiling.  Mirrors the control stream being broadcast to every subtask (S/package.scala:63-65). */
int fjp_model_replicate(fjp_engine* dst, const fjp_model* src, fjp_model** out, char* err, size_t errcap);

/* Model facts the Scala side needs: getActiveFields.size / names (S/api/PmmlModel.scala:128,
 * S/api/converter/VectorConverter.scala:97-98) and modelName (S/api/PmmlModel.scala:90). */
int fjp_model_num_fields(const fjp_model* m);
const char* fjp_model_field_name(const fjp_model* m, int i);
const char* fjp_model_name(const fjp_model* m);
const char* fjp_model_family(const fjp_model* m);   /* "ClusteringModel", "TreeModel", "MiningModel", ... */
int fjp_model_num_labels(const fjp_model* m);       /* class labels / cluster ids */
const char* fjp_model_label(const fjp_model* m, int i);
int fjp_model_num_segments(const fjp_model* m);     /* trees of an ensemble (0 otherwise) */
/* One-line description of the compiled layout (format, key encoding, depth, bytes). */
const char* fjp_model_describe(const fjp_model* m);
size_t fjp_model_device_bytes(const fjp_model* m);

/* ---- device model table keyed by ModelId(name, version) ---------------------------------------------------------
 * Replaces servingModels: WeakHashMap[Int, PmmlModel] (S/api/functions/EvaluationCoFunction.scala:66-67), keyed on the
 * full id instead of its 32-bit hashCode (S/models/core/ModelId.scala:62).  Readers never block: the table is an
 * immutable snapshot swapped on every change; a deleted model's memory is released when the last in-flight batch
 * that uses it has finished.
 * fjp_table_add: 0 = added, 1 = id already present, ignored (S/api/managers/MetadataManager.scala:59-69).
 * fjp_table_del: 0 = removed, 1 = id was not present (S/api/managers/MetadataManager.scala:80-86,
 *                S/api/managers/ModelsManager.scala:51-57).
 * fjp_table_get: retained handle or NULL (the caller releases it). */
int fjp_table_add(fjp_engine* e, const char* name, int64_t version, fjp_model* m);
int fjp_table_del(fjp_engine* e, const char* name, int64_t version);
fjp_model* fjp_table_get(fjp_engine* e, const char* name, int64_t version);
int fjp_table_size(const fjp_engine* e);
uint64_t fjp_table_epoch(const fjp_engine* e);

/* ---- pinned host buffers for micro-batches ------------------------------------------------------------------- */
void* fjp_host_alloc(size_t bytes);
void fjp_host_free(void* p);

/* ---- scoring -------------------------------------------------------------------------------------------------
 * Replaces, for n events at once, PmmlModel.predict (S/api/PmmlModel.scala:109-119).
 *   X        n x vec_size doubles, row-major: the Vector payloads (dense values; for a SparseVector the absent slots
 *            may hold anything)                                        S/api/converter/VectorConverter.scala:68-86
 *   present  n x ceil(vec_size/32) words, bit i = slot i is present; NULL = all present (DenseVector).  A cleared bit
 *            is the SparseVector's absent index, i.e. a missing value  S/api/converter/VectorConverter.scala:100-103
 *   replace  n doubles, NaN = None; NULL = None for all rows (replaceNan) S/api/PmmlModel.scala:148
 *   score    n doubles out (NaN unless status == 0);  status n bytes out;  aux n int32 out or NULL:
 *            1-based cluster position / node id of a single tree / index of the winning label.
 * `m` == NULL scores against the empty model: every status is 5 (S/api/PmmlModel.scala:67).
 * Host-pointer variant: synchronous; copies are pipelined with the kernels in chunks. */
int fjp_score_batch(fjp_engine* e, fjp_model* m, const double* X, const uint32_t* present, const double* replace,
                    long n, int vec_size, double* score, uint8_t* status, int32_t* aux);
/* Same, with the verification outputs: node_ids n x node_stride (PMML node id reached in every segment),
 * probs n x prob_stride (class probabilities, or the winning distance for clustering).  Either may be NULL. */
int fjp_score_batch_ex(fjp_engine* e, fjp_model* m, const double* X, const uint32_t* present, const double* replace,
                       long n, int vec_size, double* score, uint8_t* status, int32_t* aux,
                       int32_t* node_ids, int node_stride, double* probs, int prob_stride);
/* Float rows: half the bytes over PCIe for models whose tests never need more than one float per value.
 * A tree ensemble compiled with the float key encoding looks at a double feature only through RU(x) (all its tests on that
 * field are `x <= t`) or RD(x) (`x < t`) with float thresholds t, and at a float-typed field through (float)x, which is what
 * JPMML's cast to the field's dataType does as well.  fjp_model_float_row_modes fills, per vector slot, how the binding must
 * turn the Vector's double into the float it ships: 0 = plain cast (float field), 1 = round up, 2 = round down; it returns
 * vec_size, or 0 when the model needs the doubles (modes may be NULL to just ask).  Rows prepared that way score bit-identically
 * to the double rows through fjp_score_batch (S/api/converter/VectorConverter.scala:68-86 is where the reference copies the
 * Vector's values; here the copy narrows them).  Same conventions as fjp_score_batch otherwise; no replaceNan. */
int fjp_model_float_row_modes(const fjp_model* m, uint8_t* modes, int cap);
int fjp_score_batch_f32(fjp_engine* e, fjp_model* m, const float* X, const uint32_t* present, long n, int vec_size,
                        double* score, uint8_t* status, int32_t* aux);
/* SparseVector rows in CSR form: row i holds indices[row_ptr[i] .. row_ptr[i+1]) with their values; every index that is
 * not listed is a missing value, an explicit 0.0 is present (S/api/converter/VectorConverter.scala:80-86, 100-103).  The rows
 * are expanded on the device (O(nnz), where the reference's toDenseData is O(size * nnz), :100-103); only the non-zeros
 * cross PCIe.  Indices outside [0, vec_size) are ignored, as in the reference.  Host pointers, synchronous. */
int fjp_score_sparse(fjp_engine* e, fjp_model* m, const int64_t* row_ptr, const int32_t* indices, const double* values,
                     const double* replace, long n, int vec_size, double* score, uint8_t* status, int32_t* aux);
/* Sparse scoring that allocates nothing.  [No counterpart in the reference.]
 * For callers that may not touch the device allocator while they score: a subtask that shares its GPU with other jobs
 * under a fixed memory budget (a cudaMalloc may fail, or stall the whole device), or one on a latency-bound path.  The
 * other host-pointer entry points grow their staging arenas with cudaMalloc / cudaFree on demand; this one never does.
 *   - every pointer is a DEVICE pointer owned by the caller: CSR inputs (row_ptr n+1 entries, indices, values as for
 *     fjp_score_sparse) and the outputs score[n], status[n];
 *   - the dense rows the CSR input is expanded into live in an area of FJP_NOALLOC_SCRATCH_BYTES built into the library
 *     (present on every GPU that loads it, until exit).  The call is refused, with a message in fjp_last_error(), when
 *     n * (vec_size*8 + ceil(vec_size/32)*4) exceeds it, when vec_size is not the model's field count, or without a model;
 *   - the area is shared by all calls and keeps the previous call's contents (nothing needs clearing: each call rewrites
 *     what it reads).  Calls whose GPU work overlaps -- two streams, two threads -- corrupt each other: serialise them;
 *   - the first call on a device also looks up the area's device address (once per device), and the first call with a
 *     model of a new family loads that family's kernels (the CUDA runtime loads code lazily, which takes device memory):
 *     make one warm-up call per model at start-up, before any stream capture on that device;
 *   - asynchronous on `cuda_stream` (a cudaStream_t; NULL = the default stream). */
#define FJP_NOALLOC_SCRATCH_BYTES (8u << 20)
int fjp_score_sparse_noalloc(fjp_engine* e, fjp_model* m, const int64_t* d_row_ptr, const int32_t* d_indices,
                             const double* d_values, long n, int vec_size, double* d_score, uint8_t* d_status,
                             void* cuda_stream);
/* Device-pointer variant: asynchronous on `cuda_stream` (a cudaStream_t, NULL = the engine's stream).  Models with derived
 * fields, and tree ensembles scoring a batch large enough for the two-kernel throughput path, take a work area from the
 * device's stream-ordered allocator (cudaMallocAsync / cudaFreeAsync on that stream; fjp_engine_create raises the default
 * pool's release threshold so the block is reused from call to call). */
int fjp_score_device(fjp_engine* e, fjp_model* m, const double* dX, const uint32_t* dpresent, const double* dreplace,
                     long n, int vec_size, double* dscore, uint8_t* dstatus, int32_t* daux, void* cuda_stream);
/* Double-buffered host variant: submit returns a ticket at once (the buffers must stay valid and should be
 * pinned); wait blocks until that batch's results are in `score`/`status`/`aux`.  A caller may keep up to
 * fjp_max_inflight() tickets open.  fjp_submit never blocks: the engine's stream slots are shared by all host threads,
 * and when none is free it returns FJP_BUSY with *ticket = -1 -- fjp_wait your oldest open ticket (or, holding none,
 * yield) and submit again.  On every failure *ticket is -1 and nothing is left to wait for. */
#define FJP_BUSY 1
int fjp_submit(fjp_engine* e, fjp_model* m, const double* X, const uint32_t* present, const double* replace,
               long n, int vec_size, double* score, uint8_t* status, int32_t* aux, int* ticket);
/* Several micro-batches, each for its own model, behind one ticket (one stream slot, copies and kernels enqueued back to back):
 * what a dynamic operator hands over per flush -- one bucket per model id (S/api/functions/EvaluationCoFunction.scala:106-117 picks
 * the model per event) -- so that its task thread packs the next events while these are scored.  Every batch must match its
 * model's field count; FJP_BUSY as for fjp_submit; fjp_wait(ticket) completes all of them. */
typedef struct fjp_batch_desc {
  fjp_model* model; const double* X; const uint32_t* present; const double* replace; long n; int vec_size; double* score; uint8_t* status;
} fjp_batch_desc;
int fjp_submit_many(fjp_engine* e, const fjp_batch_desc* batches, int count, int* ticket);
int fjp_wait(fjp_engine* e, int ticket);
int fjp_max_inflight(void);

/* Message of the last engine-fatal error on this thread. */
const char* fjp_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* FJPMML_H */
