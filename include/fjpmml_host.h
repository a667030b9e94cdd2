/* fjpmml_host.h -- C entry points of the host-side mirror of the reference's operators (libfjpmml.so).
 *
 * The reference's operators are Scala classes run by Flink; there is no JVM in this image, so their behaviour is
 * restated in C++ (flink-jpmml_b200/csrc/host/flinkjpmml.hpp) above the scoring ABI (fjpmml.h) and exposed here so
 * that the reference's operator tests can be replayed from pytest.  A Scala/JNI shim would not bind these: it would
 * keep its own EvaluationFunction / EvaluationCoFunction and call fjpmml.h.
 * S/ = flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/
 */
#ifndef FJPMML_HOST_H
#define FJPMML_HOST_H

#include "fjpmml.h"

#ifdef __cplusplus
extern "C" {
#endif

/* return codes: the two job-fatal exceptions of the reference (S/api/exceptions/package.scala:42-43, 53) */
#define FJH_OK 0
#define FJH_ENGINE_ERROR (-1)
#define FJH_WRONG_MODEL_ID_FORMAT (-2)   /* WrongModelIdFormat  */
#define FJH_MODEL_LOADING (-3)           /* ModelLoadingException */

typedef struct fjh_eval_function fjh_eval_function;
typedef struct fjh_cofunction fjh_cofunction;

/* ModelId.fromIdentifier / ModelId.hashCode  (S/models/core/ModelId.scala:36-40, 47-51, 62). */
int fjh_model_id_parse(const char* id, char* name_out, size_t name_cap, int64_t* version_out);
int32_t fjh_model_id_hash(const char* name, int64_t version);

/* PmmlModel.predict for ONE event, synchronously (S/api/PmmlModel.scala:109-119).  indices == NULL: DenseVector of
 * `size` values; otherwise SparseVector(size, indices[nnz], values[nnz]).  replace_nan: NaN = None. */
int fjh_predict(fjp_engine* e, fjp_model* m, const double* values, int size, const int* indices, int nnz, double replace_nan,
                double* score, uint8_t* status);

/* ---- fixed-model operator: stream.evaluate(reader)(f) / quickEvaluate(reader)   (S/package.scala:76-82, 138-142,
 * S/api/functions/EvaluationFunction.scala:39-50).  open() loads the model: NULL + message = ModelLoadingException
 * (the job fails, T/RichDataStreamSpec.scala:82-89).  Events are buffered into a pinned micro-batch of `max_batch`
 * rows; predictions are emitted in arrival order when the batch fills or on flush/close. */
fjh_eval_function* fjh_eval_function_open(fjp_engine* e, const char* path, int max_batch, char* err, size_t errcap);
const char* fjh_eval_function_model_name(fjh_eval_function* op);
int fjh_eval_function_process(fjh_eval_function* op, const double* values, int size, const int* indices, int nnz, double replace_nan);
/* n DenseVector events of `size` values each, in arrival order (the task thread's inner loop without per-event FFI). */
int fjh_eval_function_process_dense(fjh_eval_function* op, const double* X, long n, int size);
/* The same without the operator's copy (one host core moves about 12 GB/s, which is what bounds process_dense): the caller
 * -- a JNI shim holding the event's double[] -- writes up to *granted_rows DenseVector payloads of `size` doubles each
 * straight into the pinned micro-batch at the returned address (GetDoubleArrayRegion into it) and commits how many it
 * wrote.  NULL (granted 0): `size` is not the model's field count, fall back to fjh_eval_function_process (EmptyScore,
 * S/api/PmmlModel.scala:136-141).  No other call on the operator between reserve and commit. */
double* fjh_eval_function_reserve(fjh_eval_function* op, int size, int want_rows, int* granted_rows);
int fjh_eval_function_commit(fjh_eval_function* op, int rows);
int fjh_eval_function_flush(fjh_eval_function* op);
long fjh_eval_function_poll(fjh_eval_function* op, double* scores, uint8_t* status, long cap);
void fjh_eval_function_close(fjh_eval_function* op);

/* ---- dynamic operator: events.withSupportStream(ctrl).evaluate(f)   (S/package.scala:63-65, 107-119,
 * S/api/functions/EvaluationCoFunction.scala:55-133).
 * process_event  = processElement1: binds the event to the model its id names at this moment (cache, else metadata
 *                  -> lazy load, else the empty model), buffers it.  FJH_WRONG_MODEL_ID_FORMAT / FJH_MODEL_LOADING are
 *                  the job-fatal outcomes (T/api/functions/EvaluationCoFunctionSpec.scala:132-144,
 *                  T/RichConnectedStreamSpec.scala:288-300).
 * process_control = processElement2 (:69-74): Del evicts the cached model, then the metadata map is updated
 *                  (Add on a known id is ignored, S/api/managers/MetadataManager.scala:59-69).
 * snapshot / restore = snapshotState (:76-79) / initializeState (:81-96): the metadata map only; models reload lazily
 *                  (loadModel :98-104, fromMetadata :106-117). */
fjh_cofunction* fjh_cofunction_create(fjp_engine* e, int max_batch);
int fjh_cofunction_process_event(fjh_cofunction* op, const char* model_id, const double* values, int size, const int* indices, int nnz,
                                 double replace_nan, char* err, size_t errcap);
/* n DenseVector events in arrival order; event i carries the model id string ids[id_index[i]].
 * Unlike process_event, this call does not stop the stream for a model whose background load (started by the Add) has not
 * finished: such events are parked, keep their place in the output order (poll hands out predictions only up to the first
 * of them) and are scored when the load completes; a failed load surfaces as FJH_MODEL_LOADING from the call that settles
 * them (a later process_dense, flush or snapshot) -- job-fatal like the reference's first-event failure
 * (T/RichConnectedStreamSpec.scala:288-300). */
int fjh_cofunction_process_dense(fjh_cofunction* op, const char* const* ids, const int32_t* id_index, const double* X, long n, int size,
                                 char* err, size_t errcap);
int fjh_cofunction_process_control(fjh_cofunction* op, int is_add, const char* name, int64_t version, const char* path, int64_t occurred_on);
int fjh_cofunction_flush(fjh_cofunction* op, char* err, size_t errcap);
long fjh_cofunction_poll(fjh_cofunction* op, double* scores, uint8_t* status, long cap);
long fjh_cofunction_snapshot(fjh_cofunction* op, char* buf, long cap);
int fjh_cofunction_restore(fjh_cofunction* op, const char* const* snapshots, int n_snapshots);
int fjh_cofunction_metadata_size(fjh_cofunction* op);
int fjh_cofunction_cached_models(fjh_cofunction* op);
long fjh_cofunction_parked_events(fjh_cofunction* op);   /* events that were parked behind a model load so far (statistic) */
double fjh_cofunction_flush_seconds(fjh_cofunction* op); /* seconds the task thread spent submitting micro-batches and waiting for them (statistic) */
void fjh_cofunction_destroy(fjh_cofunction* op);

#ifdef __cplusplus
}
#endif
#endif /* FJPMML_HOST_H */
