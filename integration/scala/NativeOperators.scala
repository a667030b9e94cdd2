/*
 * SOURCE ONLY (no JVM in the build image).  Batching versions of the two operators of
 *   flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/functions/EvaluationFunction.scala:39-50
 *   flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/functions/EvaluationCoFunction.scala:55-133
 * on top of Fjpmml.  Public names and the user-facing DSL (package.scala:58-143) stay as they are; the compiled and
 * tested statement of this logic is flink-jpmml_b200/csrc/host/flinkjpmml.hpp + tests/test_gpu_operators.py.
 */
package io.radicalbit.flink.pmml.scala.native

import io.radicalbit.flink.pmml.scala.api.exceptions.ModelLoadingException
import io.radicalbit.flink.pmml.scala.api.reader.ModelReader
import io.radicalbit.flink.pmml.scala.models.control.{AddMessage, DelMessage, ServingMessage}
import io.radicalbit.flink.pmml.scala.models.core.{ModelId, ModelInfo}
import io.radicalbit.flink.pmml.scala.models.input.BaseEvent
import io.radicalbit.flink.pmml.scala.models.prediction.{EmptyScore, Prediction, Score}
import org.apache.flink.api.common.functions.RichFlatMapFunction
import org.apache.flink.api.common.state.{ListState, ListStateDescriptor}
import org.apache.flink.api.common.typeinfo.{TypeHint, TypeInformation}
import org.apache.flink.configuration.Configuration
import org.apache.flink.ml.math.Vector
import org.apache.flink.runtime.state.{FunctionInitializationContext, FunctionSnapshotContext}
import org.apache.flink.streaming.api.checkpoint.CheckpointedFunction
import org.apache.flink.streaming.api.functions.co.CoProcessFunction
import org.apache.flink.util.Collector

import scala.collection.JavaConverters._
import scala.collection.mutable
import scala.concurrent.{Await, ExecutionContext, Future}
import scala.concurrent.duration.Duration
import scala.util.{Failure, Success, Try}

/** What PmmlModel holds instead of a JPMML evaluator: a compiled, device-resident model (0 = the empty model). */
final class NativeModel(val engine: Long, val handle: Long) {
  lazy val fields: Int = if (handle == 0L) 0 else Fjpmml.modelNumFields(handle)
  def isEmpty: Boolean = handle == 0L
}

object NativeModel {
  def load(engine: Long, reader: ModelReader): NativeModel =
    Try(Fjpmml.modelCompile(engine, reader.buildDistributedPath.getBytes("UTF-8"))) match {
      case Success(h) => new NativeModel(engine, h)
      case Failure(e) => throw new ModelLoadingException(e.getMessage, e)
    }

  def prediction(score: Double, status: Byte): Prediction =
    if (status == Fjpmml.StatusScore) Prediction(Score(score)) else Prediction(EmptyScore)
}

/** Fixed model: two pinned micro-batches alternate; predictions are emitted in arrival order when a batch completes.
  * `emit` receives the event together with its prediction (quickEvaluate's `(Prediction, V)` or a user function). */
abstract class BatchedEvaluationFunction[IN, OUT](reader: ModelReader, maxBatch: Int = 16384)
    extends RichFlatMapFunction[IN, OUT]
    with CheckpointedFunction {

  protected def vectorOf(event: IN): Vector
  protected def emit(event: IN, prediction: Prediction): OUT

  @transient private var engine: Long = 0L
  @transient private var model: NativeModel = _
  @transient private var batches: Array[MicroBatch] = _
  @transient private var events: Array[mutable.ArrayBuffer[(IN, Int)]] = _ // (event, row or -1 for a size mismatch)
  @transient private var tickets: Array[Int] = _
  @transient private var current: Int = 0
  @transient private var lastOut: Collector[OUT] = _ // the operator's collector: the same object for every flatMap call

  override def open(parameters: Configuration): Unit = {
    engine = Fjpmml.engineCreate(getRuntimeContext.getIndexOfThisSubtask % gpusPerHost)
    model = NativeModel.load(engine, reader) // ModelLoadingException fails the job here, as today
    batches = Array.fill(2)(new MicroBatch(model.fields, maxBatch))
    events = Array.fill(2)(mutable.ArrayBuffer.empty[(IN, Int)])
    tickets = Array(-1, -1)
  }

  protected def gpusPerHost: Int = 8

  override def flatMap(value: IN, out: Collector[OUT]): Unit = {
    lastOut = out
    val v = vectorOf(value)
    val row = if (v.size != model.fields) -1 else batches(current).append(v, None)
    events(current) += ((value, row))
    if (batches(current).isFull || events(current).size >= maxBatch) handOver(out)
  }

  private def handOver(out: Collector[OUT]): Unit = {
    val b = current
    if (batches(b).rows > 0) {
      def trySubmit(): Int = Fjpmml.submit(engine, model.handle, batches(b).x, batches(b).present, batches(b).replace,
                                           batches(b).rows, model.fields, batches(b).score, batches(b).status)
      tickets(b) = trySubmit()
      while (tickets(b) < 0) { // FJP_BUSY: give back the older batch's slot (results stay in its buffers), else yield
        if (tickets(1 - b) >= 0) { Fjpmml.await(engine, tickets(1 - b)); tickets(1 - b) = -1 } else Thread.`yield`()
        tickets(b) = trySubmit()
      }
    }
    collect(1 - b, out) // the older batch first: keeps arrival order
    current = 1 - b
  }

  private def collect(b: Int, out: Collector[OUT]): Unit = {
    if (tickets(b) >= 0) { Fjpmml.await(engine, tickets(b)); tickets(b) = -1 }
    events(b).foreach {
      case (event, row) =>
        val p = if (row < 0) Prediction(EmptyScore)
        else NativeModel.prediction(batches(b).score.getDouble(row * 8), batches(b).status.get(row))
        out.collect(emit(event, p))
    }
    events(b).clear(); batches(b).clear()
  }

  /** A checkpoint barrier must not overtake buffered events: everything that arrived before the barrier is scored and
    * emitted before the snapshot is taken, so the operator holds no in-flight events at the barrier and needs no state
    * of its own (the reference's contract: state is consistent at snapshotState,
    * S/api/functions/EvaluationCoFunction.scala:76-79).  Flink calls snapshotState on the task thread, between two
    * flatMap calls, so emitting through the operator's collector here is the same as emitting from flatMap. */
  override def snapshotState(context: FunctionSnapshotContext): Unit = drain()

  /** Scores and emits both micro-batches, oldest first (arrival order). */
  private def drain(): Unit =
    if (lastOut != null) {
      if (events(current).nonEmpty) handOver(lastOut) // submits the batch being packed, collects the older one
      collect(1 - current, lastOut)                   // ... and the one just submitted
    }

  override def initializeState(context: FunctionInitializationContext): Unit = ()

  override def close(): Unit = {
    if (batches != null) { drain(); batches.foreach(_.release()) }
    if (model != null && !model.isEmpty) Fjpmml.modelRelease(model.handle)
    if (engine != 0L) Fjpmml.engineDestroy(engine)
  }
}

/** Dynamic models: metadata map + device model table.  An event is bound, on arrival, to the model its id names at that
  * moment, and buffered in that model's bucket; control messages therefore act between events without a flush. */
abstract class BatchedEvaluationCoFunction[EVENT <: BaseEvent, CTRL <: ServingMessage, OUT](maxBatch: Int = 16384)
    extends CoProcessFunction[EVENT, CTRL, OUT]
    with CheckpointedFunction {

  protected def vectorOf(event: EVENT): Vector
  protected def emit(event: EVENT, prediction: Prediction): OUT

  @transient private var engine: Long = 0L
  @transient protected var servingMetadata: Map[ModelId, ModelInfo] = Map.empty // checkpointed exactly as today
  @transient private var loading: mutable.Map[ModelId, Future[NativeModel]] = _
  @transient private var buckets: mutable.LinkedHashMap[Long, MicroBatch] = _
  @transient private var pending: mutable.ArrayBuffer[(EVENT, Long, Int)] = _ // (event, model handle or 0, row / -1)
  @transient private var lastOut: Collector[OUT] = _
  @transient private var snapshotMetadata: ListState[java.util.HashMap[ModelId, ModelInfo]] = _

  /** Buffered events are scored and emitted before the metadata map is written, exactly where the reference writes it
    * (S/api/functions/EvaluationCoFunction.scala:76-79): no event is in flight at a checkpoint barrier. */
  override def snapshotState(context: FunctionSnapshotContext): Unit = {
    if (lastOut != null && pending.nonEmpty) flush(lastOut)
    snapshotMetadata.clear()
    snapshotMetadata.add(new java.util.HashMap[ModelId, ModelInfo](servingMetadata.asJava))
  }

  /** Union list state, as in the reference (:81-96); models are re-loaded lazily on the first event for their id. */
  override def initializeState(context: FunctionInitializationContext): Unit = {
    servingMetadata = Map.empty
    snapshotMetadata = context.getOperatorStateStore.getUnionListState(
      new ListStateDescriptor("metadata-snapshot", TypeInformation.of(new TypeHint[java.util.HashMap[ModelId, ModelInfo]]() {})))
    if (context.isRestored) snapshotMetadata.get().asScala.foreach(m => servingMetadata ++= m.asScala)
  }

  override def open(parameters: Configuration): Unit = {
    engine = Fjpmml.engineCreate(getRuntimeContext.getIndexOfThisSubtask % 8)
    loading = mutable.Map.empty; buckets = mutable.LinkedHashMap.empty; pending = mutable.ArrayBuffer.empty
  }

  override def processElement1(event: EVENT, ctx: CoProcessFunction[EVENT, CTRL, OUT]#Context, out: Collector[OUT]): Unit = {
    lastOut = out
    val id = ModelId.fromIdentifier(event.modelId) // WrongModelIdFormat is fatal, as in the reference
    var handle = Fjpmml.tableGet(engine, id.name, id.version)
    if (handle == 0L && servingMetadata.contains(id)) {
      // the load was started at the Add; its failure surfaces here, on the first event (RichConnectedStreamSpec:288-300)
      val model = loading.remove(id) match {
        case Some(f) => Await.result(f, Duration.Inf)
        case None    => NativeModel.load(engine, ModelReader(servingMetadata(id).path))
      }
      Fjpmml.tableAdd(engine, id.name, id.version, model.handle)
      handle = model.handle
    }
    val v = vectorOf(event)
    if (handle == 0L || v.size != Fjpmml.modelNumFields(handle)) pending += ((event, 0L, -1))
    else {
      val bucket = buckets.getOrElseUpdate(handle, new MicroBatch(v.size, maxBatch))
      pending += ((event, handle, bucket.append(v, None)))
      if (bucket.isFull) flush(out)
    }
    if (pending.size >= maxBatch) flush(out)
  }

  override def processElement2(control: CTRL, ctx: CoProcessFunction[EVENT, CTRL, OUT]#Context, out: Collector[OUT]): Unit =
    control match {
      case add: AddMessage =>
        if (!servingMetadata.contains(add.modelId)) {
          servingMetadata += add.modelId -> add.modelInfo
          loading(add.modelId) = Future(NativeModel.load(engine, ModelReader(add.path)))(ExecutionContext.global)
        } // Add on a known id is ignored (MetadataManager.scala:59-69)
      case del: DelMessage =>
        Fjpmml.tableDel(engine, del.name, del.version) // buffered events keep their reference; HBM is freed when they drain
        loading.remove(del.modelId)
        servingMetadata -= del.modelId
    }

  protected def flush(out: Collector[OUT]): Unit = {
    val open = mutable.Queue.empty[Int] // own tickets, oldest first
    buckets.foreach {
      case (handle, b) if b.rows > 0 =>
        var t = Fjpmml.submit(engine, handle, b.x, b.present, b.replace, b.rows, b.fields, b.score, b.status)
        while (t < 0) { // FJP_BUSY
          if (open.nonEmpty) Fjpmml.await(engine, open.dequeue()) else Thread.`yield`()
          t = Fjpmml.submit(engine, handle, b.x, b.present, b.replace, b.rows, b.fields, b.score, b.status)
        }
        open.enqueue(t)
      case _ =>
    }
    open.foreach(Fjpmml.await(engine, _))
    pending.foreach {
      case (event, handle, row) =>
        val p = if (row < 0) Prediction(EmptyScore)
        else { val b = buckets(handle); NativeModel.prediction(b.score.getDouble(row * 8), b.status.get(row)) }
        out.collect(emit(event, p))
    }
    pending.clear(); buckets.values.foreach(_.clear())
  }
}
