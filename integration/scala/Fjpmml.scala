/*
 * SOURCE ONLY -- there is no JVM in the build image, so nothing under integration/ is compiled or tested here.
 * It is the reference-side half of the boundary described in INTEGRATION.md: what a flink-jpmml maintainer would add
 * to bind libfjpmml.so.  The same logic exists, compiled and tested, in C++ (flink-jpmml_b200/csrc/host/flinkjpmml.hpp).
 */
package io.radicalbit.flink.pmml.scala.native

import java.nio.{ByteBuffer, ByteOrder}

/** One `@native` method per entry point of include/fjpmml.h.  Handles are the C pointers carried as `Long`. */
object Fjpmml {
  System.loadLibrary("fjpmml_jni")

  @native def engineCreate(device: Int): Long
  @native def engineDestroy(engine: Long): Unit
  @native def statusCounts(engine: Long): Array[Long] // fjp_engine_status_counts: Flink metrics instead of WARN lines

  /** fjp_model_compile: throws ModelLoadingException(message) on a non-zero return. */
  @native def modelCompile(engine: Long, pmml: Array[Byte]): Long
  @native def modelRelease(model: Long): Unit
  @native def modelNumFields(model: Long): Int
  @native def modelFieldName(model: Long, index: Int): String
  @native def modelName(model: Long): String

  @native def tableAdd(engine: Long, name: String, version: Long, model: Long): Int
  @native def tableDel(engine: Long, name: String, version: Long): Int
  @native def tableGet(engine: Long, name: String, version: Long): Long

  /** fjp_host_alloc wrapped with NewDirectByteBuffer: pinned memory the GPU copies from / to directly. */
  @native def hostAlloc(bytes: Long): ByteBuffer
  @native def hostFree(buffer: ByteBuffer): Unit

  /** fjp_submit: returns the ticket; buffers must stay untouched until `await(ticket)`.  -1 = FJP_BUSY: the engine's stream
    * slots are all taken (they are shared by the subtasks of this JVM) -- `await` an own ticket, or yield, and submit again. */
  @native def submit(engine: Long,
                     model: Long,
                     x: ByteBuffer,
                     present: ByteBuffer,
                     replace: ByteBuffer,
                     rows: Long,
                     vecSize: Int,
                     score: ByteBuffer,
                     status: ByteBuffer): Int
  @native def await(engine: Long, ticket: Int): Unit

  val StatusScore: Byte = 0
}

/** A pinned micro-batch of rows for one model: dense slots, present bits, per-row replaceNan, results. */
final class MicroBatch(val fields: Int, val capacity: Int) {
  private val words = (fields + 31) / 32
  private def pinned(bytes: Long) = Fjpmml.hostAlloc(math.max(bytes, 8L)).order(ByteOrder.nativeOrder())

  val x: ByteBuffer = pinned(capacity.toLong * fields * 8)
  val present: ByteBuffer = pinned(capacity.toLong * words * 4)
  val replace: ByteBuffer = pinned(capacity.toLong * 8)
  val score: ByteBuffer = pinned(capacity.toLong * 8)
  val status: ByteBuffer = pinned(capacity.toLong)
  var rows: Int = 0

  def isFull: Boolean = rows >= capacity

  /** DenseVector: every slot present.  SparseVector: only the listed indices (an explicit 0.0 is present). */
  def append(vector: org.apache.flink.ml.math.Vector, replaceNan: Option[Double]): Int = {
    val row = rows
    val xBase = row.toLong * fields * 8
    val pBase = row.toLong * words * 4
    vector match {
      case dense: org.apache.flink.ml.math.DenseVector =>
        var i = 0
        while (i < fields) { x.putDouble((xBase + i * 8L).toInt, dense.data(i)); i += 1 }
        var w = 0
        while (w < words) { present.putInt((pBase + w * 4L).toInt, -1); w += 1 }
      case sparse: org.apache.flink.ml.math.SparseVector =>
        var i = 0
        while (i < fields) { x.putDouble((xBase + i * 8L).toInt, 0.0); i += 1 }
        var w = 0
        while (w < words) { present.putInt((pBase + w * 4L).toInt, 0); w += 1 }
        var k = 0
        while (k < sparse.indices.length) {
          val slot = sparse.indices(k)
          if (slot >= 0 && slot < fields) {
            x.putDouble((xBase + slot * 8L).toInt, sparse.data(k))
            val at = (pBase + (slot >> 5) * 4L).toInt
            present.putInt(at, present.getInt(at) | (1 << (slot & 31)))
          }
          k += 1
        }
    }
    replace.putDouble(row * 8, replaceNan.getOrElse(Double.NaN))
    rows += 1
    row
  }

  def clear(): Unit = rows = 0

  def release(): Unit = Seq(x, present, replace, score, status).foreach(Fjpmml.hostFree)
}
