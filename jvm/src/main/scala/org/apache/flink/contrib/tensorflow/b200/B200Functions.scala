package org.apache.flink.contrib.tensorflow.b200

import java.nio.ByteBuffer

import org.apache.flink.api.common.functions.{RichFlatMapFunction, RichFunction, RichMapFunction}
import org.apache.flink.configuration.Configuration
import org.apache.flink.contrib.tensorflow.models.{Model, RichModel}
import org.apache.flink.runtime.state.{FunctionInitializationContext, FunctionSnapshotContext}
import org.apache.flink.streaming.api.checkpoint.CheckpointedFunction
import org.apache.flink.util.Collector

/**
  * `ModelAwareFunction` (FT/common/functions/util/ModelAwareFunction.scala:10-19) for device-bound models.
  *
  * The reference's mix-in never looks at the runtime context (its `open(parameters)` ignores its argument, :15); a GPU
  * model has to know which device it serves, so this one reads `getIndexOfThisSubtask` and binds the model to
  * GPU `index % gpus` before opening it.  `gpus` comes from the job configuration key "ftf.gpus" (default 8, one box).
  */
trait B200ModelAwareFunction {
  this: RichFunction =>

  def model: Model[_]

  override def open(parameters: Configuration): Unit = {
    model match {
      case bound: DeviceBound => bound.bind(getRuntimeContext.getIndexOfThisSubtask, parameters.getInteger("ftf.gpus", 8))
      case _ =>
    }
    model match {
      case rich: RichModel[_] => rich.open()          // ModelUtils.openModel (FT/common/functions/util/ModelUtils.scala:13-17)
      case _ =>
    }
  }

  override def close(): Unit = model match {
    case rich: RichModel[_] => rich.close()           // ModelUtils.closeModel (:23-27)
    case _ =>
  }
}

/** `AbstractMapFunction` (FT/common/functions/AbstractMapFunction.scala:12-14): the blocking 1-in/1-out call. */
abstract class B200MapFunction[IN, OUT] extends RichMapFunction[IN, OUT] with B200ModelAwareFunction

/**
  * The micro-batched operator behind a 1-in/1-out map (`mapWithModel`, FT/streaming/package.scala:23-34), built on the hook
  * the reference already has for it (`AbstractFlatMapFunction`, FT/common/functions/AbstractFlatMapFunction.scala:12-14).
  *
  * Every record is handed to `Native.submit` with a tag (its upload starts at once); finished records come back from
  * `Native.poll` in submission order and are emitted then.  The obligations micro-batching adds to the reference's strictly
  * synchronous map (SURVEY 8b):
  *   - order: results are emitted in submission order (the runner guarantees it per handle);
  *   - latency bound: a batch closes at `maxBatch` records or `maxDelayUs` after its first record (`ftf_config`);
  *   - checkpoints: `snapshotState` flushes and emits everything in flight BEFORE the barrier goes downstream, so that no
  *     record is owed across a checkpoint and the operator keeps no state; `close` does the same;
  *   - a record that fails on its own (a corrupt image file) fails its own `flatMap` result, not the batch: `onFailure`
  *     decides (default: rethrow, which fails the task as a failing `map()` would in the reference).
  *
  * `encode` turns a record into the tensor to submit; `decode` turns the polled outputs (in `outputKeys` order) into the
  * operator's output.  Records are kept in `pending` until their result arrives because pinned payloads are read in place.
  */
abstract class MicroBatchedFlatMapFunction[IN, OUT](outputKeys: Array[String])
  extends RichFlatMapFunction[IN, OUT] with B200ModelAwareFunction with CheckpointedFunction {

  /** The runner handle of the opened model. */
  protected def runner: Long

  protected def encode(in: IN): HostTensor

  protected def decode(in: IN, outputs: Array[HostTensor]): OUT

  protected def onFailure(in: IN, status: Int): Option[OUT] =
    throw new IllegalArgumentException(s"record failed with status $status: ${Native.lastError(runner)}")

  private val cap = 256
  @transient private var nextTag = 0L
  @transient private var selected = false
  @transient private var pending: java.util.ArrayDeque[IN] = _
  @transient private var out: Collector[OUT] = _
  @transient private var tags: Array[Long] = _
  @transient private var status: Array[Int] = _
  @transient private var dtypes: Array[Int] = _
  @transient private var shapes: Array[Array[Long]] = _
  @transient private var bufs: Array[ByteBuffer] = _

  override def open(parameters: Configuration): Unit = {
    super.open(parameters)
    val k = outputKeys.length
    pending = new java.util.ArrayDeque[IN]()
    tags = new Array[Long](cap); status = new Array[Int](cap)
    dtypes = new Array[Int](cap * k); shapes = new Array[Array[Long]](cap * k); bufs = new Array[ByteBuffer](cap * k)
  }

  override def flatMap(in: IN, collector: Collector[OUT]): Unit = {
    out = collector
    val t = encode(in)
    while (!Native.submit(runner, nextTag, t.dtype, t.shape, t.data)) drain(block = true)   // every slot waits to be polled
    if (!selected) { Native.pollSelect(runner, outputKeys); selected = true }               // outputs are fixed by the first record
    pending.add(in)
    nextTag += 1
    drain(block = false)
  }

  private def drain(block: Boolean): Int = {
    val k = outputKeys.length
    val n = Native.poll(runner, block, k, tags, status, dtypes, shapes, bufs)
    var i = 0
    while (i < n) {
      val in = pending.poll()                                   // submission order == emission order
      if (status(i) == Native.OK) {
        val outs = Array.tabulate(k)(j => HostTensor(dtypes(i * k + j), shapes(i * k + j), bufs(i * k + j)))
        out.collect(decode(in, outs))
      } else onFailure(in, status(i)).foreach(out.collect)
      i += 1
    }
    n
  }

  private def flushAll(): Unit = if (runner != 0L && pending != null && !pending.isEmpty) {
    Native.flush(runner)
    while (!pending.isEmpty) drain(block = true)
  }

  override def snapshotState(context: FunctionSnapshotContext): Unit = flushAll()

  override def initializeState(context: FunctionInitializationContext): Unit = ()   // stateless: nothing is owed across a checkpoint

  override def close(): Unit = {
    try flushAll() finally super.close()
  }
}

/**
  * The labeling job's map (EX/inception/inception.scala:39-46) as a micro-batched operator: normalised images in, the
  * `take` best labels out (`toTextLabels`, EX/inception/InceptionModel.scala:81-89: stable descending order, the label list
  * truncates the score row, equal scores keep ascending index order -- computed on the GPU, fetched as "top_k_indices" /
  * "top_k_scores").
  */
class LabelImagesFunction(val inception: InceptionModelB200, take: Int = 3)
  extends MicroBatchedFlatMapFunction[(String, HostTensor), (String, List[(Float, String)])](Array("top_k_indices", "top_k_scores")) {

  def model: Model[_] = inception

  protected def runner: Long = inception.handle

  protected def encode(in: (String, HostTensor)): HostTensor = in._2

  protected def decode(in: (String, HostTensor), outputs: Array[HostTensor]): (String, List[(Float, String)]) = {
    val idx = outputs(0).data.order(java.nio.ByteOrder.nativeOrder()).asIntBuffer()
    val score = outputs(1).data.order(java.nio.ByteOrder.nativeOrder()).asFloatBuffer()
    val labeled = (0 until math.min(take, idx.remaining())).toList.collect {
      case i if idx.get(i) >= 0 && idx.get(i) < inception.labels.length => (score.get(i), inception.labels(idx.get(i)))
    }
    (in._1, labeled)
  }
}
