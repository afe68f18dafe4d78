package org.apache.flink.contrib.tensorflow.b200

import org.apache.flink.contrib.tensorflow.models.RichModel
import org.apache.flink.util.Preconditions.checkState

/** The runner settings the reference hard-codes (EX/inception/ImageNormalization.scala:51-54) or has no notion of. */
final case class B200Config(maxBatch: Int = 64, maxDelayUs: Int = 2000, outH: Int = 224, outW: Int = 224,
                            mean: Float = 117f, scale: Float = 1f, computeDtype: Int = Native.BFLOAT16, topK: Int = 3,
                            classes: Int = 1000, seed: Long = 2024L) extends Serializable

/**
  * A model bound to one GPU.  `open()` of the reference takes no argument; the B200 models need the subtask index to pick
  * their device, so the operator mix-in ([[B200ModelAwareFunction]]) sets it from the runtime context before `open()`.
  */
trait DeviceBound {
  @transient protected var subtaskIndex: Int = 0
  @transient protected var visibleGpus: Int = 1

  def bind(subtaskIndex: Int, visibleGpus: Int): Unit = {
    this.subtaskIndex = subtaskIndex
    this.visibleGpus = math.max(visibleGpus, 1)
  }

  /** subtask i -> GPU i mod #GPUs (SURVEY 8e): records are sharded by Flink's partitioner, weights replicated. */
  protected def device: Int = subtaskIndex % visibleGpus
}

/**
  * `GenericModel` (FT/models/generic/GenericModel.scala:22-43) for the graphs the runner has built in: where the reference
  * loads a GraphDef and opens `new Session(graph)`, this creates one runner.  `kind` is "builtin:image_normalization"
  * (EX/inception/ImageNormalization.scala:40-85), "builtin:inception_v3" (EX/inception/InceptionModel.scala:47-67) or
  * "builtin:normalize_label" (both chained on the device).  A failed create leaves nothing behind (:24-31); close is
  * null-safe and idempotent (:34-43).
  */
abstract class B200GenericModel[Self <: B200GenericModel[Self]](kind: String, methodName: String, config: B200Config)
  extends RichModel[Self] with DeviceBound {
  that: Self =>

  @transient protected[this] var runner: Long = 0L

  override def open(): Unit = {
    checkState(runner == 0L)
    runner = Native.create(kind, null, methodName, device, config.maxBatch, config.maxDelayUs, config.outH, config.outW,
      config.mean, config.scale, config.computeDtype, config.topK, config.classes, config.seed)
  }

  override def close(): Unit = {
    if (runner != 0L) {
      Native.destroy(runner)
      runner = 0L
    }
  }
}

/**
  * `TensorFlowModel` (FT/models/savedmodel/TensorFlowModel.scala:12-50) over a SavedModel directory in the
  * models/half_plus_two layout.  One runner per signature that is used (a runner is bound to one SignatureDef, as a
  * `ModelFunction` is); the metagraph's signature map comes from `Native.describeSavedModel`, which reads saved_model.pb
  * without loading and closing a second bundle (FT/models/savedmodel/DefaultSavedModelLoader.scala:22-31).
  * "file:" paths are used in place (:43-45); copying a remote model to a local directory first is the caller's business
  * (the reference's `FileCache.copy`, :46-53).
  */
abstract class B200TensorFlowModel[Self <: B200TensorFlowModel[Self]](modelPath: String, tags: String*)
  extends RichModel[Self] with DeviceBound {
  that: Self =>

  @transient private var runners: scala.collection.mutable.Map[String, Long] = _

  /** The MetaGraphDef's tags, signatures and variables as JSON (what `metagraph` / `signatureDef(name)` are read from). */
  def describe: String = Native.describeSavedModel(modelPath, if (tags.isEmpty) "serve" else tags.mkString(","))

  override def open(): Unit = {
    checkState(runners == null)                       // TensorFlowModel.scala:40
    runners = scala.collection.mutable.Map.empty
  }

  override def close(): Unit = {
    if (runners != null) {
      runners.values.foreach(Native.destroy)
      runners = null
    }
  }

  /** The runner for one signature, created on first use with the method name checked against the SignatureDef's. */
  protected def session(signatureKey: String, methodName: String): Long = {
    require(runners != null, "a session must be provided")
    runners.getOrElseUpdate(signatureKey,
      Native.create(modelPath, signatureKey, methodName, device, 64, 2000, 0, 0, 0f, 1f, 0, -1, 0, 0L))
  }

  protected def modelFunction[T <: B200GraphMethod](signatureKey: String)(implicit method: T): B200ModelFunction[T] =
    B200ModelFunction[T](session(signatureKey, method.name))
}

/** `HalfPlusTwo` (FTT/ml/HalfPlusTwo.scala:15-28). */
class HalfPlusTwoB200(modelPath: String) extends B200TensorFlowModel[HalfPlusTwoB200](modelPath, "serve") {
  def regress_x_to_y: B200ModelFunction[RegressionMethodB200] = modelFunction[RegressionMethodB200]("regress_x_to_y")
  def regress_x_to_y2: B200ModelFunction[RegressionMethodB200] = modelFunction[RegressionMethodB200]("regress_x_to_y2")
}

/** `ImageNormalization` (EX/inception/ImageNormalization.scala:40-85) with its constants as parameters. */
class ImageNormalizationB200(config: B200Config = B200Config())
  extends B200GenericModel[ImageNormalizationB200]("builtin:image_normalization", "inception/normalize", config) {
  def normalize: B200ModelFunction[ImageNormalizationMethodB200] = B200ModelFunction[ImageNormalizationMethodB200](runner)
}

/** `InceptionModel` (EX/inception/InceptionModel.scala:47-90). */
class InceptionModelB200(val labels: Seq[String], config: B200Config = B200Config(outH = 299, outW = 299, mean = 128f, scale = 128f))
  extends B200GenericModel[InceptionModelB200]("builtin:inception_v3", "inception/label", config) {
  def label: B200ModelFunction[LabelMethodB200] = B200ModelFunction[LabelMethodB200](runner)
  private[b200] def handle: Long = runner
}
