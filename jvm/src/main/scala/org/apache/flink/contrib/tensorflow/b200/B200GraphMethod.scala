package org.apache.flink.contrib.tensorflow.b200

/**
  * `GraphMethod` (FT/graphs/GraphMethod.scala:8-36) over [[HostTensor]]: same members, same meaning -- the method name
  * that must equal the SignatureDef's, the map of signature input keys to feed, the map of fetched outputs to unpack.
  */
trait B200GraphMethod {
  type Input
  type Output

  def name: String

  def inputs(in: Input): Map[String, HostTensor]

  def outputs(outputs: Map[String, HostTensor]): Output
}

/** `RegressionMethod` (FT/ml/signatures/RegressionMethod.scala:18-36): "tensorflow/serving/regress", inputs -> outputs. */
sealed trait RegressionMethodB200 extends B200GraphMethod {
  val name = "tensorflow/serving/regress"               // SignatureConstants.java:52
  override type Input = HostTensor                      // STRING [N]: serialized tf.Examples (TensorInjections.scala:49-72)
  override type Output = HostTensor                     // FLOAT [N,1]
  def inputs(i: Input): Map[String, HostTensor] = Map("inputs" -> i)
  def outputs(o: Map[String, HostTensor]): Output = o("outputs")
}

object RegressionMethodB200 {
  implicit val impl: RegressionMethodB200 = new RegressionMethodB200 {}
}

/** `ImageNormalizationMethod` (EX/inception/ImageNormalization.scala:14-32): "inception/normalize". */
sealed trait ImageNormalizationMethodB200 extends B200GraphMethod {
  val name = "inception/normalize"
  override type Input = HostTensor                      // STRING scalar (image file) or UINT8 [H,W,3] / [N,H,W,3]
  override type Output = HostTensor                     // FLOAT [N,out_h,out_w,3]
  def inputs(i: Input): Map[String, HostTensor] = Map("inputs" -> i)
  def outputs(o: Map[String, HostTensor]): Output = o("outputs")
}

object ImageNormalizationMethodB200 {
  implicit val impl: ImageNormalizationMethodB200 = new ImageNormalizationMethodB200 {}
}

/** `LabelMethod` (EX/inception/InceptionModel.scala:19-39): "inception/label". */
sealed trait LabelMethodB200 extends B200GraphMethod {
  val name = "inception/label"
  override type Input = HostTensor                      // FLOAT [N,299,299,3]
  override type Output = HostTensor                     // FLOAT [N,classes]
  def inputs(i: Input): Map[String, HostTensor] = Map("inputs" -> i)
  def outputs(o: Map[String, HostTensor]): Output = o("outputs")
}

object LabelMethodB200 {
  implicit val impl: LabelMethodB200 = new LabelMethodB200 {}
}
