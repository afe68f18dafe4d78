package org.apache.flink.contrib.tensorflow.b200

import java.nio.ByteBuffer

import resource._

/**
  * `ModelFunction` (FT/models/ModelFunction.scala:18-78) over a B200 runner.
  *
  * Where the reference builds a `Session.Runner`, feeds every input of the SignatureDef, fetches every output and calls
  * `runAndFetchMetadata()` (:46-66), this makes ONE native call: `Native.run` takes the inputs by signature key and returns
  * the outputs by signature key.  The checks of the reference stay where they were or move into the library with the same
  * outcome: a null session / signature / input is an `IllegalArgumentException` (`require`, :35-37,41); the method name is
  * compared with the SignatureDef's when the runner is created (`Native.create(..., methodName, ...)`, ftf.h ftf_create);
  * a missing input key throws `IllegalArgumentException("An input tensor named <key> must be provided for this
  * computation.")` from `ftf_run` (:52-53); tensor names ("op:idx", FT/types/TensorName.scala:14-19) were resolved once
  * at create time instead of per call (:54,60).  Outputs are bound by NAME, not by map iteration order (:74).
  */
trait B200ModelFunction[T <: B200GraphMethod] {
  def apply(in: T#Input): ManagedResource[T#Output]
}

object B200ModelFunction {

  def apply[T <: B200GraphMethod](runner: Long)(implicit method: T): B200ModelFunction[T] = {
    require(runner != 0L, "a session must be provided")
    require(method != null, "a signatureDef must be provided")
    val outKeys: Array[String] = Native.signatureInfo(runner, 3)

    new B200ModelFunction[T] {
      def apply(in: T#Input): ManagedResource[T#Output] = {
        require(in != null, "Input must be provided")

        // a managed resource that lazily runs the model; releasing it invalidates the output views
        val runResource = makeManagedResource {
          val inputs = method.inputs(in.asInstanceOf[method.Input])
          val keys = inputs.keys.toArray
          val tensors = keys.map(inputs)
          val outDtypes = new Array[Int](outKeys.length)
          val outShapes = new Array[Array[Long]](outKeys.length)
          val outData = new Array[ByteBuffer](outKeys.length)
          Native.run(runner, keys, tensors.map(_.dtype), tensors.map(_.shape), tensors.map(_.data),
            outKeys, outDtypes, outShapes, outData)
          outKeys.indices.map(i => outKeys(i) -> HostTensor(outDtypes(i), outShapes(i), outData(i))).toMap
        }(_ => Native.release(runner))(List.empty[Class[_ <: Throwable]])

        runResource.map(outputs => method.outputs(outputs).asInstanceOf[T#Output])
      }
    }
  }
}
