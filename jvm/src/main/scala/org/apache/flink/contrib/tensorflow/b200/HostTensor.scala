package org.apache.flink.contrib.tensorflow.b200

import java.nio.{ByteBuffer, ByteOrder}

import org.apache.flink.contrib.tensorflow.types.{TensorValue, TensorValueBuilder}
import org.apache.flink.contrib.tensorflow.util.TFUtils

/**
  * What travels across the boundary in place of an `org.tensorflow.Tensor`: a dtype code (the TF `DataType` enum values of
  * FT/util/TFUtils.java:10-32), a shape and a DIRECT byte buffer in native order.  The runner reads the buffer where it
  * lies -- in place by the copy engine when it came from [[Native.hostAlloc]] -- so creating a HostTensor copies nothing,
  * where `TensorValue.toTensor` (FT/types/TensorValue.java:131-133) copies heap to native memory for every record.
  *
  * Ownership follows the reference: inputs are borrowed (the caller keeps them alive for the call, `ModelFunction` never
  * closes them, FT/models/ModelFunction.scala:47-56); outputs of a run are views of runner-owned pinned memory, valid until
  * the surrounding managed resource is released (FT/models/ModelFunction.scala:68-69) -- copy with [[toValue]] to keep them.
  *
  * SOURCE ONLY: not compiled in this repository's environment (no JVM).
  */
final case class HostTensor(dtype: Int, shape: Array[Long], data: ByteBuffer) {
  require(data.isDirect, "tensor payloads must be direct ByteBuffers")

  def numDimensions: Int = shape.length

  def numBytes: Int = data.capacity()

  /** `RichTensor.toValue` (FT/package.scala:24-32): a serializable copy; this tensor stays with its owner. */
  def toValue[K <: org.apache.flink.api.java.tuple.Tuple, V]: TensorValue[K, V] = {
    val copy = ByteBuffer.allocate(data.capacity()).order(ByteOrder.nativeOrder())
    copy.put(data.duplicate().rewind().asInstanceOf[ByteBuffer]).rewind()
    TensorValueBuilder.newBuilder().dataType(TFUtils.getDataType(dtype)).shape(shape).data(copy).build()
      .asInstanceOf[TensorValue[K, V]]
  }
}

object HostTensor {
  /**
    * A tensor over a record's payload.  A direct payload (one deserialised by [[TensorValueSerializer]]) is used in place;
    * a heap payload is copied once into a pinned buffer owned by the returned handle.
    */
  def fromValue(dtype: Int, shape: Array[Long], payload: ByteBuffer): (HostTensor, Option[ByteBuffer]) = {
    val src = payload.duplicate().order(ByteOrder.nativeOrder())
    src.rewind()
    if (src.isDirect) (HostTensor(dtype, shape, src), None)
    else {
      val pinned = Native.hostAlloc(math.max(src.remaining(), 1)).order(ByteOrder.nativeOrder())
      pinned.put(src).rewind()
      val sized = pinned.duplicate().order(ByteOrder.nativeOrder())
      sized.limit(payload.remaining())
      (HostTensor(dtype, shape, sized.slice().order(ByteOrder.nativeOrder())), Some(pinned))
    }
  }
}
