package org.apache.flink.contrib.tensorflow.b200;

import java.nio.ByteBuffer;

/**
 * The JNI surface of libftf_jni.so (jni/ftf_jni.cc), which forwards to the C ABI of libftf_b200.so (include/ftf.h).
 *
 * <p>Stands where TF-Java's {@code org.tensorflow.Session}, {@code SavedModelBundle} and {@code Tensor} stand on the
 * model-invocation path of flink-tensorflow.  A runner handle is a {@code long}; tensor payloads are direct
 * {@link ByteBuffer}s (the pinned ones from {@link #hostAlloc} are uploaded in place); a failing call throws the
 * exception TF-Java throws for the same status.
 *
 * <p>SOURCE ONLY: no JDK exists in the environment this repository is developed in; the class has not been compiled.
 */
public final class Native {
	static {
		System.loadLibrary("ftf_jni");
	}

	private Native() {
	}

	/** TF {@code DataType} codes used on the wire (FT/util/TFUtils.java:10-32) plus the two raw-image extensions. */
	public static final int FLOAT = 1, DOUBLE = 2, INT32 = 3, UINT8 = 4, STRING = 7, INT64 = 9, BOOL = 10, BFLOAT16 = 14;

	/** ftf_poll_status codes. */
	public static final int OK = 0, INVALID_ARGUMENT = 3;

	// lifecycle: RichModel.open()/close()
	public static native long create(String model, String signatureKey, String methodName, int device, int maxBatch,
		int maxDelayUs, int outH, int outW, float mean, float scale, int computeDtype, int topK, int classes, long seed);

	public static native void destroy(long runner);

	public static native String lastError(long runner);

	/** kind: 0 method name, 1 input keys, 2 input tensor names, 3 output keys, 4 output tensor names. */
	public static native String[] signatureInfo(long runner, int kind);

	public static native String describeSavedModel(String dir, String tagsCsv);

	// ModelFunction.apply
	public static native void run(long runner, String[] inKeys, int[] inDtypes, long[][] inShapes, ByteBuffer[] inData,
		String[] outKeys, int[] outDtypes, long[][] outShapes, ByteBuffer[] outData);

	public static native void release(long runner);

	// micro-batched per-record path
	public static native void pollSelect(long runner, String[] keys);

	public static native boolean submit(long runner, long tag, int dtype, long[] shape, ByteBuffer data);

	public static native int poll(long runner, boolean block, int nSelected, long[] tags, int[] status, int[] outDtypes,
		long[][] outShapes, ByteBuffer[] outData);

	public static native void flush(long runner);

	public static native long[] stats(long runner);

	// pinned record buffers
	public static native ByteBuffer hostAlloc(long nbytes);

	public static native void hostFree(ByteBuffer buffer);

	// TensorValue wire codec
	public static native long tvBinaryLength(int rank, long nbytes);

	public static native long tvEncode(int dtype, long[] shape, ByteBuffer data, ByteBuffer out);

	public static native void tvDecode(ByteBuffer wire, long[] header4, long[] dims25);

	public static native String version();
}
