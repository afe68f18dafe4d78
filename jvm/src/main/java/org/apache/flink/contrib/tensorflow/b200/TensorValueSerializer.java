package org.apache.flink.contrib.tensorflow.b200;

import java.io.IOException;
import java.nio.ByteBuffer;
import java.nio.ByteOrder;

import org.apache.flink.api.common.typeutils.TypeSerializer;
import org.apache.flink.contrib.tensorflow.types.TensorValue;
import org.apache.flink.contrib.tensorflow.types.TensorValueBuilder;
import org.apache.flink.contrib.tensorflow.util.TFUtils;
import org.apache.flink.core.memory.DataInputView;
import org.apache.flink.core.memory.DataOutputView;
import org.apache.flink.types.VersionMismatchException;

/**
 * A Flink {@link TypeSerializer} for {@link TensorValue} that speaks exactly the wire image of
 * {@code TensorValue.write/read} (FT/types/TensorValue.java:150-187):
 * {@code 0x01 | i32 dtype | i32 rank | rank x i64 dims | i32 nbytes | payload} -- header big-endian, payload native order.
 *
 * <p>The reference relies on Flink's generic {@code Value} handling, whose {@code copy(in, out)} path ends in
 * {@code TensorValue.copyInternal} (TensorValue.java:189-198): that method forwards {@code rank} BYTES instead of
 * {@code 8 * rank} and drops the rank and length integers, so a record copied between operators without being
 * deserialised is corrupt.  {@link #copy(DataInputView, DataOutputView)} below forwards the whole image.
 *
 * <p>Payloads are read into consecutive slices of a {@link PinnedArena} (one ring of {@link Native#hostAlloc} memory per
 * task thread), so a deserialised record can be handed to {@link Native#submit} and uploaded in place (no staging copy in
 * the runner), and the uploads of records that arrive one after the other are merged into one copy.
 *
 * <p>SOURCE ONLY: not compiled in this repository's environment (no JDK).
 */
public final class TensorValueSerializer extends TypeSerializer<TensorValue> {
	private static final long serialVersionUID = 1L;
	private static final byte VERSION_1 = 0x1;

	private final long arenaBytes;
	private transient PinnedArena arena;

	/**
	 * @param arenaBytes size of the pinned ring payloads are read into (for operators that feed a B200 runner; it must hold
	 *        every record that can be in flight, see {@link PinnedArena}); 0: plain direct buffers.
	 */
	public TensorValueSerializer(long arenaBytes) {
		this.arenaBytes = arenaBytes;
	}

	private ByteBuffer payloadBuffer(int nbytes) {
		if (arenaBytes <= 0 || nbytes == 0) {
			return ByteBuffer.allocateDirect(nbytes);
		}
		if (arena == null) {
			arena = new PinnedArena(arenaBytes);
		}
		return arena.allocate(nbytes);
	}

	@Override
	public boolean isImmutableType() {
		return false;
	}

	@Override
	public TypeSerializer<TensorValue> duplicate() {
		return new TensorValueSerializer(arenaBytes);   // a ring per task thread
	}

	@Override
	public TensorValue createInstance() {
		return new TensorValue();
	}

	@Override
	public TensorValue copy(TensorValue from) {
		return from.copy();   // shallow: the buffer is immutable by contract (TensorValue.java:43,213-217)
	}

	@Override
	public TensorValue copy(TensorValue from, TensorValue reuse) {
		from.copyTo(reuse);
		return reuse;
	}

	@Override
	public int getLength() {
		return -1;   // variable length
	}

	@Override
	public void serialize(TensorValue record, DataOutputView target) throws IOException {
		record.write(target);
	}

	@Override
	public TensorValue deserialize(DataInputView source) throws IOException {
		byte version = source.readByte();
		if (version != VERSION_1) {
			throw new VersionMismatchException("incompatible tensor value");
		}
		int dtype = source.readInt();
		int rank = source.readInt();
		if (rank < 0 || rank > 25) {
			throw new UnsupportedOperationException("rank " + rank);   // TFUtils.shapeOf covers Tuple0..Tuple25
		}
		long[] shape = new long[rank];
		for (int i = 0; i < rank; i++) {
			shape[i] = source.readLong();
		}
		int nbytes = source.readInt();
		ByteBuffer payload = payloadBuffer(nbytes).order(ByteOrder.nativeOrder());
		byte[] chunk = new byte[Math.min(Math.max(nbytes, 1), 1 << 16)];
		for (int done = 0; done < nbytes; ) {
			int n = Math.min(chunk.length, nbytes - done);
			source.readFully(chunk, 0, n);
			payload.put(chunk, 0, n);
			done += n;
		}
		payload.rewind();
		return TensorValueBuilder.newBuilder().dataType(TFUtils.getDataType(dtype)).shape(shape).data(payload).build();
	}

	@Override
	public TensorValue deserialize(TensorValue reuse, DataInputView source) throws IOException {
		deserialize(source).copyTo(reuse);
		return reuse;
	}

	/** The whole wire image, field by field: what {@code TensorValue.copyInternal} gets wrong. */
	@Override
	public void copy(DataInputView source, DataOutputView target) throws IOException {
		byte version = source.readByte();
		if (version != VERSION_1) {
			throw new VersionMismatchException("incompatible tensor value");
		}
		target.writeByte(version);
		target.writeInt(source.readInt());            // dtype
		int rank = source.readInt();
		target.writeInt(rank);
		target.write(source, 8 * rank);               // rank x i64 dims
		int nbytes = source.readInt();
		target.writeInt(nbytes);
		target.write(source, nbytes);                 // payload
	}

	@Override
	public boolean equals(Object obj) {
		return obj instanceof TensorValueSerializer && ((TensorValueSerializer) obj).arenaBytes == arenaBytes;
	}

	@Override
	public boolean canEqual(Object obj) {
		return obj instanceof TensorValueSerializer;
	}

	@Override
	public int hashCode() {
		return TensorValueSerializer.class.hashCode() + Long.hashCode(arenaBytes);
	}
}
