package org.apache.flink.contrib.tensorflow.b200;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;

/**
 * A ring of pinned host memory ({@link Native#hostAlloc}) that record payloads are carved from in arrival order.
 *
 * <p>Why a ring and not one allocation per record: (1) {@code cudaHostAlloc} takes a page-locking system call, far too slow
 * per record; (2) the runner merges the uploads of consecutive records that lie back to back inside ONE pinned allocation
 * into a single copy (include/ftf.h, ftf_submit) -- a 1 MB copy moves at 44 GB/s over PCIe 5 x16, a 16 MB copy at 53 --
 * and consecutive slices of this ring are exactly that.
 *
 * <p>A slice stays valid until the ring has wrapped around once: size the ring for everything that can be in flight
 * (three micro-batches inside the runner, what waits to be polled, and what the operator has deserialised ahead), e.g.
 * {@code 8 * maxBatch * recordBytes}.  Not thread-safe: one arena per task thread, like the runner itself.
 *
 * <p>SOURCE ONLY: not compiled in this repository's environment (no JDK).
 */
public final class PinnedArena implements AutoCloseable {
	private final ByteBuffer ring;
	private int cursor;

	public PinnedArena(long capacityBytes) {
		if (capacityBytes <= 0 || capacityBytes > Integer.MAX_VALUE) {
			throw new IllegalArgumentException("capacity must be in (0, 2 GiB)");
		}
		this.ring = Native.hostAlloc(capacityBytes).order(ByteOrder.nativeOrder());
	}

	/** The next {@code nbytes} of the ring (wrapping to its start when they do not fit behind the cursor). */
	public ByteBuffer allocate(int nbytes) {
		if (nbytes < 0 || nbytes > ring.capacity()) {
			throw new IllegalArgumentException("record of " + nbytes + " bytes does not fit a ring of " + ring.capacity());
		}
		if (cursor + nbytes > ring.capacity()) {
			cursor = 0;
		}
		ByteBuffer view = ring.duplicate();
		view.position(cursor).limit(cursor + nbytes);
		ByteBuffer slice = view.slice().order(ByteOrder.nativeOrder());
		cursor += nbytes;   // back to back: the runner asks that the next payload starts where the previous one ended
		return slice;
	}

	@Override
	public void close() {
		Native.hostFree(ring);
	}
}
