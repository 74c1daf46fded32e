/*
 * PageStore backed by libaps.so (HBM slab + hash table on one B200).
 * Not compiled in this repository: the build image has no JDK (see INTEGRATION.md).
 */
package alluxio.client.file.cache.store;

import alluxio.client.file.cache.PageId;
import alluxio.client.file.cache.PageInfo;
import alluxio.client.file.cache.PageStore;
import alluxio.exception.PageNotFoundException;

import com.google.common.base.Preconditions;
import org.slf4j.Logger;
import org.slf4j.LoggerFactory;

import java.io.IOException;
import java.nio.ByteBuffer;
import java.util.ArrayList;
import java.util.List;
import java.util.concurrent.ConcurrentHashMap;
import java.util.stream.Stream;

/**
 * Drop-in for {@link LocalPageStore}: same contract, pages live in GPU memory. The native side is the
 * C ABI of include/aps.h reached through the thin JNI shim integration/jni/aps_jni.c.
 */
public class GpuPageStore implements PageStore {
  private static final Logger LOG = LoggerFactory.getLogger(GpuPageStore.class);

  static {
    System.loadLibrary("aps_jni"); // links against libaps.so
  }

  // status codes of include/aps.h
  private static final int OK = 0;
  private static final int ENOTFOUND = 1;
  private static final int EINVAL = 2;

  private final long mHandle;
  private final long mCacheSize;
  private final int mPageSize;
  /** Interned file ids that are neither decimal longs nor 32-hex strings (exact, never hashed). */
  private final ConcurrentHashMap<String, Long> mInterned = new ConcurrentHashMap<>();
  private final List<String> mInternedNames = new ArrayList<>();
  /** One pinned direct buffer per thread: byte[] <-> device staging without per-call allocation. */
  private final ThreadLocal<ByteBuffer> mStaging;
  /** Every pinned buffer handed out by nativeAllocPinned, so that close() can give them back (cudaFreeHost). */
  private final List<ByteBuffer> mPinned = new ArrayList<>();
  private final String mSnapshotPath;
  private final java.util.concurrent.atomic.AtomicBoolean mClosed = new java.util.concurrent.atomic.AtomicBoolean();

  public GpuPageStore(GpuPageStoreOptions options) throws IOException {
    mCacheSize = options.getCacheSize();
    mPageSize = (int) options.getPageSize();
    long[] handle = new long[1];
    int rc = nativeOpen(options.getDevice(), mPageSize, mCacheSize, options.getSlabSlots(),
        options.getTableCapacity(), options.getMaxBatch(), options.isCrcEnabled() ? 1 : 0, handle);
    if (rc != OK) {
      throw new IOException("aps_open failed: " + nativeStrError(rc));
    }
    mHandle = handle[0];
    mStaging = ThreadLocal.withInitial(() -> allocPinned(mPageSize));
    mSnapshotPath = options.getSnapshotPath();
    if (mSnapshotPath != null && new java.io.File(mSnapshotPath).exists()) {
      // warm restart: getPages() then lists these pages and LocalCacheManager.restore adopts them.  Ids that are neither
      // decimal nor 32-hex were interned by the process that wrote the snapshot: take its table over first.
      String[] names = nativeSnapshotNames(mSnapshotPath);
      long[] counts = new long[2];
      rc = names == null ? EINVAL : OK;
      if (rc == OK) {
        synchronized (mInternedNames) {
          for (String name : names) {
            mInterned.put(name, (long) mInternedNames.size());
            mInternedNames.add(name);
          }
        }
        rc = nativeSnapshotLoad(mHandle, mSnapshotPath, counts);
      }
      if (rc != OK) {
        LOG.warn("ignoring snapshot {}: {}", mSnapshotPath, nativeStrError(rc));
      }
    }
  }

  /** A page-locked direct buffer (kernels read and write it in place); given back to the driver by close(). */
  public ByteBuffer allocPinned(int bytes) {
    ByteBuffer b = nativeAllocPinned(bytes);
    if (b == null) {
      throw new OutOfMemoryError("cudaHostAlloc(" + bytes + ")");
    }
    synchronized (mPinned) {
      mPinned.add(b);
    }
    return b;
  }

  /** PageId -> (fileHi, fileLo): disjoint ranges, so equal keys imply equal strings. */
  private long[] fileKey(String fileId) {
    if (fileId.matches("0|-?[1-9][0-9]{0,18}")) {
      try {
        return new long[] {0L, Long.parseLong(fileId)};
      } catch (NumberFormatException e) {
        // falls through to interning
      }
    }
    if (fileId.matches("[0-9a-f]{32}")) {
      long hi = Long.parseUnsignedLong(fileId.substring(0, 16), 16);
      long lo = Long.parseUnsignedLong(fileId.substring(16), 16);
      if (hi != 0L && hi != -1L) {
        return new long[] {hi, lo};
      }
    }
    long n = mInterned.computeIfAbsent(fileId, k -> {
      synchronized (mInternedNames) {
        mInternedNames.add(k);
        return (long) (mInternedNames.size() - 1);
      }
    });
    return new long[] {-1L, n};
  }

  @Override
  public void put(PageId pageId, byte[] page) throws IOException {
    if (page.length > mPageSize) {
      throw new IOException("Failed to put page " + pageId + ": " + page.length + " bytes exceed the page size " + mPageSize);
    }
    long[] k = fileKey(pageId.getFileId());
    ByteBuffer staging = mStaging.get();
    staging.clear();
    staging.put(page);
    int rc = nativePut(mHandle, k[0], k[1], pageId.getPageIndex(), staging, page.length);
    if (rc != OK) {
      throw new IOException("Failed to put page " + pageId + ": " + nativeStrError(rc));
    }
  }

  @Override
  public int get(PageId pageId, int pageOffset, int bytesToRead, byte[] buffer, int bufferOffset)
      throws IOException, PageNotFoundException {
    Preconditions.checkArgument(pageOffset >= 0, "page offset should be non-negative");
    Preconditions.checkArgument(buffer.length >= bufferOffset,
        "page offset %s should be less or equal than buffer length %s", bufferOffset, buffer.length);
    long[] k = fileKey(pageId.getFileId());
    ByteBuffer staging = mStaging.get();
    int room = Math.min(buffer.length - bufferOffset, mPageSize);
    int[] outLen = new int[1];
    int rc = nativeGet(mHandle, k[0], k[1], pageId.getPageIndex(), pageOffset,
        Math.max(0, bytesToRead), staging, room, outLen);
    if (rc == ENOTFOUND) {
      throw new PageNotFoundException(pageId.toString());
    }
    if (rc == EINVAL) {
      throw new IllegalArgumentException("page offset " + pageOffset + " exceeded page size");
    }
    if (rc != OK) {
      throw new IOException("Failed to read page " + pageId + ": " + nativeStrError(rc));
    }
    staging.position(0);
    staging.get(buffer, bufferOffset, outLen[0]);
    return outLen[0];
  }

  @Override
  public void delete(PageId pageId) throws IOException, PageNotFoundException {
    long[] k = fileKey(pageId.getFileId());
    int rc = nativeDelete(mHandle, k[0], k[1], pageId.getPageIndex());
    if (rc == ENOTFOUND) {
      throw new PageNotFoundException(pageId.toString());
    }
    if (rc != OK) {
      throw new IOException("Failed to delete page " + pageId + ": " + nativeStrError(rc));
    }
  }

  @Override
  public Stream<PageInfo> getPages() throws IOException {
    // long[3n + ...]: (fileHi, fileLo, pageIndex, length) per resident page
    long[] flat = nativeIterate(mHandle);
    if (flat == null) {
      throw new IOException("Failed to list the pages of the GPU store");
    }
    List<PageInfo> pages = new ArrayList<>(flat.length / 4);
    for (int i = 0; i + 3 < flat.length; i += 4) {
      String fileId;
      if (flat[i] == 0L) {
        fileId = Long.toString(flat[i + 1]);
      } else if (flat[i] == -1L) {
        synchronized (mInternedNames) {
          fileId = mInternedNames.get((int) flat[i + 1]);
        }
      } else {
        fileId = String.format("%016x%016x", flat[i], flat[i + 1]);
      }
      pages.add(new PageInfo(new PageId(fileId, flat[i + 2]), flat[i + 3]));
    }
    return pages.stream();
  }

  @Override
  public long getCacheSize() {
    return mCacheSize;
  }

  @Override
  public void close() {
    if (!mClosed.compareAndSet(false, true)) {
      return; // idempotent (the native close is as well)
    }
    if (mSnapshotPath != null) {
      String[] names;
      synchronized (mInternedNames) {
        names = mInternedNames.toArray(new String[0]);
      }
      int rc = nativeSnapshotSave(mHandle, mSnapshotPath, names);
      if (rc != OK) {
        LOG.warn("could not save snapshot {}: {}", mSnapshotPath, nativeStrError(rc));
      }
    }
    nativeClose(mHandle);
    synchronized (mPinned) {
      for (ByteBuffer b : mPinned) {
        nativeFreePinned(b);
      }
      mPinned.clear();
    }
  }

  // ---- batched calls (additive; the path bench.py measures).  One JNI call and one kernel launch per batch; the pages
  // travel through ONE direct ByteBuffer: a pinned one from allocPinned (fastest), or any ByteBuffer.allocateDirect.

  private long[] keysOf(PageId[] ids) {
    long[] keys = new long[3 * ids.length];
    for (int i = 0; i < ids.length; i++) {
      long[] k = fileKey(ids[i].getFileId());
      keys[3 * i] = k[0];
      keys[3 * i + 1] = k[1];
      keys[3 * i + 2] = ids[i].getPageIndex();
    }
    return keys;
  }

  /**
   * Reads bytesToRead[i] bytes of page ids[i] from pageOffsets[i] into dst at dstOffsets[i], all in one call.
   *
   * @return per request the bytes read; -1 where the page is absent (PageNotFoundException in the single-page call),
   *         -2 where the offset lies beyond the page (IllegalArgumentException there); outCrc (optional) receives the
   *         CRC32C of the bytes returned
   */
  public int[] getBatch(PageId[] ids, int[] pageOffsets, int[] bytesToRead, ByteBuffer dst, long[] dstOffsets,
      int[] outCrc) throws IOException {
    Preconditions.checkArgument(dst.isDirect(), "the destination must be a direct buffer");
    int n = ids.length;
    int[] outLen = new int[n];
    int[] crc = outCrc != null ? outCrc : new int[n];
    int rc = nativeGetBatch(mHandle, n, keysOf(ids), pageOffsets, bytesToRead, dst, dstOffsets, outLen, crc);
    if (rc != OK) {
      throw new IOException("Failed to read " + n + " pages: " + nativeStrError(rc));
    }
    for (int i = 0; i < n; i++) {
      if (outLen[i] < -EINVAL) {
        throw new IOException("Failed to read page " + ids[i] + ": " + nativeStrError(-outLen[i]));
      }
    }
    return outLen;
  }

  /** Stores lengths[i] bytes at src + srcOffsets[i] as page ids[i] (ids distinct), all in one call. */
  public void putBatch(PageId[] ids, ByteBuffer src, long[] srcOffsets, int[] lengths) throws IOException {
    Preconditions.checkArgument(src.isDirect(), "the source must be a direct buffer");
    int n = ids.length;
    int[] status = new int[n];
    int rc = nativePutBatch(mHandle, n, keysOf(ids), src, srcOffsets, lengths, status, new int[n]);
    if (rc != OK) {
      throw new IOException("Failed to put " + n + " pages: " + nativeStrError(rc));
    }
    for (int i = 0; i < n; i++) {
      if (status[i] != OK) {
        throw new IOException("Failed to put page " + ids[i] + ": " + nativeStrError(status[i]));
      }
    }
  }

  /** Deletes the pages (ids distinct) in one call. @return per id whether it was present */
  public boolean[] deleteBatch(PageId[] ids) throws IOException {
    int n = ids.length;
    int[] status = new int[n];
    int rc = nativeDeleteBatch(mHandle, n, keysOf(ids), status);
    if (rc != OK) {
      throw new IOException("Failed to delete " + n + " pages: " + nativeStrError(rc));
    }
    boolean[] present = new boolean[n];
    for (int i = 0; i < n; i++) {
      if (status[i] != OK && status[i] != ENOTFOUND) {
        throw new IOException("Failed to delete page " + ids[i] + ": " + nativeStrError(status[i]));
      }
      present[i] = status[i] == OK;
    }
    return present;
  }

  /**
   * MD5 of n byte ranges of a pinned buffer, many at once (the arithmetic of ChecksumCommand.calculateChecksum).
   *
   * @return the digests, 16 bytes each
   */
  public static byte[] md5Batch(int device, ByteBuffer pinned, long[] offsets, long[] lengths) throws IOException {
    byte[] out = new byte[16 * offsets.length];
    int rc = nativeMd5Batch(device, offsets.length, pinned, offsets, lengths, out);
    if (rc != OK) {
      throw new IOException("md5 batch failed: " + nativeStrError(rc));
    }
    return out;
  }

  // ---- JNI (integration/jni/aps_jni.c)
  private static native int nativeOpen(int device, int pageSize, long cacheBytes, long slabSlots,
      long tableCapacity, int maxBatch, int crc, long[] handleOut);
  private static native int nativeClose(long handle);
  private static native int nativePut(long handle, long fileHi, long fileLo, long pageIndex,
      ByteBuffer pinned, int len);
  private static native int nativeGet(long handle, long fileHi, long fileLo, long pageIndex,
      int pageOffset, int bytesToRead, ByteBuffer pinned, int room, int[] outLen);
  private static native int nativeDelete(long handle, long fileHi, long fileLo, long pageIndex);
  private static native long[] nativeIterate(long handle);
  private static native ByteBuffer nativeAllocPinned(int bytes);
  private static native String nativeStrError(int code);
  private static native int nativeSnapshotSave(long handle, String path, String[] internedNames);
  private static native String[] nativeSnapshotNames(String path);
  private static native void nativeFreePinned(ByteBuffer pinned);
  private static native int nativeGetBatch(long handle, int n, long[] keys, int[] pageOffsets, int[] bytesToRead,
      ByteBuffer buffer, long[] offsets, int[] outLen, int[] outCrc);
  private static native int nativePutBatch(long handle, int n, long[] keys, ByteBuffer buffer, long[] offsets,
      int[] lengths, int[] outStatus, int[] outCrc);
  private static native int nativeDeleteBatch(long handle, int n, long[] keys, int[] outStatus);
  private static native int nativeMd5Batch(int device, int n, ByteBuffer pinned, long[] offsets, long[] lengths,
      byte[] digests);
  private static native int nativeSnapshotLoad(long handle, String path, long[] loadedAndRejected);
}
