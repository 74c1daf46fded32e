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
  private final String mSnapshotPath;

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
    mStaging = ThreadLocal.withInitial(() -> nativeAllocPinned(mPageSize));
    mSnapshotPath = options.getSnapshotPath();
    if (mSnapshotPath != null && new java.io.File(mSnapshotPath).exists()) {
      // warm restart: getPages() then lists these pages and LocalCacheManager.restore adopts them
      long[] counts = new long[2];
      rc = nativeSnapshotLoad(mHandle, mSnapshotPath, counts);
      if (rc != OK) {
        LOG.warn("ignoring snapshot {}: {}", mSnapshotPath, nativeStrError(rc));
      }
    }
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
    if (mSnapshotPath != null) {
      int rc = nativeSnapshotSave(mHandle, mSnapshotPath);
      if (rc != OK) {
        LOG.warn("could not save snapshot {}: {}", mSnapshotPath, nativeStrError(rc));
      }
    }
    nativeClose(mHandle);
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
  private static native int nativeSnapshotSave(long handle, String path);
  private static native int nativeSnapshotLoad(long handle, String path, long[] loadedAndRejected);
}
