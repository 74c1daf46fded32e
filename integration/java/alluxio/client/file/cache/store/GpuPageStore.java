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

  /** One native store per GPU; a page lives in the store of GPU nativeShardOf(key, mHandles.length). */
  private final long[] mHandles;
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
    // alluxio.user.client.cache.gpu.device: a CUDA ordinal, or -1 for one shard of the cache on every visible GPU.
    // The cache stays ONE cache: LocalCacheManager, its meta store, its evictor and its byte budget are untouched.
    int[] devices;
    if (options.getDevice() >= 0) {
      devices = new int[] {options.getDevice()};
    } else {
      devices = new int[Math.max(1, nativeDeviceCount())];
      for (int d = 0; d < devices.length; d++) {
        devices[d] = d;
      }
    }
    int shards = devices.length;
    long pages = (mCacheSize + mPageSize - 1) / mPageSize;
    // hash routing balances only on average: every shard gets its share of the budget plus a quarter plus 64 slots
    long share = (pages + shards - 1) / shards;
    long slots = options.getSlabSlots() != 0 || shards == 1 ? options.getSlabSlots() : share + share / 4 + 64;
    long shardBytes = shards == 1 ? mCacheSize : slots * (long) mPageSize;
    mHandles = new long[shards];
    for (int i = 0; i < shards; i++) {
      long[] handle = new long[1];
      int rc = nativeOpen(devices[i], mPageSize, shardBytes, slots, options.getTableCapacity(),
          options.getMaxBatch(), options.isCrcEnabled() ? 1 : 0, handle);
      if (rc != OK) {
        for (int j = 0; j < i; j++) {
          nativeClose(mHandles[j]);
        }
        throw new IOException("aps_open failed on GPU " + devices[i] + ": " + nativeStrError(rc));
      }
      mHandles[i] = handle[0];
    }
    for (int i = 0; shards > 1 && i < shards; i++) {
      int rc = nativeShardAttachLocal(mHandles, i); // lets a batched get of shard 0 read every owner over NVLink
      if (rc != OK) {
        for (long handle : mHandles) {
          nativeClose(handle);
        }
        throw new IOException("aps_shard_attach_local failed: " + nativeStrError(rc));
      }
    }
    mStaging = ThreadLocal.withInitial(() -> allocPinned(mPageSize));
    // a snapshot is one file per store; with several shards the cache simply starts cold
    mSnapshotPath = shards == 1 ? options.getSnapshotPath() : null;
    if (mSnapshotPath != null && new java.io.File(mSnapshotPath).exists()) {
      // warm restart: getPages() then lists these pages and LocalCacheManager.restore adopts them.  Ids that are neither
      // decimal nor 32-hex were interned by the process that wrote the snapshot: take its table over first.
      String[] names = nativeSnapshotNames(mSnapshotPath);
      long[] counts = new long[2];
      int rc = names == null ? EINVAL : OK;
      if (rc == OK) {
        synchronized (mInternedNames) {
          for (String name : names) {
            mInterned.put(name, (long) mInternedNames.size());
            mInternedNames.add(name);
          }
        }
        rc = nativeSnapshotLoad(mHandles[0], mSnapshotPath, counts);
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

  /** The store that owns the page: hash(PageId) mod shards, the same routing the native side uses. */
  private long handleOf(long[] k, long pageIndex) {
    return mHandles.length == 1 ? mHandles[0]
        : mHandles[nativeShardOf(k[0], k[1], pageIndex, mHandles.length)];
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
    int rc = nativePut(handleOf(k, pageId.getPageIndex()), k[0], k[1], pageId.getPageIndex(), staging,
        page.length);
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
    int rc = nativeGet(handleOf(k, pageId.getPageIndex()), k[0], k[1], pageId.getPageIndex(), pageOffset,
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
    int rc = nativeDelete(handleOf(k, pageId.getPageIndex()), k[0], k[1], pageId.getPageIndex());
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
    List<PageInfo> pages = new ArrayList<>();
    for (long handle : mHandles) {
      long[] flat = nativeIterate(handle);
      if (flat == null) {
        throw new IOException("Failed to list the pages of the GPU store");
      }
      addPages(flat, pages);
    }
    return pages.stream();
  }

  private void addPages(long[] flat, List<PageInfo> pages) {
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
      int rc = nativeSnapshotSave(mHandles[0], mSnapshotPath, names);
      if (rc != OK) {
        LOG.warn("could not save snapshot {}: {}", mSnapshotPath, nativeStrError(rc));
      }
    }
    for (long handle : mHandles) {
      if (handle != 0) {
        nativeClose(handle);
      }
    }
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

  /** Owner shard of every request (3 longs each in keys). */
  private int[] ownersOf(long[] keys) {
    int[] owners = new int[keys.length / 3];
    for (int i = 0; mHandles.length > 1 && i < owners.length; i++) {
      owners[i] = nativeShardOf(keys[3 * i], keys[3 * i + 1], keys[3 * i + 2], mHandles.length);
    }
    return owners;
  }

  /** Positions of the requests whose page lives on the given shard. */
  private static int[] ownedBy(int[] owners, int shard) {
    int[] mine = new int[owners.length];
    int count = 0;
    for (int i = 0; i < owners.length; i++) {
      if (owners[i] == shard) {
        mine[count++] = i;
      }
    }
    return java.util.Arrays.copyOf(mine, count);
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
    // several shards: ONE kernel on shard 0 that probes every owner's table and reads its slab in place (NVLink)
    int rc = mHandles.length == 1
        ? nativeGetBatch(mHandles[0], n, keysOf(ids), pageOffsets, bytesToRead, dst, dstOffsets, outLen, crc)
        : nativeGetBatchSharded(mHandles[0], n, keysOf(ids), pageOffsets, bytesToRead, dst, dstOffsets, outLen, crc);
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
    long[] keys = keysOf(ids);
    int[] owners = ownersOf(keys);
    for (int shard = 0; shard < mHandles.length; shard++) { // puts go to the owner: one call per shard that owns any
      int[] mine = ownedBy(owners, shard);
      if (mine.length == 0) {
        continue;
      }
      long[] k = new long[3 * mine.length];
      long[] off = new long[mine.length];
      int[] len = new int[mine.length];
      int[] st = new int[mine.length];
      for (int j = 0; j < mine.length; j++) {
        System.arraycopy(keys, 3 * mine[j], k, 3 * j, 3);
        off[j] = srcOffsets[mine[j]];
        len[j] = lengths[mine[j]];
      }
      int rc = nativePutBatch(mHandles[shard], mine.length, k, src, off, len, st, new int[mine.length]);
      if (rc != OK) {
        throw new IOException("Failed to put " + mine.length + " pages: " + nativeStrError(rc));
      }
      for (int j = 0; j < mine.length; j++) {
        status[mine[j]] = st[j];
      }
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
    long[] keys = keysOf(ids);
    int[] owners = ownersOf(keys);
    for (int shard = 0; shard < mHandles.length; shard++) {
      int[] mine = ownedBy(owners, shard);
      if (mine.length == 0) {
        continue;
      }
      long[] k = new long[3 * mine.length];
      int[] st = new int[mine.length];
      for (int j = 0; j < mine.length; j++) {
        System.arraycopy(keys, 3 * mine[j], k, 3 * j, 3);
      }
      int rc = nativeDeleteBatch(mHandles[shard], mine.length, k, st);
      if (rc != OK) {
        throw new IOException("Failed to delete " + mine.length + " pages: " + nativeStrError(rc));
      }
      for (int j = 0; j < mine.length; j++) {
        status[mine[j]] = st[j];
      }
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
  private static native int nativeDeviceCount();
  private static native int nativeShardOf(long fileHi, long fileLo, long pageIndex, int shards);
  private static native int nativeShardAttachLocal(long[] handles, int myShard);
  private static native int nativeGetBatchSharded(long handle, int n, long[] keys, int[] pageOffsets, int[] bytesToRead,
      ByteBuffer buffer, long[] offsets, int[] outLen, int[] outCrc);
}
