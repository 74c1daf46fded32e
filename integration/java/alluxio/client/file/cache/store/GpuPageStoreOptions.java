/*
 * Options of the GPU page store (new file; sits beside LocalPageStoreOptions).
 * Not compiled in this repository: the build image has no JDK (see INTEGRATION.md).
 */
package alluxio.client.file.cache.store;

import com.google.common.base.MoreObjects;

/** Options used to instantiate {@link GpuPageStore}. */
public class GpuPageStoreOptions extends PageStoreOptions {
  private int mDevice = 0;
  private long mSlabSlots = 0;      // 0 = derive from cache size (+12.5% for short last pages)
  private long mTableCapacity = 0;  // 0 = 2 x slots, rounded up to a power of two
  private int mMaxBatch = 4096;
  private boolean mCrcEnabled = true;
  /** alluxio.user.client.cache.gpu.snapshot.path: where close() saves and open loads the pages; null = volatile. */
  private String mSnapshotPath = null;

  public GpuPageStoreOptions setDevice(int device) { mDevice = device; return this; }
  public GpuPageStoreOptions setSlabSlots(long slots) { mSlabSlots = slots; return this; }
  public GpuPageStoreOptions setTableCapacity(long entries) { mTableCapacity = entries; return this; }
  public GpuPageStoreOptions setMaxBatch(int n) { mMaxBatch = n; return this; }
  public GpuPageStoreOptions setCrcEnabled(boolean on) { mCrcEnabled = on; return this; }
  public GpuPageStoreOptions setSnapshotPath(String path) { mSnapshotPath = path; return this; }
  public String getSnapshotPath() { return mSnapshotPath; }
  public int getDevice() { return mDevice; }
  public long getSlabSlots() { return mSlabSlots; }
  public long getTableCapacity() { return mTableCapacity; }
  public int getMaxBatch() { return mMaxBatch; }
  public boolean isCrcEnabled() { return mCrcEnabled; }

  @Override
  public PageStoreType getType() { return PageStoreType.GPU; }

  @Override
  public String toString() {
    return MoreObjects.toStringHelper(this).add("CacheSize", mCacheSize).add("PageSize", mPageSize)
        .add("Device", mDevice).add("SlabSlots", mSlabSlots).add("TableCapacity", mTableCapacity)
        .add("MaxBatch", mMaxBatch).add("Crc", mCrcEnabled).toString();
  }
}
