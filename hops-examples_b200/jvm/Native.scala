package io.hops.examples.featurestore_tour.native

import java.nio.{ByteBuffer, ByteOrder}

/** JNI surface over include/fsx.h for the featurestore_tour job (NOT compiled in this repository's
  * image: no JDK is available there; the C ABI underneath is what the tests exercise).
  *
  * Handles are opaque longs.  Buffers are direct, native-order ByteBuffers owned by the caller; nothing
  * is pinned across calls.  Every method throws RuntimeException(fsx_last_error) when the library
  * returns a non-zero status -- there is no CPU fallback behind these calls.
  */
object Native {
  System.loadLibrary("fsx_jni") // libfsx_jni.so, linked against libfsx.so

  // dtype / aggregate / join codes of include/fsx.h
  val I32 = 0; val I64 = 1; val F32 = 2; val F64 = 3
  val SUM = 0; val COUNT = 1; val COUNT_STAR = 2; val MIN = 3; val MAX = 4; val AVG = 5
  val INNER = 0; val LEFT_SEMI = 1; val LEFT_ANTI = 2

  @native def ctxCreate(device: Int): Long
  @native def ctxDestroy(ctx: Long): Unit
  @native def columnUpload(ctx: Long, dtype: Int, n: Long, data: ByteBuffer, validity: ByteBuffer): Long
  @native def columnDownload(ctx: Long, col: Long, data: ByteBuffer, validity: ByteBuffer): Unit
  @native def columnLength(col: Long): Long
  @native def columnFree(ctx: Long, col: Long): Unit
  @native def groupbyAgg(ctx: Long, key: Long, vals: Array[Long], aggs: Array[Int], expectedGroups: Long): Long
  @native def tableSortBy(ctx: Long, table: Long, column: Int): Unit
  @native def tableNrows(table: Long): Long
  @native def tableColumn(table: Long, i: Int): Long
  @native def tableFree(ctx: Long, table: Long): Unit
  @native def finalizeF32(ctx: Long, sum: Long, count: Long): Long
  @native def hashJoin(ctx: Long, leftKey: Long, rightKey: Long, joinType: Int): Long
  @native def joinGather(ctx: Long, join: Long, side: Int, col: Long): Long
  @native def joinFree(ctx: Long, join: Long): Unit
  /** moments: ncols * 56 bytes (fsx_moments); counts: ncols*nbins longs; edges: ncols*(nbins+1) doubles;
    * nbuckets: ncols ints; corr: ncols*ncols doubles (null = skip that part) */
  @native def statsProfile(ctx: Long, cols: Array[Long], nbins: Int, moments: ByteBuffer, counts: ByteBuffer,
                           edges: ByteBuffer, nbuckets: ByteBuffer, corr: ByteBuffer): Unit

  /** fsx_groupby_host: a batch in (pinned) direct buffers -> table (key, sum, avg, count, min, max); 0 when the library
    * declines the buffers (pageable memory, integer values, sparse keys) and the caller takes columnUpload + groupbyAgg */
  @native def groupbyHost(ctx: Long, keyDtype: Int, keys: ByteBuffer, valDtype: Int, vals: ByteBuffer, n: Long, chunkRows: Long): Long
  /** fsx_stats_distinct -> (nRows, nNonNull, nDistinct, nUnique) in the first four longs, entropy bits in the fifth */
  @native def statsDistinct(ctx: Long, col: Long, out5: ByteBuffer): Unit
  @native def statsQuantiles(ctx: Long, col: Long, probs: Array[Double]): Array[Double]
  @native def statsValueCounts(ctx: Long, col: Long): Long
  @native def columnCombineKeys(ctx: Long, a: Long, b: Long): Long
  @native def groupbyCountDistinct(ctx: Long, key: Long, value: Long): Long
  /** sum(value) over (partition by partition order by order range between lo and hi); partition = 0 for none */
  @native def windowRangeSum(ctx: Long, partition: Long, order: Long, value: Long, lo: Long, hi: Long): Long

  def direct(bytes: Long): ByteBuffer = ByteBuffer.allocateDirect(bytes.toInt).order(ByteOrder.nativeOrder())
}
