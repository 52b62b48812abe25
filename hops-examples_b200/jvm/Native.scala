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

  // ---- multi-GPU: the driver thread calls each operator once, the exchange between the GPUs runs inside libfsx ----------
  val ROUTE_AUTO = 0; val ROUTE_DENSE = 1; val ROUTE_PARTIAL = 2; val ROUTE_ROWS = 3
  val SHUFFLE = 0; val BROADCAST = 1
  /** fsx_ctx_create(devices, n): one context per device + an NCCL communicator; the returned handle is member 0 */
  @native def ctxCreateGroup(devices: Array[Int]): Long
  /** fsx_ctx_member: the context of devices(i), for columnUpload / columnDownload of that device's shard */
  @native def ctxMember(ctx: Long, i: Int): Long
  /** one executor JVM per GPU instead: rank 0 calls commUniqueId, every rank commInitRank with the same 128 bytes */
  @native def commUniqueId(): Array[Byte]
  @native def commInitRank(ctx: Long, uniqueId: Array[Byte], rank: Int, world: Int): Unit
  /** fsx_dist_groupby_agg: keys(m) / vals(m) = member m's shard -> one table (key, sum, avg, count, min, max) per member */
  @native def distGroupbyAgg(ctx: Long, keys: Array[Long], vals: Array[Long], route: Int, expectedGroups: Long): Array[Long]
  /** fsx_dist_groupby_host: the same for batches in (pinned) direct buffers, one per member */
  @native def distGroupbyHost(ctx: Long, keyDtype: Int, keys: Array[ByteBuffer], valDtype: Int, vals: Array[ByteBuffer],
                              n: Array[Long], expectedGroups: Long): Array[Long]
  /** fsx_dist_join_groupby_sum -> one table (g, sum, count) per member (the full result on each) */
  @native def distJoinGroupbySum(ctx: Long, factKey: Array[Long], factVal: Array[Long], dimKey: Array[Long], dimGroup: Array[Long],
                                 variant: Int): Array[Long]
  /** fsx_dist_hash_join: left / right = nLeft / nRight columns per member, flattened member by member; column 0 is the key */
  @native def distHashJoin(ctx: Long, left: Array[Long], nLeft: Int, right: Array[Long], nRight: Int, variant: Int): Array[Long]
  /** fsx_dist_stats_profile: cols = ncols columns per member, flattened member by member; buffers as statsProfile */
  @native def distStatsProfile(ctx: Long, cols: Array[Long], ncols: Int, nbins: Int, moments: ByteBuffer, counts: ByteBuffer,
                               edges: ByteBuffer, nbuckets: ByteBuffer, corr: ByteBuffer): Unit

  /** fsx_dist_stats_quantiles: exact order statistics of one column sharded over the members (cols(m) = member m's shard) */
  @native def distStatsQuantiles(ctx: Long, cols: Array[Long], probs: Array[Double]): Array[Double]
  /** fsx_dist_stats_approx_distinct: the members' HyperLogLog sketches merged by maximum */
  @native def distStatsApproxDistinct(ctx: Long, cols: Array[Long]): Long

  // ---- scans and string columns (INTEGRATION.md 4e) ------------------------------------------------------------------
  /** fsx_stats_approx_distinct: Deequ's approximateNumDistinctValues (HyperLogLog, 4096 registers) */
  @native def statsApproxDistinct(ctx: Long, col: Long): Long
  /** fsx_csv_decode: the bytes after the header line -> table handle; typesOut(c) = 1 int, 2 bigint, 3 double, 4 string */
  @native def csvDecode(ctx: Long, body: ByteBuffer, nbytes: Long, ncols: Int, typesOut: Array[Int]): Long
  /** fsx_parquet_schema: "name:type,..." of a .parquet file held in a direct buffer ('?' marks an OPTIONAL column) */
  @native def parquetSchema(file: ByteBuffer, nbytes: Long): String
  /** fsx_parquet_decode_column: an int32 / int64 / float / double column of the file -> column handle */
  @native def parquetDecodeColumn(ctx: Long, file: ByteBuffer, nbytes: Long, column: String): Long
  /** fsx_parquet_decode_column_files: the same over the part files of one table (rows run through the files in order) */
  @native def parquetDecodeColumnFiles(ctx: Long, files: Array[ByteBuffer], nbytes: Array[Long], column: String): Long
  /** fsx_parquet_decode_strings: a dictionary-encoded string column -> handle of the i32 codes; the dictionary as UTF-8 in
    * dictBytes with dictOffsets (longs); sizes(0) = distinct strings, sizes(1) = bytes.  Returns 0 when a buffer was too
    * small: allocate from sizes and call again. */
  @native def parquetDecodeStrings(ctx: Long, file: ByteBuffer, nbytes: Long, column: String, dictBytes: ByteBuffer,
                                   dictOffsets: ByteBuffer, sizes: Array[Long]): Long

  // ---- the remaining single-GPU operators a driver may want (INTEGRATION.md 4a, 4e) -----------------------------------
  /** fsx_join_groupby_sum: fact JOIN dim ON factKey = dimKey GROUP BY dimGroup -> table (g, sum(factVal), count), never
    * materialising the join */
  @native def joinGroupbySum(ctx: Long, factKey: Long, factVal: Long, dimKey: Long, dimGroup: Long, expectedGroups: Long): Long
  /** fsx_dictionary_encode_utf8 (host code): n strings as UTF-8 bytes + n + 1 offsets (longs) + validity bitmap (or null)
    * -> codesOut (n ints, first-appearance order), firstRowOut (row of each distinct string); returns the distinct count */
  @native def dictionaryEncodeUtf8(bytes: ByteBuffer, offsets: ByteBuffer, validity: ByteBuffer, n: Long, codesOut: ByteBuffer,
                                   firstRowOut: ByteBuffer): Int
  /** fsx_last_timing of the previous call on ctx: (totalMs, h2dMs, kernelMs, d2hMs, mainKernelMs, h2dBytes, d2hBytes, launches) */
  @native def lastTiming(ctx: Long): Array[Double]
  /** fsx_groupby_sum_count_noalloc: for an executor whose device arena already owns its HBM -- every argument is a DEVICE
    * address inside that arena (keys int32, vals float64, outputs sized maxGroups <= noallocMaxGroups, ngroupsOut one long),
    * stream = a cudaStream_t or 0; the library allocates nothing (its group table is a fixed array on the GPU) */
  @native def groupbySumCountNoalloc(keysDev: Long, valsDev: Long, n: Long, maxGroups: Long, keysOutDev: Long, sumOutDev: Long,
                                     countOutDev: Long, ngroupsOutDev: Long, stream: Long): Unit
  @native def noallocMaxGroups(): Int

  def direct(bytes: Long): ByteBuffer = ByteBuffer.allocateDirect(bytes.toInt).order(ByteOrder.nativeOrder())
}
