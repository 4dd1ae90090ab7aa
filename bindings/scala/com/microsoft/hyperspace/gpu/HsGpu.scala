/*
 * The JVM face of libhsgpu: native method declarations matching bindings/jni/hsgpu_jni.c one to one, and the
 * per-(thread, device) context registry that mirrors the reference's thread-local HyperspaceContext
 * (src/main/scala/com/microsoft/hyperspace/Hyperspace.scala:168-204).
 *
 * Compile-gated: it needs Scala 2.11/2.12, Spark 2.4.2 and a JDK, none of which exist in the development image;
 * the tested boundary is the C ABI (include/hsgpu.h) through the ctypes binding hyperspace-1_b200/native.py.
 */
package com.microsoft.hyperspace.gpu

import java.nio.{ByteBuffer, ByteOrder}

object HsGpu {
  System.loadLibrary("hsgpu_jni")

  // hs_type / hs_mem / hs_join_type of include/hsgpu.h
  val HS_BOOL = 1; val HS_INT8 = 2; val HS_INT16 = 3; val HS_INT32 = 4; val HS_INT64 = 5
  val HS_FLOAT32 = 6; val HS_FLOAT64 = 7; val HS_UTF8 = 8; val HS_BINARY = 9
  val HS_MEM_HOST = 0; val HS_MEM_DEVICE = 1
  val JOIN_INNER = 0; val JOIN_LEFT_OUTER = 1; val JOIN_RIGHT_OUTER = 2; val JOIN_FULL_OUTER = 3
  val JOIN_LEFT_SEMI = 4; val JOIN_LEFT_ANTI = 5
  val COLUMN_BYTES = 48 // sizeof(hs_column)

  @native def ctxCreate(device: Int): Long
  @native def ctxDestroy(ctx: Long): Unit
  @native def ctxSetOutputLocation(ctx: Long, location: Int): Unit
  @native def ctxSetDeferredResults(ctx: Long, on: Boolean): Unit
  @native def hashBucket(ctx: Long, keys: ByteBuffer, nkeys: Int, numBuckets: Int, outIds: ByteBuffer): Unit
  @native def buildIndex(ctx: Long, cols: ByteBuffer, ncols: Int, nIndexed: Int, numBuckets: Int): Long
  @native def commUniqueId(ctx: Long): Array[Byte]
  @native def commInit(ctx: Long, rank: Int, nranks: Int, id128: Array[Byte]): Unit
  @native def commDestroy(ctx: Long): Unit
  @native def buildIndexDist(ctx: Long, cols: ByteBuffer, ncols: Int, nIndexed: Int, numBuckets: Int): Long
  @native def mergeRuns(ctx: Long, colsA: ByteBuffer, offA: Array[Long], colsB: ByteBuffer, offB: Array[Long],
      ncols: Int, nIndexed: Int, numBuckets: Int): Long
  @native def joinBucketed(ctx: Long, lcols: ByteBuffer, lncols: Int, loff: Array[Long], rcols: ByteBuffer,
      rncols: Int, roff: Array[Long], nIndexed: Int, lBuckets: Int, rBuckets: Int, joinType: Int,
      leftOut: Array[Int], rightOut: Array[Int]): Long
  @native def filterScan(ctx: Long, cols: ByteBuffer, ncols: Int, off: Array[Long], nIndexed: Int,
      numBuckets: Int, terms: ByteBuffer, nterms: Int, outCols: Array[Int]): Long
  @native def parquetEncodePages(ctx: Long, col: ByteBuffer, optional: Boolean, pageRows: Array[Long],
      pageChunk: Array[Int], nchunks: Int, maxDict: Int, codec: Int): Long
  @native def tableWait(ctx: Long, table: Long): Unit
  @native def tableFree(ctx: Long, table: Long): Unit
  @native def tableRows(table: Long): Long
  @native def tableColumns(table: Long): Int
  @native def tableBucketOffsets(table: Long): Array[Long]
  @native def tableColumn(table: Long, i: Int): Array[Long] // type, length, nulls, validity, offsets, data
  @native def wrap(address: Long, bytes: Long): ByteBuffer

  /** One context per (thread, device): a context is not thread-safe, distinct contexts run concurrently. */
  private val contexts = new ThreadLocal[scala.collection.mutable.Map[Int, Long]] {
    override def initialValue(): scala.collection.mutable.Map[Int, Long] = scala.collection.mutable.Map.empty
  }
  def context(device: Int = 0): Long = contexts.get.getOrElseUpdate(device, ctxCreate(device))

  /** Writes one hs_column (include/hsgpu.h: 48 bytes) at `index` of a direct little-endian buffer. */
  def putColumn(buf: ByteBuffer, index: Int, hsType: Int, length: Long, nullCount: Long, validity: Long,
      offsets: Long, data: Long): Unit = {
    val b = buf.order(ByteOrder.LITTLE_ENDIAN)
    val at = index * COLUMN_BYTES
    b.putInt(at, hsType); b.putInt(at + 4, 0)
    b.putLong(at + 8, length); b.putLong(at + 16, nullCount)
    b.putLong(at + 24, validity); b.putLong(at + 32, offsets); b.putLong(at + 40, data)
  }
}
