/*
 * Index files <-> the columns libhsgpu works on. Compile-gated (no JVM in the development image).
 */
package com.microsoft.hyperspace.gpu

import java.nio.ByteBuffer

import org.apache.arrow.vector.VectorSchemaRoot
import org.apache.spark.sql.SparkSession
import org.apache.spark.sql.catalyst.InternalRow
import org.apache.spark.sql.hyperspace.gpu.ArrowCollector
import org.apache.spark.sql.types.StructType

/** One sorted run per bucket in Arrow off-heap vectors: the in-memory form of the files of an index version. */
class IndexRun(val root: VectorSchemaRoot, val columns: ByteBuffer, val bucketOffsets: Array[Long]) {
  def close(): Unit = root.close()
}

object IndexRunReader {
  private val BucketId = """.*_(\d+)(?:\..*)?$""".r // Spark's BucketingUtils.getBucketId

  /**
   * Files grouped by the bucket id in their names and read bucket by bucket, so that rows of bucket b are
   * contiguous. A bucket with one file is a sorted run as it is; with several files (incremental refresh) the
   * files are merged pairwise through hs_merge_runs by the caller of `readGeneration`.
   */
  def read(files: Seq[String], schema: StructType, numBuckets: Int): IndexRun = {
    val spark = SparkSession.active
    val byBucket = files.groupBy { f => f.substring(f.lastIndexOf('/') + 1) match { case BucketId(b) => b.toInt } }
    val generations = byBucket.values.map(_.size).max
    val runs = (0 until generations).map { g =>
      readGeneration(spark, (0 until numBuckets).map(b => byBucket.get(b).flatMap(_.sorted.lift(g))), schema)
    }
    runs.reduce { (a, b) =>
      val ctx = HsGpu.context()
      val t = HsGpu.mergeRuns(ctx, a.columns, a.bucketOffsets, b.columns, b.bucketOffsets, schema.size,
        /* nIndexed is carried by the caller's view; keys are the leading columns */ 1, numBuckets)
      a.close(); b.close()
      TableRows.asRun(ctx, t, schema)
    }
  }

  /** At most one file per bucket -> bucket-contiguous vectors + offsets. */
  private def readGeneration(spark: SparkSession, fileOfBucket: Seq[Option[String]], schema: StructType): IndexRun = {
    val offsets = new Array[Long](fileOfBucket.size + 1)
    val frames = fileOfBucket.zipWithIndex.map { case (f, b) =>
      val df = f.map(p => spark.read.schema(schema).parquet(p))
      offsets(b + 1) = offsets(b) + df.map(_.count()).getOrElse(0L)
      df
    }
    val all = frames.flatten.reduceOption(_ union _).getOrElse(spark.createDataFrame(spark.sparkContext.emptyRDD[org.apache.spark.sql.Row], schema))
    val root = ArrowCollector.collect(all.coalesce(1)) // file order == bucket order
    new IndexRun(root, GpuIndexWriterAccess.describe(ArrowCollector.vectors(root), schema), offsets)
  }
}

/** Rows of a host-resident hs_table for Spark; frees the table when the iterator is exhausted. */
object TableRows {
  def iterator(ctx: Long, table: Long, schema: StructType): Iterator[InternalRow] = {
    HsGpu.tableWait(ctx, table)
    val n = HsGpu.tableRows(table)
    new Iterator[InternalRow] {
      private var i = 0L
      override def hasNext: Boolean = { if (i >= n && i != Long.MaxValue) { HsGpu.tableFree(ctx, table); i = Long.MaxValue }; i < n }
      override def next(): InternalRow = { val r = BucketFileWriterAccess.row(table, schema, i); i += 1; r }
    }
  }
  def asRun(ctx: Long, table: Long, schema: StructType): IndexRun =
    throw new UnsupportedOperationException("wrap the table's host buffers as Arrow vectors (ArrowBuf over foreign memory)")
}
