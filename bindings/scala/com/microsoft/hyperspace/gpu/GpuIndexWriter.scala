/*
 * The index writer on the GPU. `write` has the signature and the effect of the block it replaces, the body of
 * CreateActionBase.write (src/main/scala/com/microsoft/hyperspace/actions/CreateActionBase.scala:122-140):
 *
 *   indexDataFrame.repartition(numBuckets, resolvedIndexedColumns...)      -> hs_build_index (K1 + K2 + K3)
 *   .write.saveWithBuckets(df, path, numBuckets, cols, SaveMode.Overwrite)  -> one Parquet file per non-empty
 *                                                                             bucket, named as Spark names it
 *
 * and is reached the same way from RefreshAction.op, RefreshIncrementalAction.op:58-71 and
 * OptimizeAction.op:85-99. Everything around it — Action.run's two-phase log protocol, IndexLogEntry, the
 * version directories — stays as it is: the writer only has to leave the same files behind.
 *
 * Compile-gated (no JVM in the development image); the same sequence of C-ABI calls is exercised by
 * hyperspace-1_b200/hyperspace.py + index_writer.py in tests/test_gpu_e2e.py.
 */
package com.microsoft.hyperspace.gpu

import java.nio.ByteBuffer
import java.util.UUID

import org.apache.arrow.vector.{FieldVector, VectorSchemaRoot}
import org.apache.spark.sql.{DataFrame, SaveMode}
import org.apache.spark.sql.types._

object GpuIndexWriter {

  private def hsType(t: DataType): Int = t match {
    case BooleanType => HsGpu.HS_BOOL
    case ByteType => HsGpu.HS_INT8
    case ShortType => HsGpu.HS_INT16
    case IntegerType | DateType => HsGpu.HS_INT32       // Spark hashes and orders DATE as its int (SURVEY A.2)
    case LongType | TimestampType => HsGpu.HS_INT64
    case FloatType => HsGpu.HS_FLOAT32
    case DoubleType => HsGpu.HS_FLOAT64
    case StringType => HsGpu.HS_UTF8
    case BinaryType => HsGpu.HS_BINARY
    case other => throw com.microsoft.hyperspace.HyperspaceException(s"hsgpu: unsupported column type $other")
  }

  /** hs_column array over the off-heap buffers of Arrow vectors (indexed columns first). */
  private[gpu] def describe(vectors: Seq[FieldVector], schema: StructType): ByteBuffer = {
    val buf = ByteBuffer.allocateDirect(vectors.size * HsGpu.COLUMN_BYTES)
    vectors.zipWithIndex.foreach { case (v, i) =>
      val t = hsType(schema(i).dataType)
      val validity = if (v.getNullCount > 0) v.getValidityBuffer.memoryAddress() else 0L
      val varlen = t == HsGpu.HS_UTF8 || t == HsGpu.HS_BINARY
      val offsets = if (varlen) v.getOffsetBuffer.memoryAddress() else 0L
      HsGpu.putColumn(buf, i, t, v.getValueCount, v.getNullCount, validity, offsets, v.getDataBuffer.memoryAddress())
    }
    buf
  }

  /**
   * @param indexDataFrame columns of the index: indexed ++ included (++ lineage), as prepareIndexDataFrame
   *                       returns them (CreateActionBase.scala:164-208)
   * @param indexedColumns resolved names of the indexed (bucket == sort) columns, first in the frame
   */
  def write(indexDataFrame: DataFrame, indexedColumns: Seq[String], numBuckets: Int, path: String,
      mode: SaveMode = SaveMode.Overwrite): Unit = {
    val schema = indexDataFrame.schema
    require(schema.fieldNames.take(indexedColumns.size).toSeq == indexedColumns, "indexed columns come first")
    val ctx = HsGpu.context()
    // rows -> Arrow off-heap vectors (Spark 2.4's ArrowConverters, the path toPandas() uses)
    val root: VectorSchemaRoot = ArrowCollector.collect(indexDataFrame)
    val table =
      try HsGpu.buildIndex(ctx, describe(ArrowCollector.vectors(root), schema), schema.size, indexedColumns.size, numBuckets)
      finally root.close()
    try {
      val offsets = HsGpu.tableBucketOffsets(table)
      val job = UUID.randomUUID().toString
      // one file per non-empty bucket: part-<task %05d>-<uuid>_<bucket %05d>.c000.snappy.parquet; after a create
      // the task id equals the bucket id (SURVEY A.4; BucketingUtils.getBucketId parses the bucket back,
      // OptimizeAction.scala:129)
      (0 until numBuckets).filter(b => offsets(b + 1) > offsets(b)).foreach { b =>
        val name = f"part-$b%05d-$job%s_$b%05d.c000.snappy.parquet"
        BucketFileWriter.write(s"$path/$name", schema, table, offsets(b), offsets(b + 1), mode)
      }
    } finally HsGpu.tableFree(ctx, table)
  }
}

object GpuIndexWriterAccess {
  def describe(vectors: Seq[FieldVector], schema: StructType): ByteBuffer = GpuIndexWriter.describe(vectors, schema)
}
