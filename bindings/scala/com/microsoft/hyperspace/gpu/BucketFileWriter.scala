/*
 * Rows [from, until) of a host-resident hs_table written as one Parquet file through Spark's own
 * ParquetFileFormat writer — the same encoder (snappy, parquet-mr 1.10, Spark's schema metadata) that
 * saveWithBuckets reaches (src/main/scala/com/microsoft/hyperspace/index/DataFrameWriterExtensions.scala:57-66),
 * so index files stay readable by the unmodified reference. Compile-gated.
 */
package com.microsoft.hyperspace.gpu

import org.apache.hadoop.mapreduce.{Job, TaskAttemptID}
import org.apache.hadoop.mapreduce.task.TaskAttemptContextImpl
import org.apache.spark.sql.{SaveMode, SparkSession}
import org.apache.spark.sql.catalyst.InternalRow
import org.apache.spark.sql.catalyst.expressions.GenericInternalRow
import org.apache.spark.sql.execution.datasources.parquet.ParquetFileFormat
import org.apache.spark.sql.types._
import org.apache.spark.unsafe.Platform
import org.apache.spark.unsafe.types.UTF8String

object BucketFileWriter {
  private case class Col(t: Int, validity: Long, offsets: Long, data: Long)

  /** Value of row i of a column, read straight from the table's pinned host buffers. */
  private[gpu] def valueAt(validity: Long, offsets: Long, data: Long, dt: DataType, i: Long): Any =
    value(Col(0, validity, offsets, data), dt, i)

  private def value(c: Col, dt: DataType, i: Long): Any = {
    if (c.validity != 0 && ((Platform.getByte(null, c.validity + (i >> 3)) >> (i & 7)) & 1) == 0) return null
    dt match {
      case BooleanType => ((Platform.getByte(null, c.data + (i >> 3)) >> (i & 7)) & 1) != 0
      case ByteType => Platform.getByte(null, c.data + i)
      case ShortType => Platform.getShort(null, c.data + 2 * i)
      case IntegerType | DateType => Platform.getInt(null, c.data + 4 * i)
      case LongType | TimestampType => Platform.getLong(null, c.data + 8 * i)
      case FloatType => Platform.getFloat(null, c.data + 4 * i)
      case DoubleType => Platform.getDouble(null, c.data + 8 * i)
      case StringType | BinaryType =>
        val a = Platform.getInt(null, c.offsets + 4 * i)
        val b = Platform.getInt(null, c.offsets + 4 * (i + 1))
        val bytes = new Array[Byte](b - a)
        Platform.copyMemory(null, c.data + a, bytes, Platform.BYTE_ARRAY_OFFSET, b - a)
        if (dt == StringType) UTF8String.fromBytes(bytes) else bytes
    }
  }

  def write(file: String, schema: StructType, table: Long, from: Long, until: Long, mode: SaveMode): Unit = {
    val spark = SparkSession.active
    val job = Job.getInstance(spark.sessionState.newHadoopConf())
    val factory = new ParquetFileFormat().prepareWrite(spark, job, Map.empty, schema)
    val attempt = new TaskAttemptContextImpl(job.getConfiguration, new TaskAttemptID())
    val out = factory.newInstance(file, schema, attempt)
    val cols = (0 until schema.size).map { i =>
      val d = HsGpu.tableColumn(table, i)
      Col(d(0).toInt, d(3), d(4), d(5))
    }
    try {
      val row = new GenericInternalRow(schema.size)
      var r = from
      while (r < until) {
        var c = 0
        while (c < schema.size) { row.update(c, value(cols(c), schema(c).dataType, r)); c += 1 }
        out.write(row: InternalRow)
        r += 1
      }
    } finally out.close()
  }
}

object BucketFileWriterAccess {
  /** Row i of a host-resident table as an InternalRow (a copy). */
  def row(table: Long, schema: StructType, i: Long): InternalRow = {
    val r = new GenericInternalRow(schema.size)
    var c = 0
    while (c < schema.size) {
      val d = HsGpu.tableColumn(table, c)
      r.update(c, BucketFileWriter.valueAt(d(3), d(4), d(5), schema(c).dataType, i))
      c += 1
    }
    r
  }
}
