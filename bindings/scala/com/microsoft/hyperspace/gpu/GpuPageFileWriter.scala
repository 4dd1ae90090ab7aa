/*
 * The index files of one version written from page bodies the device encoded (hs_parquet_encode_pages through
 * HsGpu.parquetEncodePages): the JVM counterpart of hyperspace-1_b200/parquet_plain.py. Where BucketFileWriter
 * pushes every value through Spark's ParquetFileFormat writer, this one only frames bytes: per bucket file `PAR1`,
 * per column an optional DICTIONARY_PAGE and the DATA_PAGEs (Thrift PageHeader from parquet-format, which Spark
 * 2.4 ships), the FileMetaData footer with Spark's schema JSON under the key ParquetReadSupport looks for, the
 * footer length and `PAR1`. File names and the one-file-per-non-empty-bucket layout are those of saveWithBuckets
 * (src/main/scala/com/microsoft/hyperspace/index/DataFrameWriterExtensions.scala:57-66); the bucket id is parsed
 * back from the name by BucketingUtils.getBucketId. Compile-gated (no JDK in the development image); the page
 * layout itself is pinned by tests/test_host_cpu.py and tests/test_gpu_parquet.py against pyarrow's reader.
 */
package com.microsoft.hyperspace.gpu

import java.io.OutputStream
import java.nio.ByteBuffer
import java.util.{ArrayList => JList, Arrays, UUID}

import org.apache.hadoop.fs.Path
import org.apache.parquet.format._
import org.apache.spark.sql.SparkSession
import org.apache.spark.sql.types._
import org.apache.spark.unsafe.Platform

object GpuPageFileWriter {
  val PageRows: Long = 1L << 20
  private val Magic = "PAR1".getBytes("US-ASCII")

  /** Parquet physical type and converted type of a Spark type (the ones hs_parquet_encode_pages takes). */
  private def physical(dt: DataType): (Type, Option[ConvertedType]) = dt match {
    case BooleanType => (Type.BOOLEAN, None)
    case ByteType => (Type.INT32, Some(ConvertedType.INT_8))
    case ShortType => (Type.INT32, Some(ConvertedType.INT_16))
    case IntegerType => (Type.INT32, None)
    case DateType => (Type.INT32, Some(ConvertedType.DATE))
    case LongType => (Type.INT64, None)
    case TimestampType => (Type.INT64, Some(ConvertedType.TIMESTAMP_MICROS))
    case FloatType => (Type.FLOAT, None)
    case DoubleType => (Type.DOUBLE, None)
    case StringType => (Type.BYTE_ARRAY, Some(ConvertedType.UTF8))
    case BinaryType => (Type.BYTE_ARRAY, None)
    case other => throw new UnsupportedOperationException(s"no Parquet mapping for $other")
  }

  /** Bytes [from, until) of a BINARY result column (offsets int32) as a fresh array. */
  private def body(offsets: Long, data: Long, row: Int): Array[Byte] = {
    val a = Platform.getInt(null, offsets + 4L * row)
    val b = Platform.getInt(null, offsets + 4L * (row + 1))
    val out = new Array[Byte](b - a)
    Platform.copyMemory(null, data + a, out, Platform.BYTE_ARRAY_OFFSET, b - a)
    out
  }

  private final class Encoded(val pages: Array[Array[Byte]], val dicts: Array[Array[Byte]], val entries: Array[Long],
      val usize: Array[Long])

  /**
   * table: a host-resident hs_table holding the sorted index (column i = schema(i)); bucketOffsets: rows of bucket
   * b are [off(b), off(b + 1)). columns(i): the hs_column struct of column i as a direct ByteBuffer (what
   * GpuIndexWriter already builds for hs_build_index's result; INT8 / INT16 columns widened to INT32 first).
   * maxDict = 0: PLAIN only; snappy: Spark's default codec. Returns the files written.
   */
  def write(spark: SparkSession, ctx: Long, schema: StructType, columns: IndexedSeq[ByteBuffer], bucketOffsets: Array[Long],
      dir: String, maxDict: Int = 4096, snappy: Boolean = true): Seq[String] = {
    val nb = bucketOffsets.length - 1
    // pages never cross a bucket: one device call per column encodes the pages of all files
    val pageRows = new scala.collection.mutable.ArrayBuffer[Long]
    val pageChunk = new scala.collection.mutable.ArrayBuffer[Int]
    val chunkBucket = new scala.collection.mutable.ArrayBuffer[Int]
    pageRows += 0L
    for (b <- 0 until nb if bucketOffsets(b + 1) > bucketOffsets(b)) {
      var a = bucketOffsets(b)
      while (a < bucketOffsets(b + 1)) {
        a = math.min(bucketOffsets(b + 1), a + PageRows)
        pageRows += a
        pageChunk += chunkBucket.length
      }
      chunkBucket += b
    }
    if (chunkBucket.isEmpty) return Seq.empty
    val npages = pageChunk.length
    val nchunks = chunkBucket.length
    val codec = if (snappy) 1 else 0
    val encoded = columns.map { col =>
      val t = HsGpu.parquetEncodePages(ctx, col, true, pageRows.toArray, pageChunk.toArray, nchunks, maxDict, codec)
      try {
        HsGpu.tableWait(ctx, t)
        val bodies = HsGpu.tableColumn(t, 0)          // type, length, nulls, validity, offsets, data
        val counts = HsGpu.tableColumn(t, 1)
        val rows = (0 until npages + nchunks).map(r => body(bodies(4), bodies(5), r)).toArray
        val cnt = (0 until npages + nchunks).map(r => Platform.getLong(null, counts(5) + 8L * r)).toArray
        val us =
          if (snappy) {
            val u = HsGpu.tableColumn(t, 2)
            (0 until npages + nchunks).map(r => Platform.getLong(null, u(5) + 8L * r)).toArray
          } else rows.map(_.length.toLong)
        new Encoded(rows.take(npages), rows.drop(npages), cnt.drop(npages), us)
      } finally HsGpu.tableFree(ctx, t)
    }
    val job = UUID.randomUUID()
    val suffix = if (snappy) ".c000.snappy.parquet" else ".c000.parquet"
    val fs = new Path(dir).getFileSystem(spark.sessionState.newHadoopConf())
    val firstPage = new Array[Int](nchunks + 1)
    for (p <- 0 until npages) firstPage(pageChunk(p) + 1) = p + 1
    for (c <- 1 to nchunks if firstPage(c) == 0) firstPage(c) = firstPage(c - 1)
    (0 until nchunks).map { chunk =>
      val b = chunkBucket(chunk)
      val path = new Path(dir, f"part-$b%05d-$job%s_$b%05d$suffix%s")
      val out = fs.create(path, false)
      try {
        out.write(Magic)
        var pos = 4L
        val chunks = new JList[ColumnChunk]
        for (ci <- schema.indices) {
          val e = encoded(ci)
          val (phys, _) = physical(schema(ci).dataType)
          val start = pos
          var upos = 0L
          var values = 0L
          var dataOff = -1L
          val entries = e.entries(chunk).toInt
          if (entries > 0) {
            val d = e.dicts(chunk)
            val h = new PageHeader(PageType.DICTIONARY_PAGE, e.usize(npages + chunk).toInt, d.length)
            h.setDictionary_page_header(new DictionaryPageHeader(entries, Encoding.PLAIN_DICTIONARY))
            val hl = writeHeader(h, out)
            out.write(d)
            pos += hl + d.length
            upos += hl + e.usize(npages + chunk)
            dataOff = pos
          }
          for (p <- firstPage(chunk) until firstPage(chunk + 1)) {
            val pg = e.pages(p)
            val n = (pageRows(p + 1) - pageRows(p)).toInt
            val h = new PageHeader(PageType.DATA_PAGE, e.usize(p).toInt, pg.length)
            h.setData_page_header(new DataPageHeader(n, if (entries > 0) Encoding.PLAIN_DICTIONARY else Encoding.PLAIN,
              Encoding.RLE, Encoding.BIT_PACKED))
            val hl = writeHeader(h, out)
            out.write(pg)
            pos += hl + pg.length
            upos += hl + e.usize(p)
            values += n
          }
          val encs = Arrays.asList(if (entries > 0) Encoding.PLAIN_DICTIONARY else Encoding.PLAIN, Encoding.RLE, Encoding.BIT_PACKED)
          val md = new ColumnMetaData(phys, encs, Arrays.asList(schema(ci).name),
            if (snappy) CompressionCodec.SNAPPY else CompressionCodec.UNCOMPRESSED, values, upos, pos - start,
            if (dataOff >= 0) dataOff else start)
          if (dataOff >= 0) md.setDictionary_page_offset(start)
          val cc = new ColumnChunk(start)
          cc.setMeta_data(md)
          chunks.add(cc)
        }
        val rows = bucketOffsets(b + 1) - bucketOffsets(b)
        val elems = new JList[SchemaElement]
        val root = new SchemaElement("spark_schema")
        root.setNum_children(schema.size)
        elems.add(root)
        for (f <- schema) {
          val (phys, conv) = physical(f.dataType)
          val se = new SchemaElement(f.name)
          se.setType(phys)
          se.setRepetition_type(FieldRepetitionType.OPTIONAL)
          conv.foreach(se.setConverted_type)
          elems.add(se)
        }
        val rg = new RowGroup(chunks, chunks.toArray(new Array[ColumnChunk](0)).map(_.getMeta_data.getTotal_uncompressed_size).sum, rows)
        val fmd = new FileMetaData(1, elems, rows, Arrays.asList(rg))
        val kv = new KeyValue("org.apache.spark.sql.parquet.row.metadata")
        kv.setValue(schema.json)
        fmd.setKey_value_metadata(Arrays.asList(kv))
        fmd.setCreated_by("hsgpu (GpuPageFileWriter; page bodies by hs_parquet_encode_pages)")
        val before = out.getPos
        Util.writeFileMetaData(fmd, out)
        val flen = (out.getPos - before).toInt
        out.write(Array[Byte](flen.toByte, (flen >> 8).toByte, (flen >> 16).toByte, (flen >> 24).toByte))
        out.write(Magic)
      } finally out.close()
      path.toString
    }
  }

  /** PageHeader in Thrift compact protocol; returns its length. */
  private def writeHeader(h: PageHeader, out: OutputStream): Int = {
    val buf = new java.io.ByteArrayOutputStream(64)
    Util.writePageHeader(h, buf)
    out.write(buf.toByteArray)
    buf.size
  }
}
