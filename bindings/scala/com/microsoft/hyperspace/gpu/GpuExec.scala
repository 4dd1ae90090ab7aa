/*
 * The physical operators GpuHyperspaceStrategy plans. Each reads the index files of the relation the
 * reference's rules put into the plan (IndexHadoopFsRelation: index/plans/logical/IndexHadoopFsRelation.scala:29-50;
 * the bucket id of a file comes from its name, as in Spark's bucketed FileSourceScanExec and
 * actions/OptimizeAction.scala:129), hands bucket-contiguous Arrow columns to libhsgpu and returns rows.
 * One task per GPU; with one GPU that is one partition. Compile-gated (no JVM in the development image); the
 * same calls run in hyperspace-1_b200/rules.py + dataframe.py under tests/test_gpu_e2e.py.
 */
package com.microsoft.hyperspace.gpu

import java.nio.{ByteBuffer, ByteOrder}

import org.apache.spark.rdd.RDD
import org.apache.spark.sql.catalyst.InternalRow
import org.apache.spark.sql.catalyst.expressions._
import org.apache.spark.sql.execution.{LeafExecNode, SparkPlan}
import org.apache.spark.sql.execution.datasources.LogicalRelation
import org.apache.spark.sql.types._

import com.microsoft.hyperspace.index.plans.logical.IndexHadoopFsRelation

/** A filter condition compiled to the postfix hs_pred_term program of include/hsgpu.h (three-valued logic). */
case class GpuPredicate(terms: Seq[(Int, Int, Any)]) // (hs_pred_op, column, literal)

object GpuPredicate {
  private val EQ = 1; private val NE = 2; private val LT = 3; private val LE = 4; private val GT = 5
  private val GE = 6; private val IS_NULL = 7; private val IS_NOT_NULL = 8; private val IN = 9
  private val AND = 16; private val OR = 17; private val NOT = 18

  /** None when the condition uses anything the scan kernel does not evaluate: Spark's FilterExec runs then. */
  def compile(cond: Expression, output: Seq[Attribute]): Option[GpuPredicate] = {
    def col(a: Attribute): Option[Int] = Some(output.indexWhere(_.exprId == a.exprId)).filter(_ >= 0)
    def leaf(op: Int, a: Attribute, v: Any): Option[Seq[(Int, Int, Any)]] = col(a).map(c => Seq((op, c, v)))
    def go(e: Expression): Option[Seq[(Int, Int, Any)]] = e match {
      case EqualTo(a: Attribute, Literal(v, _)) if v != null => leaf(EQ, a, v)
      case LessThan(a: Attribute, Literal(v, _)) if v != null => leaf(LT, a, v)
      case LessThanOrEqual(a: Attribute, Literal(v, _)) if v != null => leaf(LE, a, v)
      case GreaterThan(a: Attribute, Literal(v, _)) if v != null => leaf(GT, a, v)
      case GreaterThanOrEqual(a: Attribute, Literal(v, _)) if v != null => leaf(GE, a, v)
      case IsNull(a: Attribute) => leaf(IS_NULL, a, null)
      case IsNotNull(a: Attribute) => leaf(IS_NOT_NULL, a, null)
      case In(a: Attribute, list) if list.forall(l => l.isInstanceOf[Literal] && l.dataType.isInstanceOf[IntegralType]) =>
        leaf(IN, a, list.map(_.asInstanceOf[Literal].value.asInstanceOf[Number].longValue))
      case Not(EqualTo(a: Attribute, Literal(v, _))) if v != null => leaf(NE, a, v)
      case Not(c) => go(c).map(_ :+ ((NOT, 0, null)))
      case And(l, r) => for (x <- go(l); y <- go(r)) yield (x ++ y) :+ ((AND, 0, null))
      case Or(l, r) => for (x <- go(l); y <- go(r)) yield (x ++ y) :+ ((OR, 0, null))
      case _ => None
    }
    go(cond).filter(_.size <= 32).map(GpuPredicate(_))
  }

  /** hs_pred_term[] (40 bytes each) in a direct buffer; literal bytes and IN lists live in `keep`. */
  def encode(p: GpuPredicate, keep: scala.collection.mutable.Buffer[ByteBuffer]): ByteBuffer = {
    val b = ByteBuffer.allocateDirect(p.terms.size * 40).order(ByteOrder.LITTLE_ENDIAN)
    p.terms.zipWithIndex.foreach { case ((op, c, lit), i) =>
      val at = i * 40
      b.putInt(at, op); b.putInt(at + 4, c)
      lit match {
        case n: java.lang.Number if !n.isInstanceOf[java.lang.Float] && !n.isInstanceOf[java.lang.Double] =>
          b.putLong(at + 8, n.longValue); b.putDouble(at + 16, n.doubleValue)
        case n: java.lang.Number => b.putDouble(at + 16, n.doubleValue)
        case s: org.apache.spark.unsafe.types.UTF8String =>
          val d = ByteBuffer.allocateDirect(math.max(1, s.numBytes)); d.put(s.getBytes); keep += d
          b.putLong(at + 24, HsGpuBuffers.address(d)); b.putInt(at + 32, s.numBytes)
        case vs: Seq[_] =>
          val d = ByteBuffer.allocateDirect(8 * math.max(1, vs.size)).order(ByteOrder.LITTLE_ENDIAN)
          vs.foreach(v => d.putLong(v.asInstanceOf[Long])); keep += d
          b.putLong(at + 24, HsGpuBuffers.address(d)); b.putInt(at + 32, vs.size)
        case _ =>
      }
    }
    b
  }
}

object HsGpuBuffers {
  /** Address of a direct buffer (sun.nio.ch.DirectBuffer), for pointers stored inside structs. */
  def address(b: ByteBuffer): Long = b.asInstanceOf[sun.nio.ch.DirectBuffer].address()
}

/**
 * Filter over an index scan (the plan of FilterIndexRule.apply, index/rules/FilterIndexRule.scala:43-85):
 * hs_filter_scan over the whole index as one view (useBucketSpec = false there, so no bucket structure is
 * promised to Spark; the library still prunes `key = literal` to one bucket).
 */
case class GpuFilteredIndexScanExec(relation: LogicalRelation, index: IndexHadoopFsRelation, predicate: GpuPredicate,
    output: Seq[Attribute]) extends LeafExecNode {
  override protected def doExecute(): RDD[InternalRow] = {
    val schema = relation.schema
    val outCols = output.map(a => relation.output.indexWhere(_.exprId == a.exprId)).toArray
    val files = index.location.inputFiles
    val numBuckets = index.bucketSpec.map(_.numBuckets).getOrElse(1)
    val nIndexed = index.bucketSpec.map(_.bucketColumnNames.size).getOrElse(1)
    sparkContext.parallelize(Seq(0), 1).mapPartitions { _ =>
      val ctx = HsGpu.context()
      val run = IndexRunReader.read(files, schema, numBuckets) // bucket-contiguous Arrow columns + offsets
      val keep = scala.collection.mutable.Buffer[ByteBuffer]()
      val table = HsGpu.filterScan(ctx, run.columns, schema.size, run.bucketOffsets, nIndexed, numBuckets,
        GpuPredicate.encode(predicate, keep), predicate.terms.size, outCols)
      run.close()
      TableRows.iterator(ctx, table, StructType(outCols.map(schema(_))))
    }
  }
}

/**
 * Bucketed sort-merge join without exchange (the plan of JoinIndexRule.apply, index/rules/JoinIndexRule.scala:57-90,
 * over two relations made by RuleUtils.transformPlanToUseIndexOnlyScan with useBucketSpec = true,
 * index/rules/RuleUtils.scala:264-292): bucket i of the left index joins bucket i of the right one inside
 * hs_join_bucketed; several files per bucket (after an incremental refresh,
 * actions/RefreshIncrementalAction.scala:129-144) are merged by IndexRunReader with hs_merge_runs first —
 * the place of the per-side SortExec Spark plans in that case.
 */
case class GpuBucketedSortMergeJoinExec(left: LogicalRelation, leftIndex: IndexHadoopFsRelation, right: LogicalRelation,
    rightIndex: IndexHadoopFsRelation, joinType: Int, keys: Seq[(AttributeReference, AttributeReference)],
    output: Seq[Attribute]) extends LeafExecNode {
  override protected def doExecute(): RDD[InternalRow] = {
    val (ls, rs) = (left.schema, right.schema)
    val lOut = output.map(a => left.output.indexWhere(_.exprId == a.exprId)).filter(_ >= 0).toArray
    val rOut = output.map(a => right.output.indexWhere(_.exprId == a.exprId)).filter(_ >= 0).toArray
    val (lf, rf) = (leftIndex.location.inputFiles, rightIndex.location.inputFiles)
    val (lb, rb) = (leftIndex.bucketSpec.get.numBuckets, rightIndex.bucketSpec.get.numBuckets)
    val nk = keys.size
    sparkContext.parallelize(Seq(0), 1).mapPartitions { _ =>
      val ctx = HsGpu.context()
      val (l, r) = (IndexRunReader.read(lf, ls, lb), IndexRunReader.read(rf, rs, rb))
      val table = HsGpu.joinBucketed(ctx, l.columns, ls.size, l.bucketOffsets, r.columns, rs.size, r.bucketOffsets,
        nk, lb, rb, joinType, lOut, rOut)
      l.close(); r.close()
      TableRows.iterator(ctx, table, StructType(lOut.map(ls(_)) ++ rOut.map(rs(_))))
    }
  }
}
