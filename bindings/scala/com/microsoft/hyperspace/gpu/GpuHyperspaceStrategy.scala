/*
 * The planner hook. enableHyperspace() (src/main/scala/com/microsoft/hyperspace/package.scala:47-54) adds it
 * next to the reference's own strategy:
 *
 *   extraStrategies ++= BucketUnionStrategy :: GpuHyperspaceStrategy :: Nil
 *
 * following the precedent of index/execution/BucketUnionStrategy.scala:28-34. It fires only on plans the
 * reference's rules have already rewritten — a Join whose two sides scan IndexHadoopFsRelations with a
 * BucketSpec (JoinIndexRule.apply, index/rules/JoinIndexRule.scala:57-90, via RuleUtils.scala:264-292), or a
 * Filter over an index scan (FilterIndexRule.apply, index/rules/FilterIndexRule.scala:43-85) — and plans the
 * GPU operators for them; everything else falls through to Spark's own operators (the library has no CPU
 * fallback, HS_EUNSUPPORTED surfaces as HyperspaceException and the rules' non-fatal fallback applies).
 * Compile-gated (no JVM in the development image).
 */
package com.microsoft.hyperspace.gpu

import org.apache.spark.sql.catalyst.expressions.{And, AttributeReference, EqualTo, Expression}
import org.apache.spark.sql.catalyst.plans._
import org.apache.spark.sql.catalyst.plans.logical.{Filter, Join, LogicalPlan, Project}
import org.apache.spark.sql.execution.{SparkPlan, SparkStrategy}
import org.apache.spark.sql.execution.datasources.LogicalRelation

import com.microsoft.hyperspace.index.plans.logical.IndexHadoopFsRelation

object GpuHyperspaceStrategy extends SparkStrategy {

  /** A linear plan ending in an index scan that carries a BucketSpec. */
  private object BucketedIndexScan {
    def unapply(plan: LogicalPlan): Option[(LogicalRelation, IndexHadoopFsRelation)] = plan match {
      case l @ LogicalRelation(rel: IndexHadoopFsRelation, _, _, _) if rel.bucketSpec.isDefined => Some((l, rel))
      case Project(_, child) => unapply(child)
      case _ => None
    }
  }
  private object IndexScan {
    def unapply(plan: LogicalPlan): Option[(LogicalRelation, IndexHadoopFsRelation)] = plan match {
      case l @ LogicalRelation(rel: IndexHadoopFsRelation, _, _, _) => Some((l, rel))
      case _ => None
    }
  }

  private def hsJoinType(t: JoinType): Option[Int] = t match {
    case Inner => Some(HsGpu.JOIN_INNER)
    case LeftOuter => Some(HsGpu.JOIN_LEFT_OUTER)
    case RightOuter => Some(HsGpu.JOIN_RIGHT_OUTER)
    case FullOuter => Some(HsGpu.JOIN_FULL_OUTER)
    case LeftSemi => Some(HsGpu.JOIN_LEFT_SEMI)
    case LeftAnti => Some(HsGpu.JOIN_LEFT_ANTI)
    case _ => None
  }

  /** cond is a conjunction of attr = attr (the shape JoinIndexRule.isApplicable insists on, :118-124). */
  private def equiKeys(cond: Expression): Option[Seq[(AttributeReference, AttributeReference)]] = cond match {
    case EqualTo(a: AttributeReference, b: AttributeReference) => Some(Seq((a, b)))
    case And(l, r) => for (x <- equiKeys(l); y <- equiKeys(r)) yield x ++ y
    case _ => None
  }

  override def apply(plan: LogicalPlan): Seq[SparkPlan] = plan match {
    case j @ Join(BucketedIndexScan(l, lrel), BucketedIndexScan(r, rrel), joinType, Some(cond)) =>
      (for (jt <- hsJoinType(joinType); keys <- equiKeys(cond)) yield
        GpuBucketedSortMergeJoinExec(l, lrel, r, rrel, jt, keys, j.output) :: Nil).getOrElse(Nil)
    case Filter(cond, IndexScan(l, rel)) =>
      GpuPredicate.compile(cond, l.output).map(p => GpuFilteredIndexScanExec(l, rel, p, l.output) :: Nil).getOrElse(Nil)
    case Project(projectList, Filter(cond, IndexScan(l, rel))) if projectList.forall(_.isInstanceOf[AttributeReference]) =>
      GpuPredicate.compile(cond, l.output)
        .map(p => GpuFilteredIndexScanExec(l, rel, p, projectList.map(_.toAttribute)) :: Nil).getOrElse(Nil)
    case _ => Nil
  }
}
