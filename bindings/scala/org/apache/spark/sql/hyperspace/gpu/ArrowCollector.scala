/*
 * Rows of a DataFrame as Arrow off-heap vectors on the driver. Lives under org.apache.spark.sql, like the
 * reference's own shims (src/main/scala/org/apache/spark/sql/hyperspace/utils/package.scala:22-32), because
 * Spark 2.4.2's ArrowWriter / ArrowUtils are private[sql]. Compile-gated (no JVM in the development image).
 */
package org.apache.spark.sql.hyperspace.gpu

import scala.collection.JavaConverters._

import org.apache.arrow.vector.{FieldVector, VectorSchemaRoot}
import org.apache.spark.sql.DataFrame
import org.apache.spark.sql.execution.arrow.{ArrowUtils, ArrowWriter}

object ArrowCollector {
  /** All rows in arrival order, one vector per column. The caller closes the root. */
  def collect(df: DataFrame): VectorSchemaRoot = {
    val timeZone = df.sparkSession.sessionState.conf.sessionLocalTimeZone
    val arrowSchema = ArrowUtils.toArrowSchema(df.schema, timeZone)
    val root = VectorSchemaRoot.create(arrowSchema, ArrowUtils.rootAllocator)
    val writer = ArrowWriter.create(root)
    df.queryExecution.toRdd.toLocalIterator.foreach(writer.write)
    writer.finish()
    root
  }
  def vectors(root: VectorSchemaRoot): Seq[FieldVector] = root.getFieldVectors.asScala
}
