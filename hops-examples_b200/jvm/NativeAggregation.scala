package io.hops.examples.featurestore_tour.native

import java.nio.ByteBuffer
import org.apache.spark.sql.{DataFrame, Row, SparkSession}
import org.apache.spark.sql.types._

/** What replaces the `groupBy("team_id").sum()` / `.count()` / `.join(count, "team_id")` / `.map(...)`
  * block inside the reference's compute*FeatureGroup functions (ComputeFeatures.scala:148-160,206-218,
  * 245-265).  The function signatures of ComputeFeatures stay as they are; their bodies call this helper
  * and then run the unchanged feature-group builder chain and `fg.save(df)`.
  * NOT compiled in this repository's image (no JDK); see INTEGRATION.md.
  */
object NativeAggregation {
  import Native._

  /** One (sum column, avg name, sum name) triple per aggregated input column. */
  case class Spec(source: String, dtype: Int, avgName: String, sumName: String)

  /** team_id + float32 features in the given field order, sorted by team_id. */
  def aggregateByTeam(spark: SparkSession, ctx: Long, teamId: ByteBuffer, n: Long,
                      inputs: Map[String, ByteBuffer], specs: Seq[Spec], fieldOrder: Seq[String]): DataFrame = {
    val key = columnUpload(ctx, I32, n, teamId, null)
    val vals = specs.map(s => columnUpload(ctx, s.dtype, n, inputs(s.source), null))
    val table = groupbyAgg(ctx, key, (vals :+ 0L).toArray, (specs.map(_ => SUM) :+ COUNT_STAR).toArray, 0L)
    tableSortBy(ctx, table, 0)
    val count = tableColumn(table, specs.size + 1)
    val feats = specs.zipWithIndex.flatMap { case (s, i) =>
      val sum = tableColumn(table, i + 1)
      Seq(s.avgName -> finalizeF32(ctx, sum, count), s.sumName -> finalizeF32(ctx, sum, 0L))
    }.toMap
    val g = tableNrows(table).toInt
    val columns: Seq[(String, Array[Any])] = fieldOrder.map { name =>
      if (name == "team_id") {
        val b = direct(4L * g); columnDownload(ctx, tableColumn(table, 0), b, null)
        name -> Array.tabulate[Any](g)(i => b.getInt(4 * i))
      } else {
        val b = direct(4L * g); columnDownload(ctx, feats(name), b, null)
        name -> Array.tabulate[Any](g)(i => b.getFloat(4 * i))
      }
    }
    feats.values.foreach(columnFree(ctx, _)); vals.foreach(columnFree(ctx, _)); columnFree(ctx, key); tableFree(ctx, table)
    val schema = StructType(fieldOrder.map(f => StructField(f, if (f == "team_id") IntegerType else FloatType, nullable = false)))
    val rows = (0 until g).map(r => Row.fromSeq(columns.map(_._2(r))))
    spark.createDataFrame(spark.sparkContext.parallelize(rows), schema)
  }
}
