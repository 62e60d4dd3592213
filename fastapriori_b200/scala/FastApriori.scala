package com.hazzacheng.AR

/**
  * Spark-facing wrappers with the reference's original signatures (SURVEY.md 8 f-4), for users who do have Spark:
  *
  *   new FastApriori(minSupport, numPartitions).run(sc, data)        reference FastApriori.scala:18,31-34
  *   new AssociationRules(freqItemset, freqItems, itemToRank).run(sc, userRDD)   reference AssociationRules.scala:17-23
  *
  * The RDDs are only the way the lines arrive: they are collected to the driver, tokens are mapped to dense ids
  * there, and every pass over transactions, bitmaps, candidates, rules and baskets is a call into
  * libfastapriori_b200.so through the JNA trait `FaLib` of Main.scala — no executor touches the hot path.  The mining
  * context (`fa_ctx`) is shared by the two operators through the companion object, so the rule stage uses the
  * itemsets that are already resident on the GPU, exactly like `Main` does.
  *
  * NOT COMPILED IN THIS REPOSITORY'S IMAGE (no JVM, scalac, Spark or JNA there).  On a box that has them:
  *   scalac -cp jna.jar:spark-core.jar Main.scala FastApriori.scala
  * The call sequence is the one of the tested hosts (fastapriori_b200/fast_apriori.py, association_rules.py).
  */
import com.sun.jna.{Memory, Native, Pointer}
import com.sun.jna.ptr.{LongByReference, PointerByReference}
import org.apache.spark.SparkContext
import org.apache.spark.rdd.RDD

import scala.collection.mutable

object FaSession {
  val lib: FaLib = Native.load("fastapriori_b200", classOf[FaLib]).asInstanceOf[FaLib]
  private var ctx: Pointer = _

  /** One fa_ctx per driver JVM (device from FA_DEVICE, default 0); released by close() or at JVM exit. */
  def context: Pointer = synchronized {
    if (ctx == null) {
      val ref = new PointerByReference()
      if (lib.fa_create(sys.env.getOrElse("FA_DEVICE", "0").toInt, null, ref) != 0)
        throw new RuntimeException("fa_create failed: no usable CUDA device (there is no CPU fallback)")
      ctx = ref.getValue
      sys.addShutdownHook(close())
    }
    ctx
  }
  def close(): Unit = synchronized { if (ctx != null) { lib.fa_destroy(ctx); ctx = null } }
  def check(rc: Int): Unit =
    if (rc != 0) throw new RuntimeException("libfastapriori_b200 error " + rc + ": " + lib.fa_last_error(ctx))
  def ints(a: Array[Int]): Memory = { val m = new Memory(math.max(4L, 4L * a.length)); m.write(0, a, 0, a.length); m }
  def longs(a: Array[Long]): Memory = { val m = new Memory(math.max(8L, 8L * a.length)); m.write(0, a, 0, a.length); m }
  def toIntOpt(s: String): Option[Int] = try Some(s.toInt) catch { case _: NumberFormatException => None }
}

class FastApriori(private var minSupport: Double, private var numPartitions: Int) extends Serializable {
  import FaSession._

  def this() = this(0.3, -1)
  def setMinSupport(minSupport: Double): this.type = { this.minSupport = minSupport; this }
  def setNumPartitions(numPartitions: Int): this.type = { this.numPartitions = numPartitions; this }

  /** reference FastApriori.scala:31-44: (frequent itemsets over ranks with supports, itemToRank, freqItems). */
  def run(sc: SparkContext, data: RDD[Array[String]]): (Array[(Set[Int], Int)], mutable.HashMap[String, Int], Array[String]) = {
    val lines = data.collect()                                                // the only Spark action on this path
    val vocab = mutable.ArrayBuffer.empty[String]
    val index = mutable.HashMap.empty[String, Int]
    val offsets = new Array[Long](lines.length + 1)
    val ids = mutable.ArrayBuffer.empty[Int]
    for (i <- lines.indices) {
      lines(i).foreach { t => ids += index.getOrElseUpdate(t, { vocab += t; vocab.length - 1 }) }
      offsets(i + 1) = ids.length
    }
    val ctx = context
    check(lib.fa_load_transactions(ctx, longs(offsets), ints(ids.toArray), lines.length.toLong, vocab.length))
    val minCount = math.ceil(minSupport * lines.length.toLong).toInt          // :38-39
    val countsMem = new Memory(math.max(4L, 4L * vocab.length))
    check(lib.fa_count_items(ctx, countsMem))                                 // :55-56 (K1)
    val counts = countsMem.getIntArray(0, vocab.length)
    val freqIds = vocab.indices.filter(counts(_) >= minCount).sortWith { (a, b) =>   // :57-60, canonical tie rule
      if (counts(a) != counts(b)) counts(a) > counts(b)
      else (toIntOpt(vocab(a)), toIntOpt(vocab(b))) match {
        case (Some(x), Some(y)) => x < y
        case (Some(_), None) => true
        case (None, Some(_)) => false
        case _ => vocab(a) < vocab(b)
      }
    }.toArray
    val freqItems = freqIds.map(vocab(_))
    val itemToRank = mutable.HashMap.empty[String, Int]
    freqItems.zipWithIndex.foreach { case (s, r) => itemToRank.put(s, r) }    // :62
    val idToRank = Array.fill(vocab.length)(-1)
    freqIds.zipWithIndex.foreach { case (id, r) => idToRank(id) = r }
    check(lib.fa_build_bitmaps(ctx, ints(idToRank), freqIds.length, ints(freqIds.map(counts(_)))))   // :66-79, :195-210 (K2)
    val n = new LongByReference()
    val g = new LongByReference()
    val c = new LongByReference()
    check(lib.fa_count_pairs(ctx, minCount, n))                               // :104 (K3)
    var k = 3
    var nk = n.getValue
    while (nk >= k) {                                                         // :110-111
      check(lib.fa_gen_candidates(ctx, k, g, c))                              // :113 (K5)
      check(lib.fa_count_candidates(ctx, k, minCount, n))                     // :115 (K4)
      nk = n.getValue
      k += 1
    }
    val out = mutable.ArrayBuffer.empty[(Set[Int], Int)]
    for (kk <- 2 until k) {                                                   // :41 k >= 2 first ...
      check(lib.fa_level_size(ctx, kk, n))
      val cnt = n.getValue.toInt
      if (cnt > 0) {
        val setsMem = new Memory(4L * cnt * kk)
        val cntMem = new Memory(4L * cnt)
        check(lib.fa_get_level(ctx, kk, setsMem, cntMem))
        val sets = setsMem.getIntArray(0, cnt * kk)
        val sup = cntMem.getIntArray(0, cnt)
        for (s <- 0 until cnt) out += ((sets.slice(s * kk, (s + 1) * kk).toSet, sup(s)))
      }
    }
    for (r <- freqItems.indices) out += ((Set(r), counts(freqIds(r))))        // ... then the 1-itemsets (:83)
    (out.toArray, itemToRank, freqItems)
  }
}

class AssociationRules(private val freqItemset: Array[(Set[Int], Int)], private val freqItems: Array[String],
                       private val itemToRank: mutable.HashMap[String, Int]) extends Serializable {
  import FaSession._

  /** reference AssociationRules.scala:23-31: (user index, recommended item or "0").  Uses the itemsets the last
    * FastApriori.run left on the GPU; a `freqItemset` from elsewhere would be installed with fa_put_level first. */
  def run(sc: SparkContext, userRDD: RDD[Array[String]]): Array[(Int, String)] = {
    val users = userRDD.collect()
    val ctx = context
    val n = new LongByReference()
    check(lib.fa_gen_rules(ctx, n))                                           // :126-145 (K6)
    if (n.getValue == 0) throw new UnsupportedOperationException("empty.min") // :147
    check(lib.fa_prune_rules(ctx, n))                                         // :147-182 (K7)
    check(lib.fa_sort_rules(ctx, ints(freqItems.map(s => toIntOpt(s).getOrElse(0)))))   // :74, :116-120 (K8)
    val uoff = new Array[Long](users.length + 1)
    val uranks = mutable.ArrayBuffer.empty[Int]
    for (i <- users.indices) {                                                // :42-47 filter frequent -> rank
      users(i).foreach(t => itemToRank.get(t).foreach(uranks += _))
      uoff(i + 1) = uranks.length
    }
    val outMem = new Memory(math.max(4L, 4L * users.length))
    check(lib.fa_recommend(ctx, longs(uoff), ints(uranks.toArray), users.length.toLong, outMem))      // :51-58, :80-106 (K9)
    outMem.getIntArray(0, users.length).zipWithIndex.map { case (r, i) => (i, if (r >= 0) freqItems(r) else "0") }
  }
}
