package com.hazzacheng.AR

/**
  * com.hazzacheng.AR.Main on libfastapriori_b200.so (JNA), no Spark on the hot path.
  *
  * Same contract as the reference's Main (reference Main.scala:15-41, README.md:19-41):
  *   Main <input prefix> <output prefix> <temp (ignored)> [minSupport]
  * reads <input>D.dat / <input>U.dat, writes <output>freqItemset/part-00000 and
  * <output>recommends/part-00000 (+ _SUCCESS), minSupport defaults to 0.092.
  *
  * NOT COMPILED IN THIS REPOSITORY'S IMAGE (no JVM, scalac or JNA jar there); it is the
  * binding a maintainer adds on a box that has them:  scalac -cp jna.jar Main.scala
  * The tested twin with the identical call sequence is fastapriori_b200/main.py.
  */
import java.io.{File, PrintWriter}
import java.nio.charset.StandardCharsets
import java.nio.file.{Files, Paths}

import com.sun.jna.{Library, Memory, Native, Pointer}
import com.sun.jna.ptr.{LongByReference, PointerByReference}

import scala.collection.mutable

trait FaLib extends Library {
  def fa_create(device: Int, stream: Pointer, ctxOut: PointerByReference): Int
  def fa_destroy(ctx: Pointer): Unit
  def fa_last_error(ctx: Pointer): String
  def fa_load_transactions(ctx: Pointer, rowOffsets: Pointer, itemIds: Pointer, nRows: Long, nItems: Int): Int
  def fa_load_transactions_u16(ctx: Pointer, off: Pointer, ids16: Pointer, n: Long, nItems: Int): Int
  def fa_count_items(ctx: Pointer, countsOut: Pointer): Int
  def fa_global_rows(ctx: Pointer, nOut: LongByReference): Int
  def fa_build_bitmaps(ctx: Pointer, idToRank: Pointer, nFrequent: Int, rankCounts: Pointer): Int
  def fa_count_pairs(ctx: Pointer, minCount: Int, nOut: LongByReference): Int
  def fa_gen_candidates(ctx: Pointer, k: Int, groupsOut: LongByReference, candsOut: LongByReference): Int
  def fa_count_candidates(ctx: Pointer, k: Int, minCount: Int, nOut: LongByReference): Int
  def fa_level_counters(ctx: Pointer, k: Int, g: LongByReference, c: LongByReference): Int
  def fa_level_size(ctx: Pointer, k: Int, nOut: LongByReference): Int
  def fa_get_level(ctx: Pointer, k: Int, setsOut: Pointer, countsOut: Pointer): Int
  def fa_gen_rules(ctx: Pointer, nOut: LongByReference): Int
  def fa_prune_rules(ctx: Pointer, nOut: LongByReference): Int
  def fa_sort_rules(ctx: Pointer, rankToItemInt: Pointer): Int
  def fa_recommend(ctx: Pointer, userOffsets: Pointer, userRanks: Pointer, nUsers: Long, out: Pointer): Int
  def fa_recommend_ids(ctx: Pointer, userOffsets: Pointer, userIds: Pointer, nUsers: Long, idToRank: Pointer, nIds: Int, out: Pointer): Int
  def fa_put_level(ctx: Pointer, k: Int, sets: Pointer, counts: Pointer, n: Long): Int               // resume at the rule stage
  def fa_comm_unique_id(idOut: Pointer, errOut: Pointer, errCap: Long): Int                          // row-sharded runs: NCCL inside the library
  def fa_comm_init(ctx: Pointer, id: Pointer, rank: Int, worldSize: Int): Int
  def fa_rules_digest(ctx: Pointer, out4: Pointer): Int                                               // audit: 4 x uint64 over the kept, sorted rules
}

object Main {
  private val lib = Native.load("fastapriori_b200", classOf[FaLib]).asInstanceOf[FaLib]

  private def check(ctx: Pointer, rc: Int): Unit =
    if (rc != 0) throw new RuntimeException("libfastapriori_b200 error " + rc + ": " + lib.fa_last_error(ctx))

  private def ints(a: Array[Int]): Memory = { val m = new Memory(math.max(4L, 4L * a.length)); m.write(0, a, 0, a.length); m }
  private def longs(a: Array[Long]): Memory = { val m = new Memory(math.max(8L, 8L * a.length)); m.write(0, a, 0, a.length); m }

  /** Utils.scala:21 — `_.trim().split("\\s+")` per line; Files.readAllLines splits on \n, \r\n, \r like TextInputFormat. */
  private def readLines(path: String): Array[Array[String]] = {
    import scala.collection.JavaConverters._
    Files.readAllLines(Paths.get(path), StandardCharsets.UTF_8).asScala.map(_.trim().split("\\s+")).toArray
  }

  private def writePart(dir: String, lines: Iterator[String]): Unit = {
    val d = new File(dir)
    if (d.exists()) throw new java.io.IOException("Output directory " + dir + " already exists") // saveAsTextFile semantics
    d.mkdirs()
    val w = new PrintWriter(new File(d, "part-00000"), "UTF-8")
    try lines.foreach { l => w.print(l); w.print('\n') } finally w.close()
    new File(d, "_SUCCESS").createNewFile()
  }

  private def toIntOpt(s: String): Option[Int] = try Some(s.toInt) catch { case _: NumberFormatException => None }

  def main(args: Array[String]): Unit = {
    val input = args(0)
    val output = args(1)
    val minSupport = if (args.length > 3) args(3).toDouble else sys.env.getOrElse("FA_MIN_SUPPORT", "0.092").toDouble
    val data = readLines(input + "D.dat")
    val users = readLines(input + "U.dat")

    // dense item ids
    val vocab = mutable.ArrayBuffer.empty[String]
    val index = mutable.HashMap.empty[String, Int]
    val offsets = new Array[Long](data.length + 1)
    val ids = mutable.ArrayBuffer.empty[Int]
    for (i <- data.indices) {
      data(i).foreach { t => ids += index.getOrElseUpdate(t, { vocab += t; vocab.length - 1 }) }
      offsets(i + 1) = ids.length
    }

    val ctxRef = new PointerByReference()
    if (lib.fa_create(0, null, ctxRef) != 0) throw new RuntimeException("fa_create failed: no usable CUDA device")
    val ctx = ctxRef.getValue
    try {
      val time1 = System.currentTimeMillis()
      check(ctx, lib.fa_load_transactions(ctx, longs(offsets), ints(ids.toArray), data.length.toLong, vocab.length))
      val count = data.length.toLong
      val minCount = math.ceil(minSupport * count).toInt                       // FastApriori.scala:39
      val countsMem = new Memory(math.max(4L, 4L * vocab.length))
      check(ctx, lib.fa_count_items(ctx, countsMem))                           // K1
      val counts = countsMem.getIntArray(0, vocab.length)
      // FastApriori.scala:57-62: frequent items by count desc; canonical tie rule (Int value, then bytes)
      val freqIds = vocab.indices.filter(counts(_) >= minCount).sortWith { (a, b) =>
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
      freqItems.zipWithIndex.foreach { case (s, r) => itemToRank.put(s, r) }
      val idToRank = Array.fill(vocab.length)(-1)
      freqIds.zipWithIndex.foreach { case (id, r) => idToRank(id) = r }
      check(ctx, lib.fa_build_bitmaps(ctx, ints(idToRank), freqIds.length, ints(freqIds.map(counts(_)))))   // K2

      val n = new LongByReference()
      val g = new LongByReference()
      val c = new LongByReference()
      check(ctx, lib.fa_count_pairs(ctx, minCount, n))                         // K3
      println("==== 2 freq items " + n.getValue)
      var k = 3
      var nk = n.getValue
      while (nk >= k) {                                                        // FastApriori.scala:110-111
        check(ctx, lib.fa_gen_candidates(ctx, k, g, c))                        // K5
        println("==== " + k + " candidate items " + g.getValue)
        check(ctx, lib.fa_count_candidates(ctx, k, minCount, n))               // K4
        nk = n.getValue
        println("==== " + k + " freq items " + nk)
        k += 1
      }
      val maxK = k - 1

      // Utils.scala:29-41 saveFreqItemset
      val lines = mutable.ArrayBuffer.empty[String]
      freqItems.foreach(lines += _)
      for (kk <- 2 to maxK) {
        check(ctx, lib.fa_level_size(ctx, kk, n))
        val cnt = n.getValue.toInt
        if (cnt > 0) {
          val setsMem = new Memory(4L * cnt * kk)
          val cntMem = new Memory(4L * cnt)
          check(ctx, lib.fa_get_level(ctx, kk, setsMem, cntMem))
          val sets = setsMem.getIntArray(0, cnt * kk)
          for (s <- 0 until cnt) lines += (kk - 1 to 0 by -1).map(j => freqItems(sets(s * kk + j))).mkString(" ")
        }
      }
      writePart(output + "freqItemset", lines.sorted.iterator)
      println("==== Total time for get freqItemsets " + (System.currentTimeMillis() - time1))

      // AssociationRules.scala:23-31
      val time2 = System.currentTimeMillis()
      check(ctx, lib.fa_gen_rules(ctx, n))                                     // K6
      if (n.getValue == 0) throw new UnsupportedOperationException("empty.min") // AssociationRules.scala:147
      check(ctx, lib.fa_prune_rules(ctx, n))                                   // K7
      println("==== Size association rules " + n.getValue)
      check(ctx, lib.fa_sort_rules(ctx, ints(freqItems.map(s => toIntOpt(s).getOrElse(0)))))   // K8
      val uoff = new Array[Long](users.length + 1)
      val uranks = mutable.ArrayBuffer.empty[Int]
      for (i <- users.indices) {
        users(i).foreach(t => itemToRank.get(t).foreach(uranks += _))
        uoff(i + 1) = uranks.length
      }
      val outMem = new Memory(math.max(4L, 4L * users.length))
      check(ctx, lib.fa_recommend(ctx, longs(uoff), ints(uranks.toArray), users.length.toLong, outMem))   // K9
      val rec = outMem.getIntArray(0, users.length)
      writePart(output + "recommends", rec.iterator.map(r => if (r >= 0) freqItems(r) else "0"))   // Utils.scala:43-49
      println("==== Total time for get recommends " + (System.currentTimeMillis() - time2))
    } finally lib.fa_destroy(ctx)
  }
}
