"""Writes the inputs of the hand-derived known-answer vector.

The reference ships no fixtures (pom.xml:55 names a test directory that does not exist) and
cannot run here, so this vector was derived by hand from the Scala semantics
(FastApriori.scala:31-248, AssociationRules.scala:23-188, Utils.scala:29-49) and the
expected outputs next to it were typed from that derivation, not produced by any code in
this repository.  Running this script only regenerates D.dat / U.dat.
"""
import os

here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "appendix_b")
os.makedirs(here, exist_ok=True)
d = ["10 20 30"] * 6 + ["10 20"] * 6 + ["10 30"] * 6 + ["20 40"] + ["30"] * 2 + ["10"] * 3
u = ["20 30", "10", "10 20", "10 20 30", "40 99", "30", "20 30", ""]
open(os.path.join(here, "D.dat"), "w").write("\n".join(d) + "\n")
open(os.path.join(here, "U.dat"), "w").write("\n".join(u) + "\n")
