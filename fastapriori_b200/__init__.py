"""fastapriori_b200 — B200-native mining hot path of relife957/FastApriori.

Host mirror of the reference's operator interface (FastApriori, AssociationRules, Utils, Main)
over the C-ABI of include/fastapriori_b200.h; all compute runs in hand-written sm_100a kernels.
"""
from . import utils as Utils
from .association_rules import AssociationRules, NoRulesError
from .context import MiningContext
from .data import TokenizedLines
from .fast_apriori import FastApriori, FreqItemsets

__all__ = ["FastApriori", "AssociationRules", "Utils", "MiningContext", "TokenizedLines", "FreqItemsets", "NoRulesError"]
