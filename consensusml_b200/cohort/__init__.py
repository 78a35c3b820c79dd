"""Cohort-level operations: filters, label recoding, splits, re-balancing."""
from .ops import *  # noqa: F401,F403
