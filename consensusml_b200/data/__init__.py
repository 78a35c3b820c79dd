"""Data containers."""
from .cohort import Cohort  # noqa: F401
