"""One-hot encoding with train-vocabulary semantics - the reference's ``scripts/dummies.py`` (lines 1-28),
with its missing ``import pandas as pd`` (a NameError at first call) fixed."""
from __future__ import annotations

from typing import Iterable, Sequence

import pandas as pd


def make_dummies(test_col: pd.Series, train_unique_vals: Iterable, col_name: str) -> pd.DataFrame:
    """One 0/1 column per value seen in training; a level unseen in training yields an all-zero row.
    (dummies.py:1-13)"""
    return pd.DataFrame({f"{col_name}_{val}": (test_col == val).astype(int) for val in train_unique_vals}, index=test_col.index)


def make_dummies_dataframe(data: pd.DataFrame, categories: Sequence[str]) -> pd.DataFrame:
    """Append dummy columns for every listed categorical column.  (dummies.py:15-28)"""
    out = data
    for category in categories:
        out = pd.concat([out, make_dummies(data[category], data[category].unique(), category)], axis=1)
    return out
