"""Set-algebra consensus of per-model gene selections.

Restates the notebook cells that intersect the XGBoost / random-forest / lasso importance sets
(scripts/model_walkthrough.ipynb raw lines 1796-2131 and 4878-4991: pairwise intersections 48 / 290 / 75,
three-way 11, union 4285 with a membership table)."""
from __future__ import annotations

from itertools import combinations
from typing import Dict, Iterable, List

import pandas as pd


def selection_sets(importances: Dict[str, "pd.Series"], positive_only: bool = True) -> Dict[str, set]:
    """Genes with importance > 0 per model.  ``positive_only`` keeps the reference's ``coef_ > 0`` rule for
    linear models (model_walkthrough.ipynb:1488-1489, 3914); pass False to use ``!= 0``."""
    out = {}
    for name, s in importances.items():
        mask = (s > 0) if positive_only else (s != 0)
        out[name] = set(s.index[mask])
    return out


def intersections(sets: Dict[str, set]) -> Dict[str, set]:
    """Every k-way intersection (k >= 2), keyed by ``"a&b"``."""
    names = list(sets)
    out = {}
    for k in range(2, len(names) + 1):
        for combo in combinations(names, k):
            out["&".join(combo)] = set.intersection(*(sets[c] for c in combo))
    return out


def membership_table(sets: Dict[str, set]) -> pd.DataFrame:
    """Union of all sets with one 0/1 column per model and a ``n_models`` count, most-shared genes first."""
    union = sorted(set().union(*sets.values()))
    df = pd.DataFrame({name: [int(g in s) for g in union] for name, s in sets.items()}, index=union)
    df["n_models"] = df.sum(axis=1)
    return df.sort_values(["n_models"], ascending=False, kind="stable")


def consensus_genes(sets: Dict[str, set], min_models: int) -> List[str]:
    tab = membership_table(sets)
    return tab.index[tab["n_models"] >= min_models].tolist()
