"""The per-gene standard output table (``standouttable.csv``).

The reference initialises ``standtable`` from the SummarizedExperiment's rowData
(cml_targetaml_seanalysis.Rmd:630-634), appends one column per (model, replicate) importance
(:679-690 SVM, :793-799 lasso, :1087-1091 RF, :1246-1263 XGB) and writes it at :1298.  Schema (24 columns):
hgnc_id, hgnc_symbol, ensembl_gene_id, logFC, AveExpr, t, p.unadj, p.adj.bh, b, lasso_coef_rep1..3,
rfnb_{2k,5k,10k}_MeanDecNodeImp, svm1..4_weights, xg1..5_imp.

Known reference defects are NOT reproduced: its lasso columns hold only the last coefficient of each replicate
(loop bug :787) and svm2/svm4 duplicate svm1/svm3 (:155 vs :166-169); here every column holds what its name says."""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np
import pandas as pd

from ..data import Cohort
from .gene_csv import _q, r_num

BASE_COLUMNS = ["hgnc_id", "hgnc_symbol", "ensembl_gene_id", "logFC", "AveExpr", "t", "p.unadj", "p.adj.bh", "b"]
REFERENCE_COLUMNS = BASE_COLUMNS + [f"lasso_coef_rep{i}" for i in (1, 2, 3)] + \
    [f"rfnb_{k}_MeanDecNodeImp" for k in ("2k", "5k", "10k")] + [f"svm{i}_weights" for i in range(1, 5)] + \
    [f"xg{i}_imp" for i in range(1, 6)]


class StandTable:
    def __init__(self, cohort: Cohort):
        rd = cohort.row_data
        self.df = pd.DataFrame({c: rd[c].values if c in rd else np.nan for c in BASE_COLUMNS}, index=list(cohort.gene_ids))

    def add(self, name: str, values, gene_ids: Optional[Sequence[str]] = None, fill: float = 0.0) -> "StandTable":
        """Append an importance column.  ``values`` is aligned with the table's genes, or with ``gene_ids``
        (genes not mentioned get ``fill`` - e.g. lasso zeros, genes excluded in later replicates)."""
        if isinstance(values, dict):
            gene_ids, values = list(values.keys()), list(values.values())
        v = np.asarray(values, dtype=np.float64)
        if gene_ids is None:
            if v.shape[0] != len(self.df):
                raise ValueError("values must have one entry per gene")
            self.df[name] = v
        else:
            col = pd.Series(fill, index=self.df.index, dtype=np.float64)
            col.loc[list(gene_ids)] = v
            self.df[name] = col
        return self

    def write_csv(self, path: str) -> None:
        """R ``write.csv`` layout: quoted header with a leading empty name, quoted row names and strings."""
        with open(path, "w", newline="") as fh:
            fh.write(",".join([_q("")] + [_q(c) for c in self.df.columns]) + "\n")
            str_cols = {c for c in self.df.columns if not pd.api.types.is_numeric_dtype(self.df[c])}
            for gid, row in zip(self.df.index, self.df.itertuples(index=False)):
                cells = [_q(gid)]
                for c, v in zip(self.df.columns, row):
                    if c in str_cols:
                        cells.append("NA" if v is None or v != v or v == "" else _q(v))
                    else:
                        cells.append(r_num(float(v)) if v == v else "NA")
                fh.write(",".join(cells) + "\n")

    @staticmethod
    def read_csv(path: str) -> pd.DataFrame:
        return pd.read_csv(path, index_col=0)
