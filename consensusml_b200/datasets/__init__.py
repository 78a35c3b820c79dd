"""Bundled fixtures: the TARGET-AML modelling matrix and the reference's stored model outputs.

These are *data* extracted from the reference's R checkpoints by
``tools/extract_reference_fixtures.py`` (sources: composite_code/rnotebook/data/*.rda, written at
cml_targetaml_seanalysis.Rmd:472,778,1069,1239).  They are the golden oracles for the test-suite and
the input of the real-data benchmark; they exist because the GPU box has no reference mount.
"""
from __future__ import annotations

import os
from functools import lru_cache

import numpy as np
import pandas as pd
import torch

from ..data import Cohort

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "target_aml")


def path(name: str) -> str:
    return os.path.join(_DIR, name)


@lru_cache(maxsize=1)
def _degfilt_npz():
    return np.load(path("degfilt_1984x137.npz"), allow_pickle=False)


def load_target_aml() -> Cohort:
    """The 1984-DEG x 137-sample TMM log2-CPM cohort (``degfilt.se``) with ``deg.risk`` labels
    (0 = low, 1 = standard/high) and the fixed 93/44 train/test split."""
    d = _degfilt_npz()
    row = pd.DataFrame({
        "hgnc_id": d["hgnc_id"], "hgnc_symbol": d["hgnc_symbol"], "ensembl_gene_id": d["ensembl_gene_id"],
        "logFC": d["logFC"], "AveExpr": d["AveExpr"], "t": d["t"], "p.unadj": d["p_unadj"],
        "p.adj.bh": d["p_adj_bh"], "b": d["b"],
    })
    col = pd.read_csv(path("degfilt_coldata.csv"), low_memory=False)
    col["deg.risk"] = d["deg_risk"].astype(int)
    col["exptset.seahack"] = d["split"]
    return Cohort(torch.from_numpy(d["assay"].copy()), list(d["gene_ids"]), list(d["sample_ids"]), row, col,
                  {"source": "sesetfilt_degseahack_targetaml.rda"})


@lru_cache(maxsize=1)
def glmnet_oracle():
    """Stored glmnet objects of the 15 exclusion replicates (coefficient paths, CV curves, ...)."""
    return np.load(path("glmnet_oracle.npz"), allow_pickle=False)


@lru_cache(maxsize=1)
def rf_oracle():
    return np.load(path("rf_oracle.npz"), allow_pickle=False)


@lru_cache(maxsize=1)
def xgb_oracle():
    return np.load(path("xgb_oracle.npz"), allow_pickle=False)


@lru_cache(maxsize=1)
def gene_annotation() -> pd.DataFrame:
    """Ensembl v95 / hg38 BioMart annotation (``dfens``; make_seobj_targetaml.R:103-117)."""
    return pd.read_csv(path("gene_annotation_ens95.csv.gz"))
