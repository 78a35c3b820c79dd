"""Headless dumps of the arrays behind the reference's figures (no graphics stack on this box: pictures are left
out, their data are written as CSV / NPZ).

Reference plots: volcano (JSmith_code/Differential_Expression_and_Lasso.Rmd:33-85), DEG heatmaps with clustering
(cml_targetaml_seanalysis.Rmd:509-623), lasso performance curves (:806-881), Spearman correlation heatmaps and
densities (:883-1017), RF / XGB importance bars and SVM weight bars (:697-713, 1098-1143, 1267-1293), per-sample log2-CPM
boxplots by cohort (JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:120-140)."""
from __future__ import annotations

import os
from typing import Dict, Optional, Sequence

import numpy as np
import pandas as pd
import torch

from ..stats import correlation as C


def volcano_data(logfc: np.ndarray, p_value: np.ndarray, adj_p: np.ndarray, gene_ids: Sequence[str], lfc: float = 1.0,
                 fdr: float = 0.05) -> pd.DataFrame:
    """x = log2 fold change, y = -log10 p, class in {up, down, ns} (the three colours of ``volcano_plot``)."""
    cls = np.where((adj_p < fdr) & (logfc > lfc), "up", np.where((adj_p < fdr) & (logfc < -lfc), "down", "ns"))
    return pd.DataFrame({"gene": list(gene_ids), "logFC": logfc, "neg_log10_p": -np.log10(np.maximum(p_value, 1e-300)),
                         "adj.P.Val": adj_p, "class": cls})


def top_quantile(values: np.ndarray, gene_ids: Sequence[str], q: float = 0.99, absolute: bool = True) -> pd.DataFrame:
    """Genes above the ``q`` quantile of (absolute) importance / weight - the bars of SEA:697-713 and :1098-1116."""
    v = np.abs(values) if absolute else values
    thr = np.quantile(v, q)
    idx = np.flatnonzero(v > thr)
    idx = idx[np.argsort(-v[idx])]
    return pd.DataFrame({"gene": [gene_ids[i] for i in idx], "value": values[idx]})


def cohort_group(sample_id: str) -> str:
    """Cohort from the TARGET project code inside a barcode (``TARGET.30.`` NBL, ``.20.`` AML, ``.50.`` WT, ``.52.`` RT, ...;
    Normalization_and_Batch_Effect_Investigation.Rmd:120-126)."""
    parts = sample_id.replace("-", ".").split(".")
    code = parts[1] if len(parts) > 1 else ""
    return {"30": "NBL", "20": "AML", "21": "AML", "50": "WT", "51": "CCSK", "52": "RT", "40": "OS"}.get(code, "NA")


def boxplot_stats(logcpm: torch.Tensor, sample_ids: Sequence[str]) -> pd.DataFrame:
    """Per-sample box-and-whisker numbers of a genes x samples log2-CPM matrix (median, hinges, 1.5 IQR whiskers, outlier
    count), ordered by median like the reference's ``reorder(USI, Log2.CPM, FUN = median)`` plot, with the cohort label."""
    x = logcpm.double()
    q = torch.quantile(x, torch.tensor([0.25, 0.5, 0.75], dtype=torch.float64, device=x.device), dim=0)
    iqr = q[2] - q[0]
    lo_f, hi_f = q[0] - 1.5 * iqr, q[2] + 1.5 * iqr
    inside = (x >= lo_f) & (x <= hi_f)
    big, small = torch.finfo(torch.float64).max, torch.finfo(torch.float64).min
    whisk_lo = torch.where(inside, x, torch.full_like(x, big)).amin(0)
    whisk_hi = torch.where(inside, x, torch.full_like(x, small)).amax(0)
    df = pd.DataFrame({"sample": list(sample_ids), "group": [cohort_group(s) for s in sample_ids], "median": q[1].cpu().numpy(),
                       "q1": q[0].cpu().numpy(), "q3": q[2].cpu().numpy(), "whisker_low": whisk_lo.cpu().numpy(),
                       "whisker_high": whisk_hi.cpu().numpy(), "n_outliers": (~inside).sum(0).cpu().numpy()})
    return df.sort_values("median", kind="stable").reset_index(drop=True)


def lasso_rep_curves(results) -> pd.DataFrame:
    """Per exclusion replicate: TPR, TNR, FDR, FOR, test error, number of genes (SEA:811-880)."""
    rows = []
    for i, r in enumerate(results, 1):
        rep = r.report.as_dict() if r.report is not None else {}
        rows.append({"rep": i, "n_genes": len(r.nonzero_coef), "test_error": r.test_error, "auc": r.test_auc,
                     **{k: rep.get(k) for k in ("tpr", "tnr", "fdr", "for")}})
    return pd.DataFrame(rows)


def correlation_heatmap(assay: torch.Tensor, gene_idx_rows: Sequence[int], gene_idx_cols: Sequence[int],
                        samples: Optional[Sequence[int]] = None):
    """Spearman block between two gene sets (rep1 vs rep2 / rep3 genes, SEA:889-969) with clustered ordering."""
    a = assay[torch.as_tensor(list(gene_idx_rows))].float()
    b = assay[torch.as_tensor(list(gene_idx_cols))].float()
    cor = C.correlation(a, b, samples=samples, exact=True)
    ro, co = C.heatmap_order(cor) if min(cor.shape) > 2 else (np.arange(cor.shape[0]), np.arange(cor.shape[1]))
    grid, dens = C.kde(cor.flatten())
    return {"cor": cor.cpu().numpy(), "row_order": ro, "col_order": co, "density_x": grid.cpu().numpy(), "density_y": dens.cpu().numpy()}


def write_all(out_dir: str, tables: Dict[str, pd.DataFrame], arrays: Optional[Dict[str, Dict[str, np.ndarray]]] = None) -> None:
    os.makedirs(out_dir, exist_ok=True)
    for name, df in tables.items():
        df.to_csv(os.path.join(out_dir, f"{name}.csv"), index=False)
    for name, arrs in (arrays or {}).items():
        np.savez_compressed(os.path.join(out_dir, f"{name}.npz"), **arrs)
