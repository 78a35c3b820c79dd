"""Sample filters, label recoding, train/test splits and class re-balancing.

Reference: tissue-code filter and label recoding (JSmith_code/Differential_Expression_and_Lasso.Rmd:119-146;
seobjects/make_seobj_targetaml.R:56-83; scripts/model_comp.py:58-60; cml_targetaml_seanalysis.Rmd:435-451), the
seed-2019 two-thirds split (Differential_Expression_and_Lasso.Rmd:159-186), the fixed id lists
(JSmith_code/TARGET_AML_{Training,Testing}_Samples.csv), the 70/30 split of VikasP/AML.R:159-165 and the ROSE
over-sampling stub (Differential_Expression_and_Lasso.Rmd:536-543)."""
from __future__ import annotations

from typing import Iterable, Optional, Sequence, Tuple

import numpy as np
import pandas as pd

from ..data import Cohort

NON_DIAGNOSTIC = ("04A", "40A", "02A", "06A", "11A")          # excluded at DEL:122
PRIMARY = ("03A", "09A")                                        # kept at MSE:56-59


def diagnostic_id(barcode: str) -> str:
    return barcode.replace("-", ".").split(".")[3] if barcode.count(".") + barcode.count("-") >= 3 else ""


def filter_diagnostic(clinical: pd.DataFrame, column: str = "Diagnostic ID", drop: Iterable[str] = NON_DIAGNOSTIC,
                      keep: Optional[Iterable[str]] = None) -> pd.DataFrame:
    ids = clinical[column].astype(str)
    if keep is not None:
        return clinical[ids.isin(list(keep))]
    pat = "|".join(drop)
    return clinical[~ids.str.contains(pat, regex=True)]


def recode_risk(clinical: pd.DataFrame, column: str = "Risk group") -> pd.DataFrame:
    """``RiskGroup.Class``: High/Standard -> "Yes", Low -> "No", Unknown -> "Unknown" (DEL:128-131), plus the numeric
    ``deg.risk`` (0 = Low, 1 = Standard/High; SEA:435-451) and ``Low Risk`` of model_comp.py:60."""
    out = clinical.copy()
    rg = out[column].astype(str)
    out["RiskGroup.Class"] = np.where(rg.str.contains("High|Standard"), "Yes", np.where(rg.str.contains("Low"), "No", "Unknown"))
    out["deg.risk"] = np.where(out["RiskGroup.Class"] == "Yes", 1, np.where(out["RiskGroup.Class"] == "No", 0, -1))
    out["Low Risk"] = (rg == "Low").astype(int)
    return out


def age_class(clinical: pd.DataFrame, column: str = "Age at Diagnosis in Days") -> pd.Series:
    age = pd.to_numeric(clinical[column], errors="coerce")
    return pd.Series(np.where(age > age.median(), "Yes", "No"), index=clinical.index)


def mll_update(clinical: pd.DataFrame, column: str = "Primary Cytogenetic Code") -> pd.Series:
    c = clinical[column].astype(str)
    return pd.Series(np.where(c == "MLL", "MLL", np.where(c == "Unknown", "Unknown", "No")), index=clinical.index)


def drop_unknown(cohort: Cohort, column: str = "deg.risk") -> Cohort:
    v = cohort.col_data[column]
    bad = v.astype(str).isin(["Unknown", "-1", "nan", "None"])
    return cohort.select_samples(~bad.to_numpy())


def dedupe_samples(ids: Sequence[str], key=lambda s: s[:16]) -> np.ndarray:
    """Keep the first sample of every patient (the reference dedupes by patient code, VikasP/AML.R:31-64)."""
    seen, keep = set(), []
    for i, s in enumerate(ids):
        k = key(s)
        if k not in seen:
            seen.add(k)
            keep.append(i)
    return np.asarray(keep)


def split_random(ids: Sequence[str], train_frac: float = 2.0 / 3.0, seed: int = 2019) -> Tuple[list, list]:
    """``sample(ids, size = n * frac)``-style split (R's RNG stream is not reproducible here; the reference's own id
    lists are what ``split_fixed`` uses)."""
    rng = np.random.default_rng(seed)
    ids = list(ids)
    k = int(len(ids) * train_frac)
    tr = set(rng.choice(len(ids), k, replace=False).tolist())
    return [s for i, s in enumerate(ids) if i in tr], [s for i, s in enumerate(ids) if i not in tr]


def split_stratified(labels: Sequence, train_frac: float = 0.7, seed: int = 2128) -> np.ndarray:
    """Boolean train mask with the class proportions preserved (the 70/30 split of VikasP/AML.R:159-165)."""
    rng = np.random.default_rng(seed)
    lab = np.asarray(labels)
    mask = np.zeros(lab.size, bool)
    for v in np.unique(lab):
        idx = np.flatnonzero(lab == v)
        k = int(round(train_frac * idx.size))
        mask[rng.choice(idx, k, replace=False)] = True
    return mask


def split_fixed(cohort: Cohort, train_ids: Iterable[str], test_ids: Iterable[str], usi=lambda s: ".".join(s.replace("-", ".").split(".")[:3]),
                column: str = "exptset.seahack") -> Cohort:
    """Attach the reference's committed train/test id lists as a split column."""
    tr = {i.replace("-", ".") for i in train_ids}
    te = {i.replace("-", ".") for i in test_ids}
    flag = []
    for s in cohort.sample_ids:
        u = usi(s)
        flag.append("train" if u in tr else ("test" if u in te else "unassigned"))
    out = Cohort(cohort.assay, cohort.gene_ids, cohort.sample_ids, cohort.row_data, cohort.col_data.copy(), dict(cohort.metadata),
                 cohort.assay_name)
    if len(out.col_data) == 0:
        out.col_data = pd.DataFrame(index=range(len(flag)))
    out.col_data[column] = flag
    return out


def holdout_then_split(n: int, holdout_frac: float = 0.2, test_size: float = 0.33, seed: int = 8):
    """model_comp.py:102-107: 20 % holdout, then a 67/33 split of the rest.  Returns index arrays (train, test, holdout)."""
    rng = np.random.default_rng(seed)
    perm = rng.permutation(n)
    h = int(round(holdout_frac * n))
    hold, rest = perm[:h], perm[h:]
    t = int(np.ceil(test_size * rest.size))
    return rest[t:], rest[:t], hold


def resample_balanced(labels: Sequence, n_out: Optional[int] = None, p: float = 0.5, method: str = "both", seed: int = 1) -> np.ndarray:
    """ROSE ``ovun.sample(method = "both" | "over" | "under", p, N, seed)``: indices of a re-balanced sample in which
    the minority class makes up a fraction ``p``."""
    rng = np.random.default_rng(seed)
    lab = np.asarray(labels)
    vals, counts = np.unique(lab, return_counts=True)
    if vals.size != 2:
        raise ValueError("binary labels expected")
    minority, majority = vals[np.argmin(counts)], vals[np.argmax(counts)]
    i_min, i_maj = np.flatnonzero(lab == minority), np.flatnonzero(lab == majority)
    if method == "both":
        n_out = lab.size if n_out is None else n_out
        n_min = rng.binomial(n_out, p)
        return np.concatenate([rng.choice(i_min, n_min, replace=True), rng.choice(i_maj, n_out - n_min, replace=True)])
    if method == "over":
        n_min = int(round(i_maj.size * p / (1 - p)))
        return np.concatenate([i_maj, rng.choice(i_min, n_min, replace=True)])
    if method == "under":
        n_maj = int(round(i_min.size * (1 - p) / p))
        return np.concatenate([i_min, rng.choice(i_maj, min(n_maj, i_maj.size), replace=False)])
    raise ValueError(method)


def qc_summary(clinical: pd.DataFrame) -> dict:
    """Clinical summaries of JSmith_code/aml-target_qc-scripts_SeanMaden.R:17-87: counts by gender, age statistics and
    the above/below-median age classifier, blast percentages, event-free survival with its <= median classifier."""
    out = {"n": len(clinical)}
    if "Gender" in clinical:
        out["gender"] = clinical["Gender"].value_counts(dropna=False).to_dict()
    for col, key in (("Age at Diagnosis in Days", "age_days"), ("Bone marrow leukemic blast percentage (%)", "bm_blast_pct"),
                     ("Peripheral blasts (%)", "pb_blast_pct"), ("Event Free Survival Time in Days", "efs_days"),
                     ("Overall Survival Time in Days", "os_days")):
        if col in clinical:
            v = pd.to_numeric(clinical[col], errors="coerce").dropna()
            if len(v):
                out[key] = {"n": int(len(v)), "mean": float(v.mean()), "median": float(v.median()), "min": float(v.min()),
                            "max": float(v.max()), "above_median": int((v > v.median()).sum())}
    if "Risk group" in clinical:
        out["risk_group"] = clinical["Risk group"].value_counts(dropna=False).to_dict()
    return out


def r_summary(values) -> dict:
    """R's ``summary()`` of a numeric vector: min, quartiles (type-7 quantiles), median, mean, max and the NA count."""
    v = pd.to_numeric(pd.Series(values), errors="coerce")
    na = int(v.isna().sum())
    v = v.dropna().to_numpy(dtype=float)
    if v.size == 0:
        return {"n": 0, "na": na}
    q1, med, q3 = np.quantile(v, [0.25, 0.5, 0.75])
    return {"n": int(v.size), "min": float(v.min()), "q1": float(q1), "median": float(med), "mean": float(v.mean()),
            "q3": float(q3), "max": float(v.max()), "na": na}


def qc_report(clinical: pd.DataFrame) -> dict:
    """Everything ``aml-target_qc-scripts_SeanMaden.R:17-71`` prints: gender and CNS-disease tables, ``summary()`` of age
    (days and years), blast percentage and survival times, the young / old classifier (age below / not below the median),
    its cross-table with gender and the survival summaries inside each age class."""
    out = {"n": len(clinical)}
    for col, key in (("Gender", "gender"), ("CNS disease", "cns_disease")):
        if col in clinical:
            out[key] = clinical[col].value_counts().sort_index().to_dict()
    age = pd.to_numeric(clinical.get("Age at Diagnosis in Days"), errors="coerce") if "Age at Diagnosis in Days" in clinical else None
    if age is not None:
        out["age_days"] = r_summary(age)
        out["age_years"] = r_summary(age / 365)
        med = float(np.nanmedian(age))
        cls = np.where(age < med, "young", "old")
        out["median_age_days"] = med
        out["age_class"] = pd.Series(cls).value_counts().to_dict()
        if "Gender" in clinical:
            out["age_class_by_gender"] = pd.crosstab(pd.Series(cls, name="class.age"), clinical["Gender"].reset_index(drop=True))
        for col, key in (("Overall Survival Time in Days", "os_days"), ("Event Free Survival Time in Days", "efs_days")):
            if col in clinical:
                out[key] = r_summary(clinical[col])
                for c in ("young", "old"):
                    out[f"{key}_{c}"] = r_summary(clinical[col].to_numpy()[cls == c])
    if "Bone marrow leukemic blast percentage (%)" in clinical:
        out["bm_blast_pct"] = r_summary(clinical["Bone marrow leukemic blast percentage (%)"])
    return out


def barcode_table(sample_ids: Sequence[str]) -> dict:
    """Split TARGET barcodes (``TARGET-20-PABCDE-09A-01R``, ``-`` or ``.`` separated) into their five fields and count the
    tissue codes (09A, 03A, ...) and portion codes (01R, ...) - the barcode matrix of the QC script (lines 74-87)."""
    rows = [str(s).replace("-", ".").split(".") for s in sample_ids]
    rows = [r for r in rows if len(r) >= 5]
    bcm = pd.DataFrame([r[:5] for r in rows], columns=["program", "project", "case", "tissue", "portion"])
    return {"matrix": bcm, "tissue": bcm["tissue"].value_counts().sort_index().to_dict(),
            "portion": bcm["portion"].value_counts().sort_index().to_dict()}


def survival_density_groups(clinical: pd.DataFrame) -> pd.DataFrame:
    """Long table (group, survival.time) of the four curves of the "Survival Time by Age Classifier" plot (lines 96-113):
    overall and event-free survival for the young and the old class.  Densities come from ``stats.correlation.kde``."""
    age = pd.to_numeric(clinical["Age at Diagnosis in Days"], errors="coerce")
    cls = np.where(age < float(np.nanmedian(age)), "young", "old")
    parts = []
    for c in ("young", "old"):
        for col, tag in (("Overall Survival Time in Days", "overallsurv"), ("Event Free Survival Time in Days", "efsurv")):
            v = pd.to_numeric(clinical[col], errors="coerce").to_numpy()[cls == c]
            parts.append(pd.DataFrame({"group": f"{c}.{tag}", "survival.time": v[~np.isnan(v)]}))
    return pd.concat(parts, ignore_index=True)


def balance_test(labels_a: Sequence, labels_b: Sequence) -> dict:
    """Chi-square independence test of two categorical columns, e.g. gender vs binarised risk
    (cml_targetaml_seanalysis.Rmd:462,470: p = 0.8044 for gender, 0.6591 for age class)."""
    import torch
    from ..ops.metrics import chisq_test
    tab = pd.crosstab(pd.Series(list(labels_a)), pd.Series(list(labels_b)))
    stat, dof, p = chisq_test(torch.as_tensor(tab.to_numpy()))
    return {"table": tab, "statistic": stat, "dof": dof, "p_value": p}
