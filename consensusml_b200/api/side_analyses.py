"""The reference's three side analyses as runnable workflows that write the files the originals wrote.

  mlseq_prepare / mlseq_zoo   VikasP/AML.R        the MLSeq / DESeq2 classifier zoo on raw counts           -> DEG_voomnsc.csv, DEG_plda.csv, DEG_plda2.csv
  wilms_glmboost              Ryan/Feature_selection_GLMTrain_WT_TARGET.R   glmboost on event-free survival   -> feature_list.csv
  km_whatever                 BuieRProj/KM_WHATEVER.Rmd   k-means sweep + logistic GLM, linear-SVM weight tails, QDA -> output/High*.csv, low*.csv

Each follows its script step by step (line references inline), including the steps that are questionable there - the test
subset of the glmboost script is its training subset (RYN:60-61), the BuieRProj ``train`` flag marks the 49 held-out patients
(BUI:52-53) - behind explicit switches, so that both the reference's numbers and honest ones can be produced.  The models are
this library's kernels (``models.discriminant``, ``models.deseq``, ``models.glmboost``, ``models.kmeans``, ``models.svm``)."""
from __future__ import annotations

import math
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import pandas as pd
import torch

from ..models import discriminant as DC
from ..models import glmboost as GB
from ..models import kmeans as KM
from ..models import svm as SV
from ..models.glm import logistic_glm


def _strip_version(ids: Sequence[str]) -> List[str]:
    return [str(g).split(".")[0] for g in ids]


def _write_r_vector(path: str, values: Sequence[str]) -> None:
    """``write.csv(character vector)``: header ``"","x"``, quoted 1-based row names."""
    with open(path, "w") as fh:
        fh.write('"","x"\n')
        for i, v in enumerate(values, 1):
            fh.write(f'"{i}","{v}"\n')


def _confusion_stats(pred: np.ndarray, actual: np.ndarray, positive) -> Dict[str, float]:
    """caret ``confusionMatrix(table(Predicted, Actual), positive = )``: accuracy, sensitivity, specificity, the table."""
    tp = int(((pred == positive) & (actual == positive)).sum())
    tn = int(((pred != positive) & (actual != positive)).sum())
    fp = int(((pred == positive) & (actual != positive)).sum())
    fn = int(((pred != positive) & (actual == positive)).sum())
    d = lambda a, b: a / b if b else float("nan")
    return {"accuracy": d(tp + tn, tp + tn + fp + fn), "sensitivity": d(tp, tp + fn), "specificity": d(tn, tn + fp),
            "table": [[tp, fp], [fn, tn]]}


# ------------------------------------------------------------------------------------------------ VikasP/AML.R
def mlseq_prepare(counts: pd.DataFrame, deg_ids: Sequence[str], clinical: pd.DataFrame, *, n_genes: int = 100, test_frac: float = 0.3,
                  seed: int = 2128, usi_col: str = "TARGET.USI", risk_col: str = "Risk.group") -> Dict:
    """VIK:14-171.  ``counts``: genes x samples raw counts (index = gene ids, columns = dotted barcodes).  Keeps the DEG rows
    (:18-29), the first column of every patient code = third dot-separated token (:31-45), the first clinical row of every
    patient (:48-63), the samples whose patient has a known risk group (:66-116), labels H = High / Standard, L = Low (:119-141),
    drops HTSeq's ``__`` summary rows (:143-151), keeps the ``n_genes`` rows of highest variance (:156-157), draws
    ceiling(0.3 n) test samples (:158-159; numpy's generator stands in for R's ``sample``) and adds the pseudo-count (:161-162)."""
    deg = set(map(str, deg_ids))
    sub = counts.loc[[g for g in counts.index if str(g) in deg]]
    code = lambda col: str(col).split(".")[2] if len(str(col).split(".")) > 2 else str(col)
    seen, keep_cols = set(), []
    for c in sub.columns:
        if code(c) not in seen:
            seen.add(code(c))
            keep_cols.append(c)
    sub = sub[keep_cols]
    clin = clinical.copy()
    import re
    clin["_code"] = [re.split(r"-|\s", str(u))[2] if len(re.split(r"-|\s", str(u))) > 2 else str(u) for u in clin[usi_col]]
    clin = clin[~clin["_code"].duplicated()]
    risk_of = dict(zip(clin["_code"], clin[risk_col].astype(str)))
    cols = [c for c in sub.columns if code(c) in risk_of and risk_of[code(c)] != "Unknown"]
    sub = sub[cols]
    labels = np.array(["L" if risk_of[code(c)] == "Low" else "H" for c in cols])
    sub = sub.loc[[g for g in sub.index if "__" not in str(g)]]
    var = sub.var(axis=1, ddof=1).sort_values(ascending=False, kind="stable")
    data = sub.loc[var.index[:n_genes]]
    n = data.shape[1]
    n_test = int(math.ceil(n * test_frac))
    rng = np.random.default_rng(seed)
    ind = np.sort(rng.choice(n, n_test, replace=False))
    tr = np.setdiff1d(np.arange(n), ind)
    return {"genes": list(data.index), "train": data.iloc[:, tr] + 1, "test": data.iloc[:, ind] + 1, "y_train": labels[tr], "y_test": labels[ind],
            "n_samples": n, "test_index": ind}


MLSEQ_METHODS = {"svmRadial": "deseq-vst", "pam": "deseq-vst", "PLDA": None, "PLDA2": None, "NBLDA": None, "voomDLDA": None, "voomNSC": None}


def mlseq_zoo(prep: Dict, out_dir: Optional[str] = None, *, methods: Optional[Dict[str, Optional[str]]] = None, folds: int = 5, repeats: int = 10,
              tune_length: int = 10, seed: int = 2128, device=None) -> Dict:
    """VIK:235-291: every classifier of the zoo tuned by ``folds`` x ``repeats`` repeated CV on the training counts, predictions
    and caret-style confusion statistics on the test counts (positive class H), ``selectedGenes`` of the sparse ones, and the
    three gene-list files the script writes (version-stripped ids)."""
    methods = dict(MLSEQ_METHODS if methods is None else methods)
    dev = torch.device(device) if device is not None else torch.device("cpu")
    ctr = torch.as_tensor(np.array(prep["train"].to_numpy(dtype=np.float64)), device=dev)
    cte = torch.as_tensor(np.array(prep["test"].to_numpy(dtype=np.float64)), device=dev)
    ytr = torch.as_tensor((prep["y_train"] == "H").astype(np.int64), device=dev)      # ref = "H": the positive class
    yte = (prep["y_test"] == "H").astype(np.int64)
    grids = {"PLDA": np.linspace(0, 15, tune_length), "PLDA2": np.linspace(0, 15, tune_length), "pam": np.linspace(0, 4, tune_length),
             "voomNSC": np.linspace(0, 4, tune_length), "svmRadial": [2.0 ** k for k in range(-2, tune_length - 2)]}
    out: Dict = {"methods": {}}
    for method, pre in methods.items():
        model, table = DC.classify(ctr, ytr, method, tune=grids.get(method), folds=folds, repeats=repeats, seed=seed, pre=pre or "deseq-vst")
        if method in ("PLDA", "PLDA2", "NBLDA") or method in DC._TRANSFORMED:
            pred = model.predict(cte)
        elif method == "pam":                                       # test counts through the TRAINING set's transform
            pred = model.predict(DC.deseq_vst(cte, fit=model.train_fit_) if model.pre_ == "deseq-vst" else DC.deseq_rlog(cte, frozen=model.train_fit_))
        else:                                                       # voom-based: log-cpm of the test counts
            lib = cte.sum(0)
            pred = model.predict(torch.log2((cte + 0.5) / (lib + 1)[None, :] * 1e6))
        pred = pred.cpu().numpy()
        rec = {"best_param": getattr(model, "best_param_", None), "cv_accuracy": max(table.values()), "test": _confusion_stats(pred, yte, 1)}
        try:
            rec["selected_genes"] = DC.selected_genes(model, prep["genes"])
        except Exception:                                          # noqa: BLE001 - a dense learner has no gene selection
            rec["selected_genes"] = list(prep["genes"])
        out["methods"][method] = rec
    if out_dir:
        os.makedirs(out_dir, exist_ok=True)
        for method, fname in (("voomNSC", "DEG_voomnsc.csv"), ("PLDA", "DEG_plda.csv"), ("PLDA2", "DEG_plda2.csv")):
            if method in out["methods"]:
                _write_r_vector(os.path.join(out_dir, fname), _strip_version(out["methods"][method]["selected_genes"]))
    return out


# ------------------------------------------------------------------------------------------------ Ryan/Feature_selection_GLMTrain_WT_TARGET.R
def wilms_glmboost(df: pd.DataFrame, *, target: str = "Event.Free.Survival.Time.in.Days", gene_cols: Optional[Sequence[str]] = None,
                   first_gene_col: int = 41, split_days: float = 313.0, cutoff: float = 331.0, threshold: float = 0.05, mstop: int = 100,
                   nu: float = 0.1, honest_test: bool = False, out_dir: Optional[str] = None, device=None) -> Dict:
    """RYN:12-108.  Dichotomise on ``target <= split_days`` (:12-18), keep the genes whose between-cohort mean difference exceeds
    ``threshold`` (:44-53), glmboost of the survival time on them using every second row (:56-62), predict the "test" rows - the
    SAME rows in the script (:60-61); ``honest_test=True`` predicts the other half instead - and count the dichotomised outcomes
    at ``cutoff`` days (:78-108).  Writes ``feature_list.csv`` (gene, coeff of the selected base learners) and the scatter data."""
    genes = list(gene_cols) if gene_cols is not None else list(df.columns[first_gene_col:])
    dev = torch.device(device) if device is not None else torch.device("cpu")
    y = torch.as_tensor(np.array(df[target].to_numpy(dtype=np.float64)), device=dev)
    x = torch.as_tensor(np.array(df[genes].to_numpy(dtype=np.float64)), device=dev)
    early = y <= split_days
    keep = GB.mean_difference_filter(x, early, threshold)
    xs = x[:, keep]
    n = x.shape[0]
    train = torch.arange(0, n, 2, device=dev)                        # seq(1, n, by = 2)
    test = torch.arange(1, n, 2, device=dev) if honest_test else train
    fit = GB.glmboost(xs[train], y[train], mstop=mstop, nu=nu)
    pred = fit.predict(xs[test])
    truth = y[test]
    pos_t, pos_p = (truth > cutoff).cpu().numpy(), (pred > cutoff).cpu().numpy()
    counts = {"cp": int(pos_t.sum()), "cn": int((~pos_t).sum()), "pp": int(pos_p.sum()), "pn": int((~pos_p).sum()),
              "tp": int((pos_t & pos_p).sum()), "fn": int((pos_t & ~pos_p).sum()), "fp": int((~pos_t & pos_p).sum()),
              "tn": int((~pos_t & ~pos_p).sum())}
    sel = fit.selected().cpu().tolist()
    features = pd.DataFrame({"gene": [genes[int(keep[j])] for j in sel], "coeff": [float(fit.coef[j]) for j in sel]})
    out = {"n_genes_after_filter": int(keep.numel()), "n_train": int(train.numel()), "n_test": int(test.numel()), "counts": counts,
           "metrics": GB.dichotomy_metrics(pred, truth, cutoff), "features": features, "intercept": fit.intercept,
           "train_equals_test": not honest_test}
    if out_dir:
        os.makedirs(out_dir, exist_ok=True)
        features.to_csv(os.path.join(out_dir, "feature_list.csv"), index=False)
        pd.DataFrame({"truth": truth.cpu().numpy(), "response": pred.cpu().numpy()}).to_csv(os.path.join(out_dir, "regression_scatter.csv"), index=False)
    return out


# ------------------------------------------------------------------------------------------------ BuieRProj/KM_WHATEVER.Rmd
def km_whatever(binary_features: Optional[pd.DataFrame] = None, risk: Optional[Sequence[int]] = None, *, expression: Optional[pd.DataFrame] = None,
                expr_risk: Optional[Sequence[int]] = None, train_flag: Optional[Sequence[int]] = None, ks: Sequence[int] = tuple(range(20, 1, -1)),
                nstart: int = 25, seed: int = 1000, out_dir: Optional[str] = None, device=None, qda_shrinkage: float = 0.1) -> Dict:
    """BUI:83-386 in its three aims.
    (1) ``binary_features`` (patients x cytogenetic / mutation indicators) with ``risk``: k-means for every k of ``ks`` with
        ``nstart`` restarts (:151-157), then ``glm(risk ~ factor(cluster), binomial)`` per k (:169-184): deviance and AIC.
    (2) ``expression`` (samples x genes) with ``expr_risk`` and ``train_flag`` (1 = the rows to TRAIN on, as the notebook uses it,
        BUI:304-312): linear SVM, weight-quantile tails at 25 / 10 / 5 / 1 / 0.1 % (:315-349), written as High4 / 10 / 20 / 100 /
        1000.csv and low*.csv (:353-362).
    (3) QDA on the same split (the notebook's aim 3 stops at the label recode, :375-386; fitted here on the top-weight genes)."""
    dev = torch.device(device) if device is not None else torch.device("cpu")
    out: Dict = {}
    if binary_features is not None:
        xb = torch.as_tensor(binary_features.to_numpy(dtype=np.float32), device=dev)
        yb = torch.as_tensor(np.asarray(risk, dtype=np.float64), device=dev)
        sweep = KM.kmeans_sweep(xb, ks, nstart=nstart, seed=seed)
        rows = []
        for k in ks:
            res = sweep[k]
            lab = res.labels.long()
            used = torch.unique(lab)
            design = torch.stack([torch.ones_like(yb)] + [(lab == c).double() for c in used[1:]], 1)
            fit = logistic_glm(design, yb)
            rows.append({"k": int(k), "tot_withinss": float(res.tot_withinss), "clusters_used": int(used.numel()),
                         "deviance": float(fit.deviance), "aic": float(fit.deviance + 2 * design.shape[1])})
        out["kmeans"] = pd.DataFrame(rows)
    if expression is not None:
        xe = torch.as_tensor(expression.to_numpy(dtype=np.float32), device=dev)
        ye = torch.as_tensor(np.asarray(expr_risk, dtype=np.int64), device=dev)
        flag = torch.as_tensor(np.asarray(train_flag, dtype=np.int64), device=dev)
        tr, te = flag == 1, flag == 0
        res = SV.run_svm(xe[tr], ye[tr], xe[te], ye[te], kerneltype="linear")
        genes = list(expression.columns)
        tails = SV.weight_quantile_tails(res.weightsvect, genes)
        out["svm"] = {"auc": float(res.auc), "error": float(res.test_error), "n_train": int(tr.sum()), "n_test": int(te.sum()),
                      "tail_sizes": {f: (len(tails["high"][f]), len(tails["low"][f])) for f in tails["high"]}}
        out["tails"] = tails
        names = {0.25: "4", 0.10: "10", 0.05: "20", 0.01: "100", 0.001: "1000"}
        if out_dir:
            os.makedirs(os.path.join(out_dir, "output"), exist_ok=True)
            for f, tag in names.items():
                _write_r_vector(os.path.join(out_dir, "output", f"High{tag}.csv"), tails["high"][f])
                _write_r_vector(os.path.join(out_dir, "output", f"low{tag}.csv"), tails["low"][f])
        top = torch.argsort(res.weightsvect.abs(), descending=True)[: max(2, min(10, int(tr.sum()) // 6))]
        q = DC.QDA(shrink=qda_shrinkage).fit(xe[tr][:, top].double(), ye[tr])
        out["qda"] = {"genes": [genes[i] for i in top.cpu().tolist()], "test_accuracy": float((q.predict(xe[te][:, top].double()) == ye[te]).float().mean())}
    return out


# columns of the TARGET-AML clinical sheet that are NOT cytogenetic / mutation indicators (BUI:86-124, R's dotted names undone)
_NON_PHENOTYPE = ["Gender", "Race", "Ethnicity", "Age at Diagnosis in Days", "First Event", "FLT3/ITD allelic ratio", "Event Free Survival Time in Days",
                  "Vital Status", "Overall Survival Time in Days", "Year of Diagnosis", "Year of Last Follow Up", "Protocol", "WBC at Diagnosis",
                  "Bone marrow leukemic blast percentage (%)", "Peripheral blasts (%)", "CNS disease", "Chloroma", "FAB Category", "Cytogenetic Code Other",
                  "Cytogenetic Complexity", "Primary Cytogenetic Code", "ISCN", "MRD at end of course 1", "MRD % at end of course 1", "MRD at end of course 2",
                  "MRD % at end of course 2", "CR status at end of course 1", "CR status at end of course 2", "Risk group", "SCT in 1st CR",
                  "Bone Marrow Site of Relapse/Induction Failure", "CNS Site of Relapse/Induction Failure", "Chloroma Site of Relapse/Induction Failure",
                  "Cytogenetic Site of Relapse/Induction Failure", "Other Site of Relapse/Induction Failure", "Comment",
                  "Refractory Timepoint sent for Induction Failure Project"]


def phenotype_matrix(clinical: pd.DataFrame, id_col: str = "TARGET USI") -> pd.DataFrame:
    """BUI:83-130: the patients x {translocations, deletions, monosomies / trisomies, gene mutations} indicator matrix the k-means of
    aim 1 runs on.  Non-phenotype columns are dropped, ``Unknown`` / ``N/A`` / ``Not done`` / ``No`` become 0 and ``Yes`` 1, anything
    that is not numeric afterwards - and every missing value, which the notebook reads as "not tested, so not expected" - is 0."""
    keep = [c for c in clinical.columns if c not in _NON_PHENOTYPE and c != id_col]
    m = clinical[keep].copy()
    rec = {"unknown": 0, "n/a": 0, "not done": 0, "no": 0, "yes": 1}
    for c in keep:
        col = m[c].map(lambda v: rec.get(v.strip().lower(), v) if isinstance(v, str) else v)
        m[c] = pd.to_numeric(col, errors="coerce").fillna(0.0)
    m.index = clinical[id_col].astype(str).to_numpy()
    return m.astype(np.float32)
