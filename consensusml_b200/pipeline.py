"""The ConsensusML pipeline end to end: cohort -> (normalise -> DE prefilter) -> model ensemble -> per-gene standard
table -> consensus ranking -> reference-format outputs.

This is the ``cml_targetaml_seanalysis.Rmd`` run (call stack in SURVEY.md 3.3) plus the ``Differential_Expression_and_
Lasso.Rmd`` front half (3.2), with stages checkpointed like the reference's ``save()`` / ``load()`` points."""
from __future__ import annotations

import os
from typing import Dict, Optional

import numpy as np
import pandas as pd
import torch

from . import consensus as CS
from . import datasets
from .data import Cohort
from .models import lasso as L
from .models import mlp as MLP
from .models import svm as S
from .models import trees as T
from .ops import metrics as M
from .report import dump as R
from .runtime.config import Config
from .runtime.log import RunLog, check_finite
from .runtime.store import ArtifactStore
from .stats import de as DE
from .stats import normalize as N


def resolve_device(cfg: Config) -> torch.device:
    if cfg.device != "auto":
        return torch.device(cfg.device)
    return torch.device("cuda" if torch.cuda.is_available() else "cpu")


def cohort_from_counts(counts: torch.Tensor, gene_ids, sample_ids, group: np.ndarray, train_mask: np.ndarray, cfg: Config,
                       log: Optional[RunLog] = None) -> Cohort:
    """Raw counts -> TMM log2-CPM restricted to the limma-voom DEGs of the TRAINING samples (DEL:217-261)."""
    log = log or RunLog()
    dev = counts.device
    with log.stage("normalize", bytes_moved=float(counts.numel() * 4 * 3)):
        norm = N.tmm_normalize(counts, cfg.normalize.min_cpm, cfg.normalize.min_samples, prior_count=cfg.normalize.prior_count)
    with log.stage("voom_de"):
        tr = torch.as_tensor(np.flatnonzero(train_mask), device=dev)
        de = DE.voom_de(counts[:, tr].contiguous(), group[train_mask], min_frac=cfg.de.min_frac, p_value=cfg.de.p_value,
                        lfc=cfg.de.lfc, span=cfg.de.span)
    deg_full = de.gene_index[de.deg_order]                          # indices into the full gene list
    kept_full = np.flatnonzero(norm.keep.cpu().numpy())
    pos = {g: i for i, g in enumerate(kept_full)}
    rows = [pos[g] for g in deg_full if g in pos]
    assay = norm.logcpm[torch.as_tensor(rows, device=dev)]
    ids = [gene_ids[kept_full[r]] for r in rows]
    tab = de.table(gene_ids)
    rd = pd.DataFrame({"hgnc_id": "", "hgnc_symbol": "", "ensembl_gene_id": [i.split(".")[0] for i in ids],
                       "logFC": tab.loc[ids, "logFC"].values, "AveExpr": tab.loc[ids, "AveExpr"].values, "t": tab.loc[ids, "t"].values,
                       "p.unadj": tab.loc[ids, "P.Value"].values, "p.adj.bh": tab.loc[ids, "adj.P.Val"].values, "b": tab.loc[ids, "B"].values})
    cd = pd.DataFrame({"deg.risk": group.astype(int), "exptset.seahack": np.where(train_mask, "train", "test")})
    return Cohort(assay.double(), ids, list(sample_ids), rd, cd, {"n_deg": len(ids), "df_prior": de.df_prior})


def run(cfg: Config, cohort: Optional[Cohort] = None, log: Optional[RunLog] = None, return_details: bool = False) -> Dict:
    """Run the ensemble + consensus on ``cohort`` (default: the bundled TARGET-AML DEG cohort).  Returns a summary dict
    and writes ``standouttable.csv``, the lasso gene CSV, ``consensus_rank.csv`` and the report dumps."""
    dev = resolve_device(cfg)
    out = cfg.io.out_dir
    log = log or RunLog(os.path.join(out, "run.jsonl"))
    store = ArtifactStore(out, resume=cfg.resume)
    cohort = cohort or datasets.load_target_aml()
    log.event("run_start", device=str(dev), genes=cohort.n_genes, samples=cohort.n_samples, config_hash=cfg.hash())
    tr, te = cohort.split()
    xtr, ytr = tr.design(torch.float32).to(dev), tr.labels().to(dev)
    xte, yte = te.design(torch.float32).to(dev), te.labels().to(dev)
    genes = cohort.gene_ids
    st = CS.StandTable(cohort)
    importances: Dict[str, np.ndarray] = {}
    scores: Dict[str, np.ndarray] = {}            # test-set class-1 probabilities per model (for one-vs-rest voting)
    summary: Dict[str, Dict] = {}
    data_key = f"{cohort.n_genes}x{cohort.n_samples}"

    if cfg.lasso.enabled:
        def lasso_stage():
            reps = L.exclusion_reps(cohort, nreps=cfg.lasso.exclusion_reps, seed=cfg.lasso.seed, alpha=cfg.lasso.alpha,
                                    nfolds=cfg.lasso.nfolds, device=dev)
            return {"reps": reps}
        with log.stage("lasso"):
            res = store.run("models/lasso", cfg.hash(["lasso", "split"]) + data_key, lasso_stage, log)["reps"]
        for i, r in enumerate(res, 1):
            st.add(f"lasso_coef_rep{i}", r.nonzero_coef)
        importances["lasso"] = st.df["lasso_coef_rep1"].to_numpy()
        r1 = res[0]
        CS.write_lasso_gene_csv(r1.nonzero_coef, os.path.join(out, "Logistic_Lasso_RiskGroup_Associated_Genes.csv"))
        summary["lasso"] = {"lambda_min": r1.cv.lambda_min, "lambda_1se": r1.cv.lambda_1se, "test_error": r1.test_error,
                            "auc": r1.test_auc, "confusion": r1.confusion.tolist(), "n_genes": [len(r.nonzero_coef) for r in res],
                            "jerr": r1.jerr}
        R.write_all(os.path.join(out, "report"), {"lasso_rep_curves": R.lasso_rep_curves(res)})

    if cfg.svm.enabled:
        with log.stage("svm"):
            k = 0
            for kern in cfg.svm.kernels:
                for wf in cfg.svm.weightfilt:
                    k += 1
                    r = S.run_svm(xtr, ytr, xte, yte, kern, weightfilt=wf, seed=cfg.svm.seed)
                    check_finite(f"svm{k}_weights", r.weightsvect)
                    # the weights of the model that is actually evaluated (the reference stored the pre-filter vector twice)
                    w = torch.zeros(len(genes), device=dev)
                    w[r.columns] = r.model.weights()
                    st.add(f"svm{k}_weights", w.cpu().numpy())
                    summary[f"svm{k}"] = {"kernel": kern, "weightfilt": wf, "test_error": r.test_error, "auc": r.auc,
                                          "precision": r.precision_test, "recall": r.recall_test}
            importances["svm"] = st.df["svm1_weights"].to_numpy()

    if cfg.rf.enabled:
        with log.stage("random_forest"):
            for nt in cfg.rf.n_trees:
                rf = T.RandomForest(n_trees=nt, proximity=cfg.rf.proximity, seed=cfg.rf.seed).fit(xtr, ytr)
                tag = f"{nt // 1000}k" if nt % 1000 == 0 else str(nt)
                st.add(f"rfnb_{tag}_MeanDecNodeImp", rf.importance_.cpu().numpy())
                pred = rf.predict(xte)
                proba = rf.predict_proba(xte)
                summary[f"rf_{tag}"] = {"oob_error": rf.oob_error_, "test_error": float((pred != yte.long()).float().mean()),
                                        "auc": M.auc(proba, yte)}
                importances["rf"] = rf.importance_.cpu().numpy()
                scores["rf"] = proba.float().cpu().numpy()

    if cfg.gbm.enabled:
        with log.stage("gbm"):
            for i, rep in enumerate(cfg.gbm.reps, 1):
                gb = T.GradientBoosting(n_estimators=int(rep["nrounds"]), learning_rate=cfg.gbm.eta, max_depth=int(rep["max_depth"]),
                                        mode="xgboost").fit(xtr, ytr)
                st.add(f"xg{i}_imp", gb.importance_["gain"].cpu().numpy())
                pr = gb.predict_proba(xte)
                summary[f"xgb{i}"] = {"train_error": gb.train_error_[-1], "test_error": float(((pr > 0.5).float() != yte).float().mean()),
                                      "auc": M.auc(pr, yte), "n_genes": int((gb.importance_["gain"] > 0).sum())}
                importances["xgb"] = gb.importance_["gain"].cpu().numpy()
                scores["xgb"] = pr.float().cpu().numpy()

    if cfg.mlp.enabled:
        with log.stage("mlp"):
            grid = MLP.make_grid(cfg.mlp.hidden, cfg.mlp.lr, cfg.mlp.weight_decay)
            res = MLP.train_grid(xtr, ytr, grid, epochs=cfg.mlp.epochs, x_val=xte, y_val=yte)
            best = int(res.val_auc.argmax())
            st.add("mlp_importance", res.input_importance(best).cpu().numpy())
            importances["mlp"] = res.input_importance(best).cpu().numpy()
            summary["mlp"] = {"best": grid[best], "auc": float(res.val_auc[best]), "val_loss": float(res.val_loss[best])}

    with log.stage("consensus"):
        st.write_csv(os.path.join(out, "standouttable.csv"))
        rank = None
        if len(importances) >= 2:
            rank = CS.aggregate(importances, genes, method=cfg.consensus.method)
            rank.to_csv(os.path.join(out, "consensus_rank.csv"))
            sets = CS.selection_sets({k: pd.Series(v, index=genes) for k, v in importances.items()}, positive_only=False)
            CS.membership_table(sets).to_csv(os.path.join(out, "membership_table.csv"))
            summary["consensus"] = {"top": rank.index[: cfg.consensus.top].tolist(),
                                    "intersections": {k: len(v) for k, v in CS.intersections(sets).items()}}
        rd = cohort.row_data
        if "logFC" in rd and "p.unadj" in rd:
            R.write_all(os.path.join(out, "report"), {"volcano": R.volcano_data(rd["logFC"].to_numpy(float), rd["p.unadj"].to_numpy(float),
                                                                                  rd["p.adj.bh"].to_numpy(float), genes)})
    log.event("run_end", summary=summary)
    if return_details:
        return {"summary": summary, "importances": importances, "scores": scores, "rank": rank}
    return summary


def multiclass_cohort_from_counts(counts: torch.Tensor, gene_ids, sample_ids, label: np.ndarray, train_mask: np.ndarray, cfg: Config,
                                  max_genes_per_class: int = 500, log: Optional[RunLog] = None) -> Cohort:
    """Raw counts + integer class labels -> TMM log2-CPM restricted to the union of the one-vs-rest limma-voom DEGs
    (training samples only; at most ``max_genes_per_class`` per class, by P value)."""
    log = log or RunLog()
    dev = counts.device
    with log.stage("normalize", bytes_moved=float(counts.numel() * 4 * 3)):
        norm = N.tmm_normalize(counts, cfg.normalize.min_cpm, cfg.normalize.min_samples, prior_count=cfg.normalize.prior_count)
    kept_full = np.flatnonzero(norm.keep.cpu().numpy())
    pos = {g: i for i, g in enumerate(kept_full)}
    tr = torch.as_tensor(np.flatnonzero(train_mask), device=dev)
    ctr = counts[:, tr].contiguous()
    chosen: Dict[int, int] = {}
    with log.stage("voom_de_one_vs_rest"):
        for c in np.unique(label):
            de = DE.voom_de(ctr, (label[train_mask] == c).astype(np.int8), min_frac=cfg.de.min_frac, p_value=cfg.de.p_value,
                            lfc=cfg.de.lfc, span=cfg.de.span)
            for g in de.gene_index[de.deg_order][:max_genes_per_class]:
                if g in pos:
                    chosen.setdefault(int(g), int(c))
    rows = [pos[g] for g in sorted(chosen)]
    ids = [gene_ids[g] for g in sorted(chosen)]
    assay = norm.logcpm[torch.as_tensor(rows, device=dev)]
    rd = pd.DataFrame({"ensembl_gene_id": [i.split(".")[0] for i in ids], "first_class": [chosen[g] for g in sorted(chosen)]})
    cd = pd.DataFrame({"class": label.astype(int), "deg.risk": 0, "exptset.seahack": np.where(train_mask, "train", "test")})
    return Cohort(assay.double(), ids, list(sample_ids), rd, cd, {"n_deg": len(ids)})


def run_multiclass(cfg: Config, cohort: Cohort, class_column: str = "class", class_names: Optional[Dict[int, str]] = None,
                   log: Optional[RunLog] = None) -> Dict:
    """Pan-cancer style multiclass run (BASELINE.json config 3): the binary ensemble is fitted one-vs-rest for every
    class, each class gets its own standard table / consensus ranking under ``out_dir/<class>/``, the per-class rankings
    are aggregated into one pan-class consensus list, and samples are assigned to the class whose one-vs-rest models
    give the highest mean probability.  (The reference's notebooks only ever fit the two-group AML risk model; its
    pan-cancer matrix is used for normalisation checks - Normalization_and_Batch_Effect_Investigation.Rmd:49-66.)"""
    import copy
    out = cfg.io.out_dir
    log = log or RunLog(os.path.join(out, "run.jsonl"))
    labels = cohort.col_data[class_column].to_numpy()
    classes = sorted(np.unique(labels).tolist())
    test_mask = (cohort.col_data["exptset.seahack"] == "test").to_numpy()
    per_class, ranks, votes = {}, {}, []
    for c in classes:
        name = (class_names or {}).get(c, str(c))
        sub = copy.copy(cohort)
        sub.col_data = cohort.col_data.copy()
        sub.col_data["deg.risk"] = (labels == c).astype(int)
        cfg_c = copy.deepcopy(cfg)
        cfg_c.io.out_dir = os.path.join(out, f"class_{name}")
        with log.stage(f"one_vs_rest[{name}]"):
            det = run(cfg_c, sub, log=RunLog(os.path.join(cfg_c.io.out_dir, "run.jsonl")), return_details=True)
        per_class[name] = det["summary"]
        if det["rank"] is not None:
            ranks[name] = det["rank"]
        if det["scores"]:
            votes.append(np.mean(np.stack(list(det["scores"].values())), 0))
    summary: Dict = {"classes": [(class_names or {}).get(c, str(c)) for c in classes], "per_class": per_class}
    if votes:
        pred = np.asarray(classes)[np.argmax(np.stack(votes, 1), 1)]
        truth = labels[test_mask]
        summary["accuracy"] = float((pred == truth).mean())
        summary["confusion"] = [[int(((truth == a) & (pred == b)).sum()) for b in classes] for a in classes]
    if ranks:
        # a gene's pan-class score is its best (smallest) consensus rank over the classes; ties broken by the mean rank
        cols = {n: r["mean_rank"] if "mean_rank" in r else r.iloc[:, 0] for n, r in ranks.items()}
        tab = pd.DataFrame({n: v.rank(method="average") for n, v in cols.items()})
        tab["best_rank"], tab["mean_rank"] = tab.min(axis=1), tab.mean(axis=1)
        tab = tab.sort_values(["best_rank", "mean_rank"])
        tab.to_csv(os.path.join(out, "consensus_rank_multiclass.csv"))
        summary["top"] = tab.index[: cfg.consensus.top].tolist()
    log.event("multiclass_end", accuracy=summary.get("accuracy"), n_classes=len(classes))
    return summary
