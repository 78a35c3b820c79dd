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
from .models import trees as T_
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


def cohort_from_counts_sharded(counts_local: torch.Tensor, gene_ids, sample_ids, group: np.ndarray, train_mask: np.ndarray, cfg: Config,
                               log: Optional[RunLog] = None) -> Cohort:
    """``cohort_from_counts`` for a SAMPLE-SHARDED count matrix under torchrun: ``counts_local`` holds all genes of this rank's
    block of samples (``parallel.dist.shard_bounds`` over ``len(sample_ids)``); ``gene_ids`` / ``sample_ids`` / ``group`` /
    ``train_mask`` are global.  Normalisation runs sample-sharded and the voom prefilter gene-sharded (stats/sharded.py:
    pass-count all-reduce, reference-column broadcast, shard transpose, five G-length all-gathers); the returned cohort -
    TMM log2-CPM of the DEGs for ALL samples - is the same on every rank, ready for ``run``."""
    import torch.distributed as dist
    from .parallel.dist import shard_bounds
    from .stats import sharded as SH
    log = log or RunLog()
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    S = len(sample_ids)
    lo, hi = shard_bounds(S, world, rank)
    assert counts_local.shape[1] == hi - lo, "counts_local must hold this rank's sample block"
    with log.stage("normalize_sharded", bytes_moved=float(counts_local.numel() * 4 * 3)):
        nrm = SH.tmm_normalize_sharded(counts_local, cfg.normalize.min_cpm, cfg.normalize.min_samples, prior_count=cfg.normalize.prior_count)
    with log.stage("voom_de_sharded"):
        # the prefilter is fitted on the TRAINING samples only (DEL:254): every rank passes its training columns
        tr_local = torch.as_tensor(np.flatnonzero(train_mask[lo:hi]), device=counts_local.device)
        tr_counts = counts_local[:, tr_local].contiguous()
        if world > 1:
            # training samples are not evenly spread over the sample blocks: re-shard them evenly (all-gather of the training
            # columns is G x n_train ints - small next to the cohort - then every rank keeps its own even block)
            parts = [None] * world
            dist.all_gather_object(parts, tr_counts.cpu())
            full = torch.cat(parts, 1)
            a, b = shard_bounds(full.shape[1], world, rank)
            tr_counts = full[:, a:b].contiguous().to(counts_local.device)
        de = SH.voom_de_sharded(tr_counts, group[train_mask], min_frac=cfg.de.min_frac, p_value=cfg.de.p_value, lfc=cfg.de.lfc,
                                span=cfg.de.span)
    deg_full = de.gene_index[de.deg_order]
    kept_full = np.flatnonzero(nrm["keep"].cpu().numpy())
    pos = {g: i for i, g in enumerate(kept_full)}
    rows = [pos[g] for g in deg_full if g in pos]
    local = nrm["logcpm"][torch.as_tensor(rows, device=counts_local.device)]               # [n_deg, S_local]
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, local.cpu())
        assay = torch.cat(parts, 1).to(counts_local.device)
    else:
        assay = local
    ids = [gene_ids[kept_full[r]] for r in rows]
    tab = de.table(gene_ids)
    rd = pd.DataFrame({"hgnc_id": "", "hgnc_symbol": "", "ensembl_gene_id": [i.split(".")[0] for i in ids],
                       "logFC": tab.loc[ids, "logFC"].values, "AveExpr": tab.loc[ids, "AveExpr"].values, "t": tab.loc[ids, "t"].values,
                       "p.unadj": tab.loc[ids, "P.Value"].values, "p.adj.bh": tab.loc[ids, "adj.P.Val"].values, "b": tab.loc[ids, "B"].values})
    cd = pd.DataFrame({"deg.risk": group.astype(int), "exptset.seahack": np.where(train_mask, "train", "test")})
    return Cohort(assay.double(), ids, list(sample_ids), rd, cd, {"n_deg": len(ids), "df_prior": de.df_prior, "sharded_over": world})


def _host(t):
    return t.detach().float().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def _rf_tag(nt: int) -> str:
    return f"{nt // 1000}k" if nt % 1000 == 0 else str(nt)


def build_tasks(cfg: Config, cohort: Cohort, dev: torch.device, store: Optional[ArtifactStore], key: str, log: RunLog):
    """The independent pieces of one ensemble run (see parallel/tasks.py).  Every task returns host-side, picklable data.
    The device copies of the design matrices are created lazily - a rank only uploads what its own tasks need."""
    from .parallel.tasks import Task
    cache: Dict[str, object] = {}

    def data():
        if "xy" not in cache:
            tr, te = cohort.split()
            cache["xy"] = (tr.design(torch.float32).to(dev, non_blocking=True), tr.labels().to(dev, non_blocking=True),
                           te.design(torch.float32).to(dev, non_blocking=True), te.labels().to(dev, non_blocking=True))
        return cache["xy"]

    def checkpointed(name, fn):
        if store is None:
            return fn
        return lambda: store.run(f"tasks/{name}", key, lambda: {"v": fn()}, log)["v"]

    tasks = []

    def add(name, cost, fn, group):
        tasks.append(Task(name, cost, checkpointed(name, fn), group))

    if cfg.lasso.enabled:
        def lasso_task():
            reps = L.exclusion_reps(cohort, nreps=cfg.lasso.exclusion_reps, seed=cfg.lasso.seed, alpha=cfg.lasso.alpha,
                                    nfolds=cfg.lasso.nfolds, device=dev)
            r1 = reps[0]
            return {"nonzero": [r.nonzero_coef for r in reps], "curves": R.lasso_rep_curves(reps),
                    "test_error": [r.test_error for r in reps], "jerr": [r.jerr for r in reps],
                    "rep1": {"lambda_min": r1.cv.lambda_min, "lambda_1se": r1.cv.lambda_1se, "test_error": r1.test_error,
                             "auc": r1.test_auc, "logloss": r1.test_logloss, "confusion": r1.confusion.tolist(),
                             "train_npasses": r1.train_npasses, "final_npasses": r1.final_npasses}}
        # the exclusion chain is sequential (each replicate removes the genes the previous ones selected): one task
        add("lasso", 4.0 + 0.75 * cfg.lasso.exclusion_reps, lasso_task, "lasso")

    if cfg.lasso.enabled and cfg.lasso.stability_boot > 0:
        # stability selection (north-star config 5): bootstrap replicates x alpha grid, each replicate a full 100-lambda path.
        # Replicates are independent: fixed chunks of them are tasks, their selection counts add up.
        from .ops import lasso as KL
        lam_grid = KL.default_lambda_grid()
        chunk = max(1, int(cfg.lasso.stability_chunk))
        for ai, a in enumerate(cfg.lasso.stability_alphas):
            for c0 in range(0, cfg.lasso.stability_boot, chunk):
                nb = min(chunk, cfg.lasso.stability_boot - c0)

                def stab_task(ai=ai, a=a, c0=c0, nb=nb):
                    from .ops.replicate import replicate_weights
                    xtr, ytr, _, _ = data()
                    n, p = xtr.shape
                    on_dev = KL._native.on_device(xtr)
                    packed = KL.pack_design(xtr) if on_dev else None
                    n_pad = packed[1] if on_dev else n
                    freq = torch.zeros(len(lam_grid), p, dtype=torch.float32, device=dev)
                    w = replicate_weights(nb, n, mode="bootstrap", seed=cfg.lasso.seed + 7919 * ai, device=dev, n_pad=n_pad, rep0=c0)
                    KL.lognet_path(xtr, ytr, w[:, :n], lam_grid, alpha=float(a), want_eta=False, packed=packed, freq_out=freq,
                                   padded_weights=w if on_dev else None, want_coef=False)
                    return {"counts": _host(freq), "n": nb}
                add(f"stab_a{ai}_b{c0}", 0.12 * nb / 250.0, stab_task, f"stab_a{ai}")

    if cfg.svm.enabled:
        k = 0
        for kern in cfg.svm.kernels:
            for wf in cfg.svm.weightfilt:
                k += 1

                def svm_task(kern=kern, wf=wf, k=k):
                    xtr, ytr, xte, yte = data()
                    r = S.run_svm(xtr, ytr, xte, yte, kern, weightfilt=wf, seed=cfg.svm.seed)
                    check_finite(f"svm{k}_weights", r.weightsvect)
                    # the weights of the model that is actually evaluated (the reference stored the pre-filter vector twice)
                    w = torch.zeros(cohort.n_genes, device=dev)
                    w[r.columns] = r.model.weights()
                    return {"weights": _host(w), "score": _host(torch.sigmoid(r.decision_test)),
                            "summary": {"kernel": kern, "weightfilt": wf, "test_error": r.test_error, "auc": r.auc,
                                        "precision": r.precision_test, "recall": r.recall_test, "iters": r.model.iters_}}
                add(f"svm{k}", 1.0 if not wf else 1.5, svm_task, "svm")

    if cfg.rf.enabled:
        chunk = max(1, int(cfg.rf.chunk))
        for nt in cfg.rf.n_trees:
            tag = _rf_tag(nt)
            for c0 in range(0, nt, chunk):
                T = min(chunk, nt - c0)

                def rf_task(nt=nt, c0=c0, T=T):
                    xtr, ytr, xte, _ = data()
                    pk = ("rf", nt)
                    if pk not in cache:          # binning + labels once per forest and rank
                        cache[pk] = T_.RandomForest(n_trees=nt, proximity=cfg.rf.proximity, seed=cfg.rf.seed).begin(xtr, ytr)
                    part = cache[pk].grow_chunk(c0, T, x_eval=xte, keep_trees=False)
                    return {kk: _host(v) for kk, v in part.items()}
                add(f"rf_{tag}_t{c0}", 1.2 * T / 1000.0, rf_task, f"rf_{tag}")

    if cfg.gbm.enabled:
        for i, rep in enumerate(cfg.gbm.reps, 1):
            def gbm_task(rep=rep):
                xtr, ytr, xte, yte = data()
                gb = T_.GradientBoosting(n_estimators=int(rep["nrounds"]), learning_rate=cfg.gbm.eta, max_depth=int(rep["max_depth"]),
                                         mode="xgboost").fit(xtr, ytr)
                pr = gb.predict_proba(xte)
                gain = gb.importance_["gain"]
                return {"gain": _host(gain), "score": _host(pr),
                        "summary": {"train_error": gb.train_error_[-1], "test_error": float(((pr > 0.5).float() != yte).float().mean()),
                                    "auc": M.auc(pr, yte), "n_genes": int((gain > 0).sum())}}
            add(f"xgb{i}", 0.05 * float(rep["nrounds"]) * min(float(rep["max_depth"]), 8.0) / 4.0 + 0.2, gbm_task, "gbm")

    if cfg.mlp.enabled:
        def mlp_task():
            xtr, ytr, xte, yte = data()
            grid = MLP.make_grid(cfg.mlp.hidden, cfg.mlp.lr, cfg.mlp.weight_decay)
            # the grid point is chosen on an inner split of the TRAINING samples; the held-out test set is only used to
            # report that single choice (lasso / SVM / RF / GBM never see the test set either)
            n = xtr.shape[0]
            perm = torch.randperm(n, generator=torch.Generator().manual_seed(cfg.mlp.seed)).to(dev)
            n_val = max(2, int(round(cfg.mlp.val_frac * n)))
            iv, it = perm[:n_val], perm[n_val:]
            sel = MLP.train_grid(xtr[it], ytr[it], grid, epochs=cfg.mlp.epochs, x_val=xtr[iv], y_val=ytr[iv])
            best = int(torch.argmin(sel.val_loss))
            res = MLP.train_grid(xtr, ytr, [grid[best]], epochs=cfg.mlp.epochs, x_val=xte, y_val=yte)
            return {"importance": _host(res.input_importance(0)), "score": _host(res.predict_proba(xte, 0)),
                    "summary": {"best": grid[best], "inner_val_loss": float(sel.val_loss[best]), "inner_val_auc": float(sel.val_auc[best]),
                                "auc": float(res.val_auc[0]), "test_loss": float(res.val_loss[0])}}
        add("mlp", 2.0, mlp_task, "mlp")
    return tasks


def run(cfg: Config, cohort: Optional[Cohort] = None, log: Optional[RunLog] = None, return_details: bool = False) -> Dict:
    """Run the ensemble + consensus on ``cohort`` (default: the bundled TARGET-AML DEG cohort).  Returns a summary dict
    and writes ``standouttable.csv``, the lasso gene CSV, ``consensus_rank.csv`` and the report dumps.

    Under ``torchrun`` (``torch.distributed`` initialised, ``cfg.dist.enabled``) the independent stages are spread over
    the ranks (parallel/tasks.py); every rank returns the same summary, rank 0 writes the files.  The result does not
    depend on the number of ranks."""
    import torch.distributed as dist
    from .parallel.tasks import run_tasks
    dev = resolve_device(cfg)
    out = cfg.io.out_dir
    multi = cfg.dist.enabled and dist.is_initialized() and dist.get_world_size() > 1
    rank = dist.get_rank() if multi else 0
    log = log or RunLog(os.path.join(out, "run.jsonl") if cfg.write_outputs else None, rank=rank)
    store = ArtifactStore(out, resume=cfg.resume, rank=rank) if cfg.checkpoint else None
    cohort = cohort or datasets.load_target_aml()
    fp = cohort.fingerprint()
    log.event("run_start", device=str(dev), genes=cohort.n_genes, samples=cohort.n_samples, config_hash=cfg.hash(),
              cohort_fingerprint=fp, world=dist.get_world_size() if multi else 1)
    genes = cohort.gene_ids
    # a checkpoint is only reused for the same configuration AND the same data (ids, labels, split and assay content)
    key = cfg.hash(["lasso", "svm", "rf", "gbm", "mlp", "split"]) + ":" + fp
    tasks = build_tasks(cfg, cohort, dev, store, key, log)
    results, report = run_tasks(tasks, enabled=cfg.dist.enabled, fail_fast=cfg.fail_fast, streams=cfg.dist.task_streams)
    from .parallel.tasks import TaskError
    failed = {n: r.error for n, r in results.items() if isinstance(r, TaskError)}
    for t in tasks:
        log.event("task", task=t.name, group=t.group, rank=report.assignment[t.name], seconds=report.seconds.get(t.name),
                  **({"error": failed[t.name]} if t.name in failed else {}))
    ok = lambda name: name not in failed                      # stages that failed are left out of the table and reported

    st = CS.StandTable(cohort)
    importances: Dict[str, np.ndarray] = {}
    scores: Dict[str, np.ndarray] = {}            # test-set class-1 probabilities per model (for one-vs-rest voting)
    summary: Dict[str, Dict] = {}
    tr_c, te_c = cohort.split()
    yte_np = te_c.labels().cpu().numpy()

    if cfg.lasso.enabled and ok("lasso"):
        lr = results["lasso"]
        for i, nz in enumerate(lr["nonzero"], 1):
            st.add(f"lasso_coef_rep{i}", nz)
        importances["lasso"] = st.df["lasso_coef_rep1"].to_numpy()
        summary["lasso"] = {**lr["rep1"], "n_genes": [len(nz) for nz in lr["nonzero"]], "test_errors": lr["test_error"],
                            "jerr": lr["jerr"][0], "genes_rep1": sorted(lr["nonzero"][0])}
    if cfg.lasso.enabled and cfg.lasso.stability_boot > 0:
        maxf = np.zeros(len(genes))
        per_alpha = {}
        for ai, a in enumerate(cfg.lasso.stability_alphas):
            names = [t.name for t in tasks if t.group == f"stab_a{ai}"]
            if not all(ok(nm) for nm in names):
                continue
            cnt = sum(results[nm]["counts"] for nm in names)
            nrep = sum(results[nm]["n"] for nm in names)
            f = (cnt / max(nrep, 1)).max(0)                          # max over the lambda grid of the selection frequency
            per_alpha[str(a)] = int((f >= cfg.lasso.stability_threshold).sum())
            maxf = np.maximum(maxf, f)
        if per_alpha:
            st.add("stability_max_freq", maxf)
            importances["stability"] = maxf
            summary["stability"] = {"n_boot": cfg.lasso.stability_boot, "alphas": list(cfg.lasso.stability_alphas),
                                    "stable_genes_per_alpha": per_alpha, "stable_genes": int((maxf >= cfg.lasso.stability_threshold).sum()),
                                    "threshold": cfg.lasso.stability_threshold}
    if cfg.svm.enabled:
        k = 0
        for kern in cfg.svm.kernels:
            for wf in cfg.svm.weightfilt:
                k += 1
                if not ok(f"svm{k}"):
                    continue
                r = results[f"svm{k}"]
                st.add(f"svm{k}_weights", r["weights"])
                summary[f"svm{k}"] = r["summary"]
                scores.setdefault("svm", r["score"])
                importances.setdefault("svm", r["weights"])
    if cfg.rf.enabled:
        from .models.trees import RandomForest
        ytr_t = tr_c.labels()
        for nt in cfg.rf.n_trees:
            tag = _rf_tag(nt)
            if not all(ok(t.name) for t in tasks if t.group == f"rf_{tag}"):
                continue
            parts = [results[t.name] for t in tasks if t.group == f"rf_{tag}"]         # task order = tree order
            rf = RandomForest(n_trees=nt, proximity=cfg.rf.proximity, seed=cfg.rf.seed)
            rf._bins, rf._yf = torch.zeros(1), ytr_t.float().cpu()                      # merge on the host (parts are host arrays)
            rf.finish(parts)
            imp = rf.importance_.numpy()
            st.add(f"rfnb_{tag}_MeanDecNodeImp", imp)
            proba = rf.eval_proba_.numpy()
            summary[f"rf_{tag}"] = {"oob_error": rf.oob_error_, "test_error": float(((proba > 0.5) != (yte_np > 0.5)).mean()),
                                    "auc": M.auc(torch.from_numpy(proba), torch.from_numpy(yte_np)), "mean_nodes": rf.mean_nodes_}
            importances["rf"] = imp
            scores["rf"] = proba
            if cfg.rf.proximity and cfg.write_outputs and rank == 0:
                R.write_all(os.path.join(out, "report"), {}, {f"rf_{tag}_proximity": {"proximity": rf.proximity_.numpy()}})
    if cfg.gbm.enabled:
        for i, _ in enumerate(cfg.gbm.reps, 1):
            if not ok(f"xgb{i}"):
                continue
            r = results[f"xgb{i}"]
            st.add(f"xg{i}_imp", r["gain"])
            summary[f"xgb{i}"] = r["summary"]
            importances["xgb"] = r["gain"]
            scores["xgb"] = r["score"]
    if cfg.mlp.enabled and ok("mlp"):
        r = results["mlp"]
        st.add("mlp_importance", r["importance"])
        importances["mlp"] = r["importance"]
        scores["mlp"] = r["score"]
        summary["mlp"] = r["summary"]

    rank_tab = None
    if len(importances) >= 2:
        rank_tab = CS.aggregate(importances, genes, method=cfg.consensus.method)
        sets = CS.selection_sets({k: pd.Series(v, index=genes) for k, v in importances.items()}, positive_only=False)
        summary["consensus"] = {"top": rank_tab.index[: cfg.consensus.top].tolist(),
                                "intersections": {k: len(v) for k, v in CS.intersections(sets).items()}}
    if failed:
        summary["failed_tasks"] = failed
    summary["schedule"] = {"critical": report.critical(), "rank_seconds": report.rank_seconds,
                           "group_seconds": {g: sum(report.seconds.get(t.name, 0.0) for t in tasks if t.group == g)
                                             for g in sorted({t.group for t in tasks})}}
    if cfg.write_outputs and rank == 0:
        with log.stage("write_outputs"):
            os.makedirs(out, exist_ok=True)
            st.write_csv(os.path.join(out, "standouttable.csv"))
            if cfg.lasso.enabled and ok("lasso"):
                CS.write_lasso_gene_csv(results["lasso"]["nonzero"][0], os.path.join(out, "Logistic_Lasso_RiskGroup_Associated_Genes.csv"))
                R.write_all(os.path.join(out, "report"), {"lasso_rep_curves": results["lasso"]["curves"]})
            if rank_tab is not None:
                rank_tab.to_csv(os.path.join(out, "consensus_rank.csv"))
                CS.membership_table(sets).to_csv(os.path.join(out, "membership_table.csv"))
            rd = cohort.row_data
            if "logFC" in rd and "p.unadj" in rd:
                R.write_all(os.path.join(out, "report"), {"volcano": R.volcano_data(rd["logFC"].to_numpy(float), rd["p.unadj"].to_numpy(float),
                                                                                      rd["p.adj.bh"].to_numpy(float), genes)})
    log.event("run_end", summary=summary)
    if return_details:
        return {"summary": summary, "importances": importances, "scores": scores, "rank": rank_tab, "standtable": st.df}
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
