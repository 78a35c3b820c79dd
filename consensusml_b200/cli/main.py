"""Command line: ``python -m consensusml_b200.cli <command> [--config file] [section.key=value ...]``.

The reference has no CLI - its entry points are "open the notebook and run the chunks" (SURVEY.md 1).  Commands:

  run         the full ensemble + consensus pipeline (default cohort: bundled TARGET-AML; with --counts/--groups: normalise,
              DE-prefilter and model your own two-group cohort, the Differential_Expression_and_Lasso.Rmd flow)
  multiclass  one-vs-rest ensemble + pan-class consensus on a counts CSV with a multi-level group column
  acquire     query the GDC for the TARGET HTSeq count files, write id map + manifest, download, verify, build the counts CSV
              (GDC_Data_Download.Rmd; --endpoint selects the API server, --clinical adds the clinical tables)
  concat      per-sample HTSeq count files -> one genes x samples counts CSV (Cat_Expn_Data.r)
  lasso       the reference's headline model only (runLasso + gene CSV)
  de          limma-voom DE on a counts CSV: writes the DEG table in the reference's format
  normalize   TMM log2-CPM of a counts CSV
  mlseq       the MLSeq / DESeq2 classifier zoo on raw counts (VikasP/AML.R): --counts, --degs, --clinical -> DEG_*.csv + test statistics
  glmboost    glmboost on event-free survival (Ryan/Feature_selection_GLMTrain_WT_TARGET.R): --table -> feature_list.csv
  kmsvm       k-means + GLM / SVM weight tails / QDA (BuieRProj/KM_WHATEVER.Rmd): --features and/or --expr -> output/High*.csv, low*.csv
  build       compile the native libraries
  config      print the resolved configuration
"""
from __future__ import annotations

import argparse
import json
import os
import sys


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="consensusml_b200", description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("command", choices=["run", "multiclass", "acquire", "concat", "lasso", "de", "normalize", "mlseq", "glmboost", "kmsvm", "build", "config"])
    ap.add_argument("--config", default=None)
    ap.add_argument("--counts", default=None, help="genes x samples counts CSV")
    ap.add_argument("--groups", default=None, help="CSV with columns sample,group (1 = non-reference level)")
    ap.add_argument("--files", default=None, help="concat: glob pattern of the per-sample count files (sample id = file name up to the first dot)")
    ap.add_argument("--degs", default=None, help="mlseq: DEG table CSV (first column = gene ids)")
    ap.add_argument("--clinical-csv", default=None, help="mlseq: clinical table (.csv or .xlsx) with TARGET.USI and Risk.group columns")
    ap.add_argument("--table", default=None, help="glmboost: samples x (clinical + genes) CSV with the survival column")
    ap.add_argument("--first-gene-col", type=int, default=41, help="glmboost: 0-based position of the first gene column (the script's column 42)")
    ap.add_argument("--honest", action="store_true", help="glmboost: predict the rows that were NOT used for fitting (the script predicts its training rows)")
    ap.add_argument("--features", default=None, help="kmsvm: patients x binary features CSV with a 'risk' column (0 / 1), or the TARGET-AML clinical .xlsx")
    ap.add_argument("--expr", default=None, help="kmsvm: samples x genes CSV with 'risk' (0 / 1) and 'train' (1 = fit on it) columns")
    ap.add_argument("--endpoint", default=None, help="acquire: GDC API base URL (default https://api.gdc.cancer.gov)")
    ap.add_argument("--clinical", action="store_true", help="acquire: also fetch the per-case clinical tables and the clinical supplements")
    ap.add_argument("--train-frac", type=float, default=2.0 / 3.0, help="run/multiclass on counts: stratified training fraction")
    ap.add_argument("overrides", nargs="*", help="section.key=value")
    args = ap.parse_intermixed_args(argv)
    from ..runtime import config as C
    cfg = C.load(args.config, args.overrides)
    if args.command == "config":
        print(json.dumps(cfg.to_dict(), indent=1))
        return 0
    if args.command == "build":
        from .. import _build
        _build.build_all()
        print(_build.ptxas_report())
        return 0
    if args.command == "acquire":
        from ..io import gdc
        url = args.endpoint or gdc.GDC_API
        counts, idmap = gdc.acquire(cfg.io.out_dir, url)
        path = os.path.join(cfg.io.out_dir, "TARGET_HTSeq_Counts.csv")
        counts.to_csv(path)
        out = {"files": int(len(idmap)), "genes": int(counts.shape[0]), "samples": int(counts.shape[1]), "csv": path}
        if args.clinical:
            full = gdc.clinical(gdc.case_ids("TARGET-AML", url), url)
            full.to_csv(os.path.join(cfg.io.out_dir, "TARGET_AML_clinical_GDC.csv"), index=False)
            sup = gdc.clinical_supplements(os.path.join(cfg.io.out_dir, "Clinical_Data"), url)
            out.update(clinical_rows=int(len(full)), clinical_supplements=len(sup))
        print(json.dumps(out))
        return 0
    import numpy as np
    import pandas as pd
    import torch
    from .. import pipeline
    # under torchrun: one process per GPU; the pipeline spreads its stages over the ranks and rank 0 reports
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = 0
    if world > 1 and args.command in ("run", "multiclass"):
        from ..parallel import dist as PD
        rank, world, ddev = PD.init()
        if cfg.device == "auto":
            cfg.device = str(ddev)
    dev = pipeline.resolve_device(cfg)
    os.makedirs(cfg.io.out_dir, exist_ok=True)
    say = print if rank == 0 else (lambda *a, **k: None)
    if args.command == "concat":
        from ..io import counts as IC
        if not args.files:
            ap.error("'concat' needs --files")
        import glob
        files = sorted(glob.glob(args.files))
        if not files:
            ap.error(f"no file matches {args.files!r}")
        mat, genes, ids = IC.concat_counts(files, regex=r"^([^.]*).*$")
        path = os.path.join(cfg.io.out_dir, "HTSeq_Counts.csv")
        IC.write_matrix_csv(path, torch.as_tensor(mat), genes, ids)
        print(json.dumps({"genes": len(genes), "samples": len(ids), "csv": path}))
        return 0
    if args.command == "run" and args.counts is None:
        summary = pipeline.run(cfg)
        say(json.dumps(summary, indent=1, default=str))
        return 0
    if args.command in ("run", "multiclass"):
        from ..cohort import ops as CO
        from ..io import counts as IC
        if args.groups is None:
            ap.error(f"'{args.command}' on counts needs --groups")
        counts, genes, samples = IC.read_counts_csv(args.counts)
        grp = pd.read_csv(args.groups).set_index("sample").loc[samples, "group"]
        levels = sorted(grp.unique().tolist())
        lab = grp.map({v: i for i, v in enumerate(levels)}).to_numpy()
        train = CO.split_stratified(lab, args.train_frac, seed=cfg.split.seed)
        if args.command == "run":
            if len(levels) != 2:
                ap.error("'run' needs a two-level group column; use 'multiclass'")
            if world > 1:
                # every rank read the CSV; it keeps its own block of samples: normalisation runs sample-sharded and the voom
                # prefilter gene-sharded (stats/sharded.py), every rank ends up with the same DEG cohort
                from ..parallel.dist import shard_bounds
                lo, hi = shard_bounds(len(samples), world, rank)
                co = pipeline.cohort_from_counts_sharded(counts[:, lo:hi].contiguous().to(dev), genes, samples, lab, train, cfg)
            else:
                co = pipeline.cohort_from_counts(counts.to(dev), genes, samples, lab, train, cfg)
            summary = pipeline.run(cfg, co)
        else:
            co = pipeline.multiclass_cohort_from_counts(counts.to(dev), genes, samples, lab, train, cfg)
            summary = pipeline.run_multiclass(cfg, co, class_names={i: str(v) for i, v in enumerate(levels)})
        say(json.dumps(summary, indent=1, default=str))
        return 0
    if args.command == "lasso":
        from .. import consensus, datasets
        from ..models import lasso
        res = lasso.run_lasso(datasets.load_target_aml(), seed=cfg.lasso.seed, alpha=cfg.lasso.alpha, nfolds=cfg.lasso.nfolds, device=dev)
        path = os.path.join(cfg.io.out_dir, "Logistic_Lasso_RiskGroup_Associated_Genes.csv")
        consensus.write_lasso_gene_csv(res.nonzero_coef, path)
        print(json.dumps({"lambda_min": res.cv.lambda_min, "test_error": res.test_error, "auc": res.test_auc,
                          "confusion": res.confusion.tolist(), "genes": len(res.nonzero_coef), "csv": path}))
        return 0
    if args.command in ("mlseq", "glmboost", "kmsvm"):
        from ..api import side_analyses as SA
        if args.command == "mlseq":
            if not (args.counts and args.degs and args.clinical_csv):
                ap.error("'mlseq' needs --counts, --degs and --clinical-csv")
            counts_df = pd.read_csv(args.counts, index_col=0)
            prep = SA.mlseq_prepare(counts_df, pd.read_csv(args.degs).iloc[:, 0].astype(str).tolist(), __import__("consensusml_b200.io.xlsx", fromlist=["read_table"]).read_table(args.clinical_csv))
            out = SA.mlseq_zoo(prep, cfg.io.out_dir, device=dev)
            print(json.dumps({m: {"best_param": r["best_param"], "cv_accuracy": r["cv_accuracy"], **r["test"], "n_selected": len(r["selected_genes"])}
                              for m, r in out["methods"].items()}, indent=1, default=float))
        elif args.command == "glmboost":
            if not args.table:
                ap.error("'glmboost' needs --table")
            out = SA.wilms_glmboost(pd.read_csv(args.table), first_gene_col=args.first_gene_col, honest_test=args.honest, out_dir=cfg.io.out_dir, device=dev)
            print(json.dumps({k: v for k, v in out.items() if k != "features"} | {"n_features": len(out["features"])}, default=float))
        else:
            if not (args.features or args.expr):
                ap.error("'kmsvm' needs --features and/or --expr")
            kw = {}
            if args.features:
                from ..io.xlsx import read_table
                if args.features.lower().endswith(".xlsx"):      # the TARGET clinical workbook itself: the notebook's recoding
                    clin = read_table(args.features, keep_default_na=False)
                    kw.update(binary_features=SA.phenotype_matrix(clin), risk=(clin["Risk group"].astype(str) != "Low").astype(int).to_numpy())
                else:
                    f = read_table(args.features)
                    kw.update(binary_features=f.drop(columns=["risk"]), risk=f["risk"].to_numpy())
            if args.expr:
                e = pd.read_csv(args.expr)
                kw.update(expression=e.drop(columns=["risk", "train"]), expr_risk=e["risk"].to_numpy(), train_flag=e["train"].to_numpy())
            out = SA.km_whatever(out_dir=cfg.io.out_dir, device=dev, **kw)
            print(json.dumps({"kmeans": out["kmeans"].to_dict("records") if "kmeans" in out else None, "svm": out.get("svm"), "qda": out.get("qda")},
                             default=float))
        return 0
    if args.counts is None:
        ap.error(f"'{args.command}' needs --counts")
    from ..io import counts as IC
    counts, genes, samples = IC.read_counts_csv(args.counts)
    counts = counts.to(dev)
    if args.command == "normalize":
        from ..stats import normalize as N
        norm = N.tmm_normalize(counts, cfg.normalize.min_cpm, cfg.normalize.min_samples, prior_count=cfg.normalize.prior_count)
        kept = [g for g, k in zip(genes, norm.keep.cpu().tolist()) if k]
        IC.write_matrix_csv(os.path.join(cfg.io.out_dir, "TMMCPM_log2_Norm_Counts.csv"), norm.logcpm, kept, samples)
        print(json.dumps({"genes_kept": len(kept), "norm_factor_range": [float(norm.norm_factors.min()), float(norm.norm_factors.max())]}))
        return 0
    if args.command == "de":
        if args.groups is None:
            ap.error("'de' needs --groups")
        from ..stats import de as DE
        g = pd.read_csv(args.groups).set_index("sample").loc[samples, "group"].to_numpy()
        res = DE.voom_de(counts, g, min_frac=cfg.de.min_frac, p_value=cfg.de.p_value, lfc=cfg.de.lfc, span=cfg.de.span)
        path = os.path.join(cfg.io.out_dir, "DEGs.csv")
        DE.write_deg_table(res.table(genes), path)
        print(json.dumps({"genes_tested": int(res.keep.sum()), "degs": int(len(res.deg_order)), "csv": path}))
        return 0
    return 1


if __name__ == "__main__":
    sys.exit(main())
