"""Extract the numeric oracles that live in the reference's ``.rda`` checkpoints into small
``.npz``/``.csv`` fixtures under ``consensusml_b200/datasets/target_aml/``.

The GPU box has no ``/root/reference`` mount, so the golden tests and the real-data benchmark read
these fixtures instead.  Only *data* is extracted (expression matrix, labels, split, glmnet
coefficient paths, CV curves, RF/XGB importance vectors) - no reference code.

Sources: composite_code/rnotebook/data/{sesetfilt_degseahack_targetaml,lasso_resultslist,
rf_noboost_2k5k10ktrees_allresultslist,xgb_resultslist,dfens_v95hg38_bmart}.rda and
composite_code/rnotebook/data/standouttable.csv (written at cml_targetaml_seanalysis.Rmd:472,778,1069,1239,1298).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from consensusml_b200.data import Cohort  # noqa: E402
from consensusml_b200.io.rdx import dgcmatrix_to_dense, read_rda, unwrap  # noqa: E402

REF = os.environ.get("CML_REFERENCE_DIR", "/root/reference")
RN = os.path.join(REF, "composite_code/rnotebook")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "consensusml_b200/datasets/target_aml")


def main():
    os.makedirs(OUT, exist_ok=True)
    se = read_rda(os.path.join(RN, "data/sesetfilt_degseahack_targetaml.rda"))["degfilt.se"]
    co = Cohort.from_summarized_experiment(se)
    rd, cd = co.row_data, co.col_data
    np.savez_compressed(
        os.path.join(OUT, "degfilt_1984x137.npz"),
        assay=co.assay.numpy(), gene_ids=np.array(co.gene_ids), sample_ids=np.array(co.sample_ids),
        deg_risk=np.array(cd["deg.risk"], dtype=np.int8), split=np.array([str(x) for x in cd["exptset.seahack"]]),
        hgnc_symbol=np.array([s if s is not None else "" for s in rd["hgnc_symbol"]]),
        hgnc_id=np.array([s if s is not None else "" for s in rd["hgnc_id"]]),
        ensembl_gene_id=np.array([s if s is not None else "" for s in rd["ensembl_gene_id"]]),
        logFC=np.array(rd["logFC"], float), AveExpr=np.array(rd["AveExpr"], float), t=np.array(rd["t"], float),
        p_unadj=np.array(rd["p.unadj"], float), p_adj_bh=np.array(rd["p.adj.bh"], float), b=np.array(rd["b"], float),
        gender=np.array([str(x) for x in cd["Gender"]]) if "Gender" in cd else np.array([]),
        binom_age=np.array([str(x) for x in cd["binom.age"]]) if "binom.age" in cd else np.array([]),
    )
    # clinical table (strings) as CSV for the cohort tests
    cd.to_csv(os.path.join(OUT, "degfilt_coldata.csv"), index=False)

    las = read_rda(os.path.join(RN, "data/lasso_resultslist.rda"))["lasso.resultslist"]
    out = {}
    for i, rep in enumerate(las):
        tf, cv, fm = rep["train.fit"], rep["cv.fit"], rep["final.model"]
        tag = f"rep{i + 1}_"
        bt = dgcmatrix_to_dense(tf["beta"])
        bf = dgcmatrix_to_dense(fm["beta"])
        genes_t = np.array(list(unwrap(tf["beta"].attrs["Dimnames"])[0]))
        nzr = np.nonzero(np.abs(bt).sum(1))[0]
        nzf = np.nonzero(np.abs(bf).sum(1))[0]
        out[tag + "genes"] = genes_t
        out[tag + "train_rows"] = nzr.astype(np.int32)
        out[tag + "train_beta"] = bt[nzr]
        out[tag + "final_rows"] = nzf.astype(np.int32)
        out[tag + "final_beta"] = bf[nzf]
        out[tag + "lambda"] = np.asarray(unwrap(tf["lambda"]), float)
        out[tag + "train_df"] = np.asarray(unwrap(tf["df"]), np.int32)
        out[tag + "train_dev_ratio"] = np.asarray(unwrap(tf["dev.ratio"]), float)
        out[tag + "train_npasses"] = np.asarray(unwrap(tf["npasses"]), np.int32)
        out[tag + "final_npasses"] = np.asarray(unwrap(fm["npasses"]), np.int32)
        out[tag + "cvm"] = np.asarray(unwrap(cv["cvm"]), float)
        out[tag + "cvsd"] = np.asarray(unwrap(cv["cvsd"]), float)
        out[tag + "cv_nzero"] = np.asarray(unwrap(cv["nzero"]), np.int32)
        out[tag + "lambda_min"] = np.asarray(unwrap(cv["lambda.min"]), float)
        out[tag + "lambda_1se"] = np.asarray(unwrap(cv["lambda.1se"]), float)
        out[tag + "confusion"] = np.asarray(unwrap(rep["confusionMatrix"]), np.int32)
        out[tag + "test_error"] = np.asarray(unwrap(rep["test.error"]), float)
        nz = rep["nonzero.coef"]
        out[tag + "nonzero_names"] = np.array(list(unwrap(nz.attrs["names"])))
        out[tag + "nonzero_coef"] = np.asarray(unwrap(nz), float)
    np.savez_compressed(os.path.join(OUT, "glmnet_oracle.npz"), **out)

    rf = read_rda(os.path.join(RN, "data/rf_noboost_2k5k10ktrees_allresultslist.rda"))["rf.returnlist"]
    rfo = {}
    for name, obj in zip(rf.names, rf.value):
        key = name.replace(".", "_")
        try:
            fit = obj[0] if not hasattr(obj, "names") or obj.names is None else obj
            names = obj.names or []
            for nm, val in zip(names, obj.value):
                if hasattr(val, "names") and val.names and "importance" in val.names:
                    imp = val["importance"]
                    rfo[key + "_importance"] = np.asarray(unwrap(imp), float).ravel()
                    rfo[key + "_genes"] = np.array(list(unwrap(imp.attrs["dimnames"])[0]))
                    rfo[key + "_err_rate"] = np.asarray(unwrap(val["err.rate"]), float)
                    rfo[key + "_ntree"] = np.asarray(unwrap(val["ntree"]), np.int32)
                    rfo[key + "_mtry"] = np.asarray(unwrap(val["mtry"]), np.int32)
                    rfo[key + "_confusion_oob"] = np.asarray(unwrap(val["confusion"]), float)
                    rfo[key + "_proximity"] = np.asarray(unwrap(val["proximity"]), np.float32)
                    rfo[key + "_ndbigtree"] = np.asarray(unwrap(val["forest"]["ndbigtree"]), np.int32)
                elif isinstance(unwrap(val), np.ndarray) and unwrap(val).shape == (2, 2):
                    rfo[key + "_" + nm.replace(".", "_")] = np.asarray(unwrap(val), float)
        except Exception as exc:  # keep going: the RF objects are only statistical oracles
            print("rf", name, "skipped:", exc)
    np.savez_compressed(os.path.join(OUT, "rf_oracle.npz"), **rfo)

    xg = read_rda(os.path.join(RN, "data/xgb_resultslist.rda"))["xg.resultslist"]
    xo = {}
    for name, obj in zip(xg.names, xg.value):
        key = name.replace(".", "_")
        def walk(o, prefix, depth=0):
            if depth > 3 or o is None:
                return
            if hasattr(o, "names") and o.names:
                if set(["Feature", "Gain"]).issubset(set(o.names)):
                    xo[prefix + "_feature"] = np.array(list(unwrap(o["Feature"])))
                    for c in ("Gain", "Cover", "Frequency"):
                        if c in o.names:
                            xo[prefix + "_" + c.lower()] = np.asarray(unwrap(o[c]), float)
                    return
                for nm, val in zip(o.names, o.value):
                    if nm in ("raw", "handle", "call", "callbacks"):
                        continue
                    walk(val, prefix + "_" + str(nm).replace(".", "_"), depth + 1)
            else:
                v = unwrap(o)
                if isinstance(v, np.ndarray) and v.dtype.kind in "fi" and v.size <= 64:
                    xo[prefix] = np.asarray(v, float)
        walk(obj, key)
    np.savez_compressed(os.path.join(OUT, "xgb_oracle.npz"), **xo)

    ann = read_rda(os.path.join(RN, "data/dfens_v95hg38_bmart.rda"))["dfens"]
    import pandas as pd
    cols = {}
    for n, v in zip(ann.names, ann.value):
        vv = unwrap(v)
        lev = v.attrs.get("levels") if hasattr(v, "attrs") else None
        if lev is not None:
            lv = list(unwrap(lev))
            vv = [lv[i - 1] if i > 0 else "" for i in np.asarray(vv).tolist()]
        cols[n] = list(vv)
    pd.DataFrame(cols).to_csv(os.path.join(OUT, "gene_annotation_ens95.csv.gz"), index=False)
    print("wrote fixtures to", OUT)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
