import filecmp
import os

import numpy as np
import pandas as pd
import pytest
import torch

from consensusml_b200 import consensus as C
from consensusml_b200.consensus import rank_aggregate as RA
from consensusml_b200.models import lasso

REF_CSV = "/root/reference/TARGET_AML_Logistic_Lasso_withDEGs_RiskGroup_Associated_Genes.csv"


def test_lasso_gene_csv_is_byte_identical_to_reference(aml, tmp_path):
    res = lasso.run_lasso(aml)
    out = tmp_path / "genes.csv"
    df = C.write_lasso_gene_csv(res.nonzero_coef, str(out))
    assert list(df["Gene.name"])[:3] == ["HOXA9", "PLCB4", "GOLGA8M"]
    assert len(df) == 13
    if os.path.exists(REF_CSV):
        # identical layout, quoting and annotation; coefficients agree to the 15th significant digit, where the
        # summation order of glmnet's Fortran and of our solver differ by an ulp
        ours, ref = out.read_text().splitlines(), open(REF_CSV).read().splitlines()
        assert len(ours) == len(ref) == 14 and ours[0] == ref[0]
        for a, b in zip(ours[1:], ref[1:]):
            fa, fb = a.split(","), b.split(",")
            assert fa[:2] == fb[:2] and fa[3:] == fb[3:]
            assert abs(float(fa[2]) - float(fb[2])) < 1e-12
    else:
        first = out.read_text().splitlines()
        assert first[0].startswith('"","Gene_ID","Mod.RG.nonzero.coef","Gene.name"')
        assert first[1] == '"1","ENSG00000078399",0.218496737198212,"HOXA9",7,27162435,27175180,"p15.2"'


def test_sets_and_membership():
    imp = {"xgb": pd.Series([0.5, 0, 0.1, 0], index=list("abcd")), "rf": pd.Series([0.2, 0.3, 0, 0], index=list("abcd")),
           "lasso": pd.Series([1.0, -0.5, 0.2, 0], index=list("abcd"))}
    s = C.selection_sets(imp)
    assert s["lasso"] == {"a", "c"}                      # positive coefficients only, like the reference
    assert C.selection_sets(imp, positive_only=False)["lasso"] == {"a", "b", "c"}
    inter = C.intersections(s)
    assert inter["xgb&rf"] == {"a"} and inter["xgb&rf&lasso"] == {"a"} and inter["xgb&lasso"] == {"a", "c"}
    tab = C.membership_table(s)
    assert tab.index[0] == "a" and tab.loc["a", "n_models"] == 3 and "d" not in tab.index


def test_rank_aggregation_properties():
    g = torch.Generator().manual_seed(0)
    p, m = 200, 5
    imp = torch.rand(p, m, generator=g, dtype=torch.float64)
    imp[7] = 10.0                                          # top in every model
    imp[11, :4] = 9.0                                      # top in 4 of 5
    ranks = RA.rank_matrix(imp)
    assert torch.allclose(ranks.sum(0), torch.full((m,), p * (p + 1) / 2, dtype=torch.float64))
    rho = RA.robust_rank_aggregation(ranks)
    assert int(torch.argmin(rho)) == 7 and rho[11] < 1e-4 and float(rho.median()) > 0.2
    perm = torch.randperm(m, generator=g)
    assert torch.allclose(RA.robust_rank_aggregation(ranks[:, perm]), rho)      # model order does not matter
    df = C.aggregate({f"m{j}": imp[:, j].numpy() for j in range(m)}, [f"g{i}" for i in range(p)], method="mean_rank")
    assert df.index[0] == "g7"
    assert C.top_k_overlap(df.index, df.index, 50) == 1.0


def test_standtable_schema_and_roundtrip(aml, tmp_path):
    from consensusml_b200.consensus.standtable import BASE_COLUMNS, REFERENCE_COLUMNS
    st = C.StandTable(aml)
    st.add("lasso_coef_rep1", {"ENSG00000078399.14": 0.2185})
    st.add("svm1_weights", np.arange(1984, dtype=float))
    assert list(st.df.columns[:9]) == BASE_COLUMNS
    out = tmp_path / "standouttable.csv"
    st.write_csv(str(out))
    back = C.StandTable.read_csv(str(out))
    assert back.shape == (1984, 11)
    assert back.loc["ENSG00000078399.14", "lasso_coef_rep1"] == 0.2185 and (back["lasso_coef_rep1"] != 0).sum() == 1
    ref = "/root/reference/composite_code/rnotebook/data/standouttable.csv"
    if os.path.exists(ref):
        r = pd.read_csv(ref, index_col=0)
        assert list(r.columns) == REFERENCE_COLUMNS
        assert list(r.index) == list(back.index)
        assert np.allclose(r["logFC"], back["logFC"])
        assert out.read_text().splitlines()[0].split(",")[:10] == open(ref).readline().rstrip("\n").split(",")[:10]
