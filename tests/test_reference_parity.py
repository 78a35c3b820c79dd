"""Side-by-side runs against the reference's OWN Python helpers, loaded unmodified from ``/root/reference/scripts`` when that
tree is present (it exists on the development box only; elsewhere these tests skip).  ``dummies.py`` lacks its pandas
import (scripts/README.md calls it unused), so ``pd`` is injected into its module namespace - nothing else is touched."""
import importlib.util
import os

import numpy as np
import pandas as pd
import pytest

from consensusml_b200.api import clean, dummies

REF = "/root/reference/scripts"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not available")


def _load(name):
    spec = importlib.util.spec_from_file_location(f"ref_{name}", os.path.join(REF, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _tables(seed=0, n=12, genes=7):
    rng = np.random.default_rng(seed)
    usi = [f"TARGET-20-PA{''.join(rng.choice(list('ABCDEFGH'), 4))}" for _ in range(n)]
    bar = [f"{u}-{rng.choice(['09A', '03A', '01A'])}-01R" for u in usi]
    manifest = pd.DataFrame({"file_id": [f"f{i}" for i in range(n + 3)],
                             "project.project_id": ["TARGET-AML"] * n + ["TARGET-NBL"] * 3,
                             "entity_submitter_id": bar + [f"TARGET-30-PAZZZ{i}-01A-01R" for i in range(3)]})
    clinical = pd.DataFrame({"TARGET USI": usi[:-2] + ["TARGET-20-PANONE"], "Risk group": rng.choice(["Low", "High", "Standard"], n - 1),
                             "Age": rng.integers(1, 20, n - 1)})
    assay = pd.DataFrame(rng.normal(size=(genes, n)), columns=[b.replace("-", ".") for b in bar])
    assay.insert(0, "gene", [f"ENSG{i:05d}" for i in range(genes)])
    return manifest, clinical, assay


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_clean_functions_equal_the_reference(seed):
    ref = _load("clean")
    manifest, clinical, assay = _tables(seed)
    with pd.option_context("mode.chained_assignment", None):
        a = ref.manifest_clinical_merge(manifest.copy(), clinical.copy(), "TARGET-AML")
    b = clean.manifest_clinical_merge(manifest.copy(), clinical.copy(), "TARGET-AML")
    pd.testing.assert_frame_equal(a.reset_index(drop=True), b.reset_index(drop=True), check_dtype=False)
    ta = ref.assay_transpose(assay.copy())
    tb = clean.assay_transpose(assay.copy())
    pd.testing.assert_frame_equal(ta, tb, check_dtype=False, check_names=False)
    assert ref.column_reorder(a) == clean.column_reorder(b)
    ma = ref.assay_clinical_merge(ta.copy(), a.copy())
    mb = clean.assay_clinical_merge(tb.copy(), b.copy())
    pd.testing.assert_frame_equal(ma, mb, check_dtype=False, check_names=False)
    assert list(mb.columns[:2]) == [mb.columns[0], "Diagnostic ID"] and set(mb["Diagnostic ID"]) <= {"09A", "03A", "01A"}


def test_dummies_equal_the_reference():
    ref = _load("dummies")
    ref.pd = pd                                   # the reference file forgot `import pandas as pd`
    train = pd.Series(["a", "b", "a", "c"])
    test = pd.Series(["b", "zzz", "a"], index=[5, 6, 7])
    a = ref.make_dummies(test, train.unique(), "city")
    b = dummies.make_dummies(test, train.unique(), "city")
    pd.testing.assert_frame_equal(a, b, check_dtype=False)
    assert b.loc[6].sum() == 0                     # level unseen in training -> all zeros
    df = pd.DataFrame({"city": ["x", "y", "x"], "phone": ["p", "p", "q"], "v": [1, 2, 3]})
    pd.testing.assert_frame_equal(ref.make_dummies_dataframe(df.copy(), ["city", "phone"]),
                                  dummies.make_dummies_dataframe(df.copy(), ["city", "phone"]), check_dtype=False)


def test_model_comp_data_preparation_equals_the_reference():
    """scripts/model_comp.py imports xgboost (not installed) at module top; a stub module lets the file load unmodified so
    that its pure pandas / sklearn functions (data_prep_columns, model_prep, model_prep_loc) can be compared with ours."""
    import sys
    import types
    from consensusml_b200.api import model_comp as ours
    stub = types.ModuleType("xgboost")
    stub.XGBClassifier = object
    had = sys.modules.get("xgboost")
    sys.modules["xgboost"] = stub
    try:
        ref = _load("model_comp")
    finally:
        if had is None:
            sys.modules.pop("xgboost", None)
        else:
            sys.modules["xgboost"] = had
    rng = np.random.default_rng(3)
    n, genes = 90, 2600
    clin = {f"c{i}": rng.integers(0, 5, n) for i in range(84)}
    clin["c1"] = rng.choice(["09A", "03A", "01A", "11A"], n, p=[0.5, 0.3, 0.15, 0.05])
    clin["c2"] = rng.choice(["Low", "Standard", "High", "Unknown"], n, p=[0.4, 0.3, 0.25, 0.05])
    df = pd.DataFrame(clin).rename(columns={"c1": "Diagnostic ID", "c2": "Risk group"})
    sd = rng.permutation(np.linspace(0.2, 6.0, genes))               # well separated variances: no ordering ties
    g = pd.DataFrame(rng.normal(size=(n, genes)) * sd, columns=[f"ENSG{i:06d}" for i in range(genes)])
    tail = pd.DataFrame({f"t{i}": rng.integers(0, 3, n) for i in range(4)})
    df = pd.concat([df, g, tail], axis=1)
    for var in ("Max", "Min"):
        a_df, a_cols = ref.data_prep_columns(df.copy(), var)
        b_df, b_cols = ours.data_prep_columns(df.copy(), var)
        pd.testing.assert_frame_equal(a_df, b_df, check_dtype=False)
        assert len(a_cols) == len(b_cols) and set(a_cols) == set(b_cols)
        assert a_cols[:1000] == b_cols[:1000]                        # same order within the first class's top-1000
    with pd.option_context("mode.chained_assignment", None):
        ra = ref.model_prep(a_df, a_cols)
    rb = ours.model_prep(b_df, a_cols)
    for x, y in zip(ra[:4], rb[:4]):
        assert np.array_equal(np.asarray(x), np.asarray(y))
    pd.testing.assert_frame_equal(ra[4], rb[4], check_dtype=False)
    names = [a_df.columns[i] for i in a_cols[:50]]
    with pd.option_context("mode.chained_assignment", None):
        la = ref.model_prep_loc(a_df, names)
    lb = ours.model_prep_loc(b_df, names)
    assert np.array_equal(np.asarray(la[0]), np.asarray(lb[0])) and np.array_equal(np.asarray(la[3]), np.asarray(lb[3]))


def test_qc_report_reproduces_the_numbers_committed_in_the_qc_script():
    """aml-target_qc-scripts_SeanMaden.R records its outputs as comments (lines 17-71); ``Clinical_Data/AML_dataframe.csv`` is
    in the reference tree, so they can be checked directly."""
    from consensusml_b200.cohort import ops as CO
    cd = pd.read_csv("/root/reference/Clinical_Data/AML_dataframe.csv", index_col=0)
    q = CO.qc_report(cd)
    assert q["gender"] == {"Female": 96, "Male": 91} and q["cns_disease"] == {"No": 175, "Yes": 12}
    a = q["age_days"]
    assert (a["min"], round(a["q1"]), a["median"], round(a["mean"]), round(a["q3"]), a["max"]) == (137, 966, 3273, 3205, 5251, 8231)
    assert abs(q["age_years"]["median"] - 8.9671) < 1e-4 and abs(q["age_years"]["max"] - 22.5507) < 1e-4
    b = q["bm_blast_pct"]
    assert (b["min"], b["q1"], b["median"], round(b["mean"], 2), b["q3"], b["max"], b["na"]) == (14.0, 58.25, 73.6, 70.92, 89.0, 100.0, 5)
    tab = q["age_class_by_gender"]
    assert tab.loc["old"].to_dict() == {"Female": 46, "Male": 48} and tab.loc["young"].to_dict() == {"Female": 50, "Male": 43}
    assert (q["os_days"]["min"], q["os_days"]["q1"], q["os_days"]["median"], q["os_days"]["max"]) == (112.0, 501.5, 1229.0, 4022.0)
    assert (q["efs_days"]["median"], q["efs_days"]["q3"]) == (392.0, 861.5)
    assert q["os_days_young"]["median"] == 1464 and q["os_days_old"]["median"] == 924 and q["efs_days_old"]["q1"] == 262.5
    bc = CO.barcode_table(cd["entity_submitter_id"])
    assert set(bc["tissue"]) <= {"01A", "02A", "03A", "04A", "06A", "09A", "11A", "40A"} and sum(bc["tissue"].values()) == len(cd)
    g = CO.survival_density_groups(cd)
    assert set(g["group"]) == {"young.overallsurv", "young.efsurv", "old.overallsurv", "old.efsurv"}


def test_gpu_student_t_tail_reproduces_the_reference_deg_table_pvalues():
    """The reference's committed limma table (1998 DEGs: t, P.Value, adj.P.Val) is consistent with one total df (residual 94 +
    prior 1.885) and 17,445 tested genes.  The eBayes kernel's Student-t tail - the same source compiled for the host
    (csrc/tdist.cuh) - must return its P.Value column."""
    import ctypes
    from scipy import optimize, stats
    from consensusml_b200 import _native
    from consensusml_b200.oracle.limma import bh_adjust
    d = pd.read_csv("/root/reference/JSmith_code/Results/TARGET_AML_High.Std.Risk_vs_LowRisk_DEGs.csv", index_col=0)
    t, p = d["t"].to_numpy(), d["P.Value"].to_numpy()
    df = optimize.brentq(lambda v: np.log(2 * stats.t.sf(abs(t[0]), v)) - np.log(p[0]), 5, 5000)
    assert 95.8 < df < 96.0                                        # 96 training samples - 2 groups + df.prior
    out = np.empty_like(t)
    dfv = np.array([df])
    lib = _native.host()
    lib.cml_host_t_two_sided.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    lib.cml_host_t_two_sided.restype = None
    lib.cml_host_t_two_sided(t.ctypes.data, dfv.ctypes.data, 0, t.size, out.ctypes.data)
    assert np.max(np.abs(out / p - 1)) < 1e-9                      # p-values from 5e-15 to 5e-3
    # BH: adj.P.Val = P.Value * n / rank (then a running minimum), with rank counted among ALL tested genes; the table drops
    # genes with |logFC| <= 1, so rank >= row number.  n = 17,445 makes the implied ranks integers, increasing down the table.
    n_tested = int(round(d["adj.P.Val"].iloc[0] / p[0]))
    assert n_tested == 17445
    rank = n_tested * p / d["adj.P.Val"].to_numpy()
    near_int = np.abs(rank - np.round(rank)) < 1e-6
    assert near_int.mean() > 0.5                                   # the other rows take their value from a later gene (running min)
    r = np.round(rank[near_int])
    assert (np.diff(r) > 0).all() and (r >= np.flatnonzero(near_int) + 1).all()
    # and our BH on a vector with those ranks filled in reproduces the column where it is rank-determined
    full = np.ones(n_tested)
    full[(r - 1).astype(int)] = p[near_int]
    order_fill = np.setdiff1d(np.arange(int(r.max())), (r - 1).astype(int))
    full[order_fill] = np.interp(order_fill, (r - 1), p[near_int])          # any monotone filler between known ranks
    ours = bh_adjust(full)[(r - 1).astype(int)]
    assert np.max(np.abs(ours / d["adj.P.Val"].to_numpy()[near_int] - 1)) < 1e-6
