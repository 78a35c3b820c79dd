import numpy as np
import pytest
import torch

from consensusml_b200 import cohort as CO
from consensusml_b200.data.cohort import MultiCohort
from consensusml_b200.models import lasso


def test_l1_logistic_matches_sklearn_objective():
    from sklearn.linear_model import LogisticRegression
    g = torch.Generator().manual_seed(0)
    n, p = 73, 300                                       # the walkthrough's 73 training samples
    x = torch.randn(n, p, generator=g, dtype=torch.float64)
    y = (torch.rand(n, generator=g, dtype=torch.float64) < torch.sigmoid(1.5 * x[:, 0] - 1.5 * x[:, 1] + x[:, 2])).double()
    ours = lasso.l1_logistic(x, y, C=1.0)
    sk = LogisticRegression(penalty="l1", solver="saga", C=1.0, max_iter=20000, tol=1e-7).fit(x.numpy(), y.numpy())

    def objective(w, b):
        z = x.numpy() @ w + b
        return np.abs(w).sum() + np.sum(np.log1p(np.exp(-(2 * y.numpy() - 1) * z)))
    o_ours = objective(ours["coef"].numpy(), float(ours["intercept"]))
    o_sk = objective(sk.coef_[0], sk.intercept_[0])
    assert o_ours <= o_sk + 1e-4                           # the path solver reaches (at least) SAGA's optimum
    assert set(ours["selected"].tolist()) >= set(np.flatnonzero(np.abs(sk.coef_[0]) > 1e-3).tolist())
    pos = lasso.l1_logistic(x, y, C=1.0, positive_only=True)
    assert set(pos["selected"].tolist()) == set(torch.nonzero(ours["coef"] > 0).flatten().tolist())


def test_qc_summary_balance_and_multicohort(aml):
    qc = CO.qc_summary(aml.col_data)
    assert qc["n"] == 137 and sum(qc["gender"].values()) == 137
    res = CO.balance_test(aml.col_data["Gender"], aml.col_data["deg.risk"])
    assert abs(res["p_value"] - 0.8044) < 5e-4             # cml_targetaml_seanalysis.Rmd:462
    assert res["table"].to_numpy().tolist() == [[29, 40], [31, 37]]
    res = CO.balance_test(aml.col_data["binom.age"], aml.col_data["deg.risk"])
    assert abs(res["p_value"] - 0.6591) < 5e-4             # :470
    mc = MultiCohort().add("deg", aml).add("deg_copy", aml)
    sub = mc.select_samples(np.arange(10))
    assert sub["deg"].n_samples == 10 and sub["deg_copy"].n_samples == 10
    with pytest.raises(ValueError):
        mc.add("bad", aml.select_samples(np.arange(5)))


def test_boxplot_stats_match_numpy_percentiles():
    """Data behind the per-sample log2-CPM boxplots of the normalisation notebook."""
    import numpy as np
    import torch
    from consensusml_b200.report.dump import boxplot_stats, cohort_group
    rng = np.random.default_rng(0)
    x = rng.normal(size=(500, 6)) + np.arange(6)[None, :] * 0.3
    x[0, 2] = 40.0                                                  # one far outlier
    ids = ["TARGET.20.PAAAAA.09A.01R", "TARGET.30.PABBBB.01A.01R", "TARGET.50.PACCCC.01A.01R", "TARGET-20-PADDDD-03A-01R",
           "TARGET.52.PAEEEE.01A.01R", "X"]
    df = boxplot_stats(torch.from_numpy(x), ids)
    assert df["median"].is_monotonic_increasing
    assert [cohort_group(i) for i in ids] == ["AML", "NBL", "WT", "AML", "RT", "NA"]
    row = df[df["sample"] == ids[2]].iloc[0]
    q1, med, q3 = np.percentile(x[:, 2], [25, 50, 75])
    assert abs(row["median"] - med) < 1e-12 and abs(row["q1"] - q1) < 1e-12 and abs(row["q3"] - q3) < 1e-12
    assert row["whisker_high"] < 40 and row["n_outliers"] >= 1
