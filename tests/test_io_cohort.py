import os

import numpy as np
import pandas as pd
import pytest
import torch

from consensusml_b200 import cohort as CO
from consensusml_b200.io import counts as IC

REF = "/root/reference"


def test_concat_counts_native_parser(tmp_path):
    rng = np.random.default_rng(0)
    genes = [f"ENSG{i:011d}.{i % 7 + 1}" for i in range(500)] + ["__no_feature", "__ambiguous"]
    mat = rng.poisson(30, (len(genes), 7))
    files = []
    for s in range(7):
        p = tmp_path / f"{s:02d}_TARGET-20-PAAAA{s}-09A-01R.htseq.counts"
        with open(p, "w") as fh:
            for g, v in zip(genes, mat[:, s]):
                fh.write(f"{g}\t{v}\n")
        files.append(str(p))
    counts, gids, sids = IC.concat_counts(files, regex=r"^\d+_(TARGET-[0-9A-Z-]+)\.htseq\.counts$")
    assert counts.shape == (500, 7) and counts.dtype == torch.int32
    assert gids == genes[:500] and sids[3] == "TARGET-20-PAAAA3-09A-01R"
    assert np.array_equal(counts.numpy(), mat[:500])
    # a file whose gene order differs is rejected with its name
    bad = tmp_path / "bad.counts"
    bad.write_text("".join(f"{g}\t1\n" for g in reversed(genes)))
    with pytest.raises(ValueError, match="bad.counts"):
        IC.concat_counts(files + [str(bad)])
    out = tmp_path / "m.csv"
    IC.write_matrix_csv(str(out), counts, gids, sids)
    c2, g2, s2 = IC.read_counts_csv(str(out))
    assert torch.equal(c2, counts) and g2 == gids and s2 == sids


def test_label_recoding_and_filters():
    df = pd.DataFrame({"Diagnostic ID": ["09A", "03A", "04A", "40A", "01A"], "Risk group": ["Low", "High", "Standard", "Unknown", "Low"],
                       "Age at Diagnosis in Days": [100, 200, 300, 400, 500], "Primary Cytogenetic Code": ["MLL", "Normal", "Unknown", "MLL", "inv(16)"]})
    kept = CO.filter_diagnostic(df)
    assert list(kept["Diagnostic ID"]) == ["09A", "03A", "01A"]
    assert list(CO.filter_diagnostic(df, keep=CO.PRIMARY)["Diagnostic ID"]) == ["09A", "03A"]
    r = CO.recode_risk(df)
    assert list(r["RiskGroup.Class"]) == ["No", "Yes", "Yes", "Unknown", "No"]
    assert list(r["deg.risk"]) == [0, 1, 1, -1, 0] and list(r["Low Risk"]) == [1, 0, 0, 0, 1]
    assert list(CO.age_class(df)) == ["No", "No", "No", "Yes", "Yes"]
    assert list(CO.mll_update(df)) == ["MLL", "No", "Unknown", "MLL", "No"]
    assert CO.dedupe_samples(["TARGET-20-PABLDZ-09A-01R", "TARGET-20-PABLDZ-03A-01R", "TARGET-20-PADYIR-09A-01R"]).tolist() == [0, 2]


def test_splits_and_resampling():
    ids = [f"s{i}" for i in range(145)]
    tr, te = CO.split_random(ids, 2 / 3, seed=2019)
    assert len(tr) == 96 and len(te) == 49 and not set(tr) & set(te)       # the reference's 96 / 49
    y = np.array([0] * 60 + [1] * 40)
    m = CO.split_stratified(y, 0.7)
    assert m.sum() == 70 and y[m].sum() == 28
    a, b, c = CO.holdout_then_split(137)
    assert len(c) == 27 and len(a) + len(b) + len(c) == 137 and len(b) == 37   # 27 / 73 / 37 as in the walkthrough
    idx = CO.resample_balanced(np.array([0] * 90 + [1] * 10), n_out=1000, p=0.5)
    frac = (np.array([0] * 90 + [1] * 10)[idx] == 1).mean()
    assert idx.size == 1000 and 0.44 < frac < 0.56


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted")
def test_fixed_split_reproduces_reference_flags(aml):
    tr = pd.read_csv(f"{REF}/JSmith_code/TARGET_AML_Training_Samples.csv")["x"].tolist()
    te = pd.read_csv(f"{REF}/JSmith_code/TARGET_AML_Testing_Samples.csv")["x"].tolist()
    assert len(tr) == 96 and len(te) == 49
    out = CO.split_fixed(aml, tr, te, column="split2")
    assert (out.col_data["split2"].to_numpy() == aml.col_data["exptset.seahack"].to_numpy()).all()   # 93 / 44 after Unknown drop
    assert (out.col_data["split2"] == "train").sum() == 93


def test_read_manifest_and_merge(tmp_path):
    """GDC id-map CSV -> frame usable by api.clean.manifest_clinical_merge (GDC_Data_Download.Rmd:219-229, scripts/clean.py:1-18)."""
    import pandas as pd
    from consensusml_b200.api import clean
    from consensusml_b200.io.counts import read_manifest
    p = tmp_path / "manifest.csv"
    pd.DataFrame({"file_id": ["f1", "f2", "f3"], "project.project_id": ["TARGET-AML", "TARGET-AML", "TARGET-NBL"],
                  "entity_submitter_id": ["TARGET-20-PABCDE-09A-01R", "TARGET-20-PAXYZW-03A-01R", "TARGET-30-PAQQQQ-01A-01R"]}).to_csv(p, index=False)
    man = read_manifest(str(p))
    assert list(man.columns) == ["file_id", "project.project_id", "entity_submitter_id"] and len(man) == 3
    clin = pd.DataFrame({"TARGET USI": ["TARGET-20-PABCDE", "TARGET-20-PAXYZW"], "Risk group": ["Low", "High"]})
    merged = clean.manifest_clinical_merge(man, clin, "TARGET-AML")
    assert len(merged) == 2 and set(merged["Risk group"]) == {"Low", "High"}


def test_native_csv_matrix_reader_matches_pandas(tmp_path):
    """csrc/host/csv_matrix.cpp: quoted ids and values, CRLF, "12.0", negative numbers, blank lines; non-integer files fall back."""
    import numpy as np
    import pandas as pd
    import torch
    from consensusml_b200.io import counts as IC
    rng = np.random.default_rng(0)
    m = rng.integers(0, 500000, size=(257, 13)).astype(np.int32)
    m[3, 4] = -7
    genes = [f"ENSG{i:011d}.{i % 17}" for i in range(257)]
    samples = [f"TARGET.20.PA{i:04d}.09A.01R" for i in range(13)]
    p1 = tmp_path / "plain.csv"
    IC.write_matrix_csv(str(p1), torch.from_numpy(m), genes, samples)
    a = IC._read_counts_csv_native(str(p1), threads=3)
    b = IC.read_counts_csv(str(p1), native=False)
    assert a is not None and torch.equal(a[0], b[0]) and a[1] == b[1] == genes and a[2] == b[2] == samples
    p2 = tmp_path / "quoted_crlf.csv"
    with open(p2, "w", newline="") as fh:                       # R's write.csv style, Windows line ends, a blank line
        fh.write(",".join(['""'] + [f'"{s}"' for s in samples]) + "\r\n")
        for i in range(257):
            fh.write(",".join([f'"{genes[i]}"'] + [f"{v}.0" if j % 2 else f'"{v}"' for j, v in enumerate(m[i])]) + "\r\n")
            if i == 100:
                fh.write("\r\n")
    c = IC.read_counts_csv(str(p2))
    assert torch.equal(c[0], torch.from_numpy(m)) and c[1] == genes and c[2] == samples
    p3 = tmp_path / "floats.csv"
    pd.DataFrame(m[:5] + 0.5, index=genes[:5], columns=samples).to_csv(p3)
    assert IC._read_counts_csv_native(str(p3)) is None          # not an integer matrix: the general reader takes over
    p4 = tmp_path / "ragged.csv"
    p4.write_text("id,a,b\nx,1,2\ny,3\n")
    assert IC._read_counts_csv_native(str(p4)) is None
    assert IC._read_counts_csv_native(str(tmp_path / "missing.csv")) is None


def test_native_csv_matrix_writer_round_trips(tmp_path):
    """Parallel writer: doubles in shortest round-trip form (exact when read back), integers, fallback for awkward ids."""
    import numpy as np
    import pandas as pd
    import torch
    from consensusml_b200.io import counts as IC
    rng = np.random.default_rng(1)
    x = rng.normal(size=(1300, 9)) * np.array([1e-12, 1e-3, 1, 1, 1e3, 1e8, 1e18, 1, 1])
    x[0, 0], x[1, 1], x[2, 2], x[3, 3] = np.nan, 1e-300, 5.0, -0.0
    rows, cols = [f"ENSG{i:05d}.3" for i in range(1300)], [f"S{j}" for j in range(9)]
    IC.write_matrix_csv(str(tmp_path / "f.csv"), torch.from_numpy(x), rows, cols, threads=3)
    d = pd.read_csv(tmp_path / "f.csv", index_col=0, float_precision="round_trip")
    assert list(d.index) == rows and list(d.columns) == cols and np.array_equal(d.to_numpy(), x, equal_nan=True)
    m = rng.integers(-5, 10 ** 9, size=(700, 4))
    IC.write_matrix_csv(str(tmp_path / "i.csv"), torch.from_numpy(m), rows[:700], cols[:4])
    back = IC.read_counts_csv(str(tmp_path / "i.csv"))
    assert torch.equal(back[0], torch.from_numpy(m).int()) and back[1] == rows[:700] and back[2] == cols[:4]
    IC.write_matrix_csv(str(tmp_path / "q.csv"), torch.from_numpy(x[:3]), ['a,b', 'c "d"', "e"], cols)     # pandas fallback quotes them
    q = pd.read_csv(tmp_path / "q.csv", index_col=0)
    assert list(q.index) == ['a,b', 'c "d"', "e"]
    with pytest.raises(OSError):
        IC.write_matrix_csv(str(tmp_path / "no_such_dir" / "x.csv"), torch.from_numpy(x[:3]), rows[:3], cols)
