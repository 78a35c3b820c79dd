"""Writers for the reference's gene-list outputs.

``write_lasso_gene_csv`` reproduces ``TARGET_AML_Logistic_Lasso_withDEGs_RiskGroup_Associated_Genes.csv``
(written at JSmith_code/Differential_Expression_and_Lasso.Rmd:449-460): non-zero coefficients sorted descending,
version-stripped Ensembl ids, BioMart annotation columns, R ``write.csv`` quoting and 15-significant-digit
numbers, row names "1".."k".  ``write_gene_list`` reproduces the ``"","x"`` single-column lists
(BuieRProj/output/*.csv, VikasP/DEG_*.csv, TARGET_AML_*_Samples.csv)."""
from __future__ import annotations

from typing import Dict, Iterable, Optional

import pandas as pd

from ..io.annotation import annotate, strip_version


def r_num(x) -> str:
    """R's ``write.csv`` rendering of a double (15 significant digits, integers without a decimal point)."""
    if x is None or (isinstance(x, float) and x != x):
        return "NA"
    if isinstance(x, (int,)) or (isinstance(x, float) and x == int(x) and abs(x) < 1e15):
        return str(int(x))
    return "%.15g" % float(x)


def _q(s) -> str:
    return '"' + str(s).replace('"', '""') + '"'


def lasso_gene_frame(nonzero_coef: Dict[str, float], coef_name: str = "Mod.RG.nonzero.coef") -> pd.DataFrame:
    items = sorted(nonzero_coef.items(), key=lambda kv: -kv[1])
    ann = annotate([g for g, _ in items])
    ann.insert(1, coef_name, [v for _, v in items])
    return ann


def write_lasso_gene_csv(nonzero_coef: Dict[str, float], path: str, coef_name: str = "Mod.RG.nonzero.coef") -> pd.DataFrame:
    df = lasso_gene_frame(nonzero_coef, coef_name)
    with open(path, "w", newline="") as fh:
        fh.write(",".join([_q("")] + [_q(c) for c in df.columns]) + "\n")
        for i, row in enumerate(df.itertuples(index=False), 1):
            cells = [_q(i)]
            for c, v in zip(df.columns, row):
                if c in ("Gene_ID", "Gene.name", "Karyotype.band"):
                    cells.append("NA" if v is None or v != v else _q(v))
                elif c == "Chromosome.scaffold.name":
                    sv = str(v)
                    cells.append(sv if sv.isdigit() else ("NA" if v is None or v != v else _q(sv)))
                else:
                    cells.append(r_num(float(v)) if v is not None and v == v else "NA")
            fh.write(",".join(cells) + "\n")
    return df


def write_gene_list(genes: Iterable[str], path: str, strip: bool = False) -> None:
    with open(path, "w", newline="") as fh:
        fh.write('"","x"\n')
        for i, g in enumerate(genes, 1):
            fh.write(f'"{i}",{_q(strip_version(g) if strip else g)}\n')


def read_gene_list(path: str):
    return pd.read_csv(path).iloc[:, 1].astype(str).tolist()
