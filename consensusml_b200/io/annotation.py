"""Gene annotation joins (symbol, chromosome, start, end, karyotype band).

The reference joins its lasso coefficients with a BioMart export
(JSmith_code/Differential_Expression_and_Lasso.Rmd:449-456) and annotates the SummarizedExperiment rows from
Ensembl v95 via biomaRt (seobjects/make_seobj_targetaml.R:103-117).  Neither service is reachable offline, so
both tables ship as fixtures: the full v95 table (``dfens``) and the 13-row BioMart export with karyotype bands."""
from __future__ import annotations

import re
from typing import Iterable

import pandas as pd

from .. import datasets


def strip_version(gene_id: str) -> str:
    """``ENSG00000078399.14`` -> ``ENSG00000078399`` (``gsub("\\\\.[0-9]{1,}", "", .)``, DEL:453)."""
    return re.sub(r"\.[0-9]+", "", gene_id)


def biomart13() -> pd.DataFrame:
    return pd.read_csv(datasets.path("lasso13_gene_symbols_biomart.csv"))


def annotate(gene_ids: Iterable[str]) -> pd.DataFrame:
    """One row per input gene (order kept) with ``Gene.name, Chromosome.scaffold.name, Gene.start..bp.,
    Gene.end..bp., Karyotype.band`` - the column names R gives the BioMart header."""
    ids = [strip_version(g) for g in gene_ids]
    out = pd.DataFrame({"Gene_ID": ids})
    bm = biomart13().rename(columns={"Gene stable ID": "Gene_ID", "Gene name": "Gene.name",
                                     "Chromosome/scaffold name": "Chromosome.scaffold.name",
                                     "Gene start (bp)": "Gene.start..bp.", "Gene end (bp)": "Gene.end..bp.",
                                     "Karyotype band": "Karyotype.band"})
    ens = datasets.gene_annotation()
    cols = {c.lower(): c for c in ens.columns}
    idc = cols.get("ensembl_gene_id")
    ens2 = pd.DataFrame({
        "Gene_ID": ens[idc].astype(str),
        "Gene.name": ens[cols.get("hgnc_symbol", idc)],
        "Chromosome.scaffold.name": ens[cols.get("seqnames", idc)],
        "Gene.start..bp.": ens[cols.get("start", idc)],
        "Gene.end..bp.": ens[cols.get("end", idc)],
    }).drop_duplicates("Gene_ID")
    ens2["Karyotype.band"] = None
    merged = out.merge(bm, on="Gene_ID", how="left")
    missing = merged["Gene.name"].isna()
    if missing.any():
        fill = out[missing.values].merge(ens2, on="Gene_ID", how="left")
        for c in ["Gene.name", "Chromosome.scaffold.name", "Gene.start..bp.", "Gene.end..bp.", "Karyotype.band"]:
            merged[c] = merged[c].astype(object)             # mixed str / number / missing: keep pandas from coercing
            merged.loc[missing.values, c] = fill[c].astype(object).values
    return merged
