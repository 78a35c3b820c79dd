"""Cohort assembly - same entry points as the reference's ``scripts/clean.py`` (lines 1-72).

Behavioural parity, not shared code: the merges below do not mutate their inputs (the reference writes new
columns into slices of the caller's frames, clean.py:16,64) and tolerate missing barcodes."""
from __future__ import annotations

import pandas as pd


def manifest_clinical_merge(manifest_df: pd.DataFrame, clinical_df: pd.DataFrame, target: str) -> pd.DataFrame:
    """Rows of the GDC manifest for one project (e.g. ``'TARGET-AML'``) inner-joined to the clinical sheet on
    ``TARGET USI`` = first 16 characters of ``entity_submitter_id``.  (clean.py:1-18)"""
    sub = manifest_df.loc[manifest_df["project.project_id"] == target].copy()
    sub["TARGET USI"] = sub["entity_submitter_id"].astype(str).str.slice(0, 16)
    return clinical_df.merge(sub, on="TARGET USI")


def assay_transpose(assay_df: pd.DataFrame) -> pd.DataFrame:
    """genes x samples (first column = gene ids) -> samples x genes with the gene ids as header and the
    barcode kept in ``entity_submitter_id``.  (clean.py:21-41)"""
    t = assay_df.T
    t.columns = t.iloc[0]
    t = t.iloc[1:].copy()
    t["entity_submitter_id"] = list(t.index)
    return t


def column_reorder(clinical_df: pd.DataFrame):
    """Column order with the last column moved to position 2.  (clean.py:43-48)"""
    cols = list(clinical_df.columns)
    return [cols[0], cols[-1]] + cols[1:-1]


def assay_clinical_merge(assay_df: pd.DataFrame, clinical_df: pd.DataFrame) -> pd.DataFrame:
    """Left-join clinical+manifest rows to the transposed assay on the barcode (``-`` -> ``.`` as R's
    ``make.names`` does to the assay header) and derive ``Diagnostic ID`` = barcode[-7:-4] (e.g. ``09A``).
    (clean.py:51-72)"""
    clin = clinical_df.copy()
    clin["entity_submitter_id"] = clin["entity_submitter_id"].astype(str).str.replace("-", ".", regex=False)
    df = clin.merge(assay_df, how="left", on="entity_submitter_id")
    df["Diagnostic ID"] = df["entity_submitter_id"].str.slice(-7, -4)
    return df[column_reorder(df)]


def parse_barcode(barcode: str) -> dict:
    """``TARGET.20.PADZCG.09A.02R`` -> project / tissue-source / case / sample-type+vial / portion+analyte."""
    parts = barcode.replace("-", ".").split(".")
    keys = ["program", "project_code", "case", "sample", "portion"]
    out = dict(zip(keys, parts))
    out["usi"] = "-".join(parts[:3])
    out["diagnostic_id"] = parts[3] if len(parts) > 3 else None
    return out
