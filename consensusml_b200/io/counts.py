"""Count-matrix ingestion: per-sample file concatenation and reference-format CSV readers.

``concat_counts`` replaces ``catExpnData(filenames, regex, cols, header, removeFirstLine, sep)``
(JSmith_code/Cat_Expn_Data.r:7-49; used at GDC_Data_Download.Rmd:306-349 to build the 60,488 x 474 matrix): a C++
mmap parser (csrc/host/count_files.cpp) fills the genes x samples int32 matrix with one thread per file.
``read_counts_csv`` reads the concatenated CSV layout the reference then writes (genes as rows, first column = id)."""
from __future__ import annotations

import ctypes
import os
import re
from typing import List, Optional, Sequence, Tuple

import numpy as np
import pandas as pd
import torch

from .. import _native


def concat_counts(filenames: Sequence[str], regex: Optional[str] = None, id_col: int = 0, value_col: int = 1, header: bool = False,
                  remove_first_line: bool = False, sep: str = "\t", drop_summary_rows: bool = True,
                  threads: Optional[int] = None) -> Tuple[torch.Tensor, List[str], List[str]]:
    """Returns (counts int32 [genes, samples], gene ids, sample ids).  ``regex`` with one capture group extracts the
    sample id from the file name (the reference's ``gsub(regex, "\\\\1", filenames)``).  HTSeq's ``__no_feature`` style
    summary rows are dropped (VikasP/AML.R:151 removes rows starting with ``__``)."""
    lib = _native.host()
    lib.cml_host_counts_scan.restype = ctypes.c_long
    lib.cml_host_counts_scan.argtypes = [ctypes.c_char_p, ctypes.c_char, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    lib.cml_host_counts_id.argtypes = [ctypes.c_long, ctypes.c_char_p, ctypes.c_int]
    lib.cml_host_counts_concat.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.c_char, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_int, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_int]
    files = [os.fspath(f) for f in filenames]
    if not files:
        raise ValueError("no files")
    skip = int(bool(header)) + int(bool(remove_first_line))
    n_genes = lib.cml_host_counts_scan(files[0].encode(), sep.encode(), skip, id_col, value_col)
    if n_genes <= 0:
        raise ValueError(f"{files[0]}: could not parse (code {n_genes})")
    buf = ctypes.create_string_buffer(256)
    genes = []
    for i in range(n_genes):
        lib.cml_host_counts_id(i, buf, 256)
        genes.append(buf.value.decode())
    out = np.zeros((n_genes, len(files)), dtype=np.int32)
    status = np.zeros(len(files), dtype=np.int64)
    arr = (ctypes.c_char_p * len(files))(*[f.encode() for f in files])
    bad = lib.cml_host_counts_concat(arr, len(files), sep.encode(), skip, id_col, value_col, out.ctypes.data, n_genes,
                                     status.ctypes.data, threads or min(16, os.cpu_count() or 1))
    if bad:
        which = [(files[i], int(status[i])) for i in np.flatnonzero(status)]
        raise ValueError(f"{bad} count file(s) do not match the gene list of {files[0]}: {which[:5]}")
    if regex is not None:
        samples = [re.sub(regex, r"\1", os.path.basename(f)) for f in files]
    else:
        samples = [os.path.basename(f) for f in files]
    if drop_summary_rows:
        keep = np.array([not g.startswith("__") for g in genes])
        out, genes = out[keep], [g for g, k in zip(genes, keep) if k]
    return torch.from_numpy(np.ascontiguousarray(out)), genes, samples


def _read_counts_csv_native(path: str, sep: str = ",", threads: Optional[int] = None):
    """mmap + thread-pool parse of an integer matrix CSV (csrc/host/csv_matrix.cpp).  Returns None when the file is not a
    plain integer matrix (the caller then uses pandas)."""
    lib = _native.host()
    lib.cml_host_csv_open.argtypes = [ctypes.c_char_p, ctypes.c_char, ctypes.POINTER(ctypes.c_long), ctypes.POINTER(ctypes.c_long)]
    lib.cml_host_csv_header.argtypes = [ctypes.c_long, ctypes.c_char_p, ctypes.c_int]
    lib.cml_host_csv_read.argtypes = [ctypes.c_char, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    lib.cml_host_csv_read.restype = ctypes.c_long
    lib.cml_host_csv_close.restype = None
    rows, cols = ctypes.c_long(0), ctypes.c_long(0)
    if lib.cml_host_csv_open(os.fspath(path).encode(), sep.encode(), ctypes.byref(rows), ctypes.byref(cols)) != 0:
        return None
    try:
        if rows.value <= 0 or cols.value <= 0:
            return None
        buf = ctypes.create_string_buffer(4096)
        samples = []
        for j in range(cols.value):
            if lib.cml_host_csv_header(j, buf, 4096) < 0:
                return None
            samples.append(buf.value.decode())
        out = np.empty((rows.value, cols.value), dtype=np.int32)
        id_cap = 64
        ids = np.zeros((rows.value, id_cap), dtype=np.uint8)
        rc = lib.cml_host_csv_read(sep.encode(), out.ctypes.data, ids.ctypes.data, id_cap, threads or min(16, os.cpu_count() or 1))
        if rc != 0:
            return None
        genes = [bytes(r).split(b"\0", 1)[0].decode() for r in ids]
        return torch.from_numpy(out), genes, samples
    finally:
        lib.cml_host_csv_close()


def read_counts_csv(path: str, id_column: int = 0, native: bool = True) -> Tuple[torch.Tensor, List[str], List[str]]:
    """Genes x samples CSV (first column gene id, header = sample barcodes), e.g. the reference's
    ``TARGET_NBL_AML_RT_WT_HTSeq_Counts.csv`` (read at Differential_Expression_and_Lasso.Rmd:91).  Plain integer matrices
    go through the parallel native parser (about 20x pandas at 60,488 x 474); anything else falls back to pandas."""
    if native and id_column == 0:
        got = _read_counts_csv_native(path)
        if got is not None:
            return got
    df = pd.read_csv(path, index_col=id_column)
    return torch.from_numpy(df.to_numpy(dtype=np.int32)), [str(i) for i in df.index], [str(c) for c in df.columns]


def write_matrix_csv(path: str, matrix: torch.Tensor, row_ids: Sequence[str], col_ids: Sequence[str], native: bool = True,
                     threads: Optional[int] = None) -> None:
    """Matrix with row and column ids -> CSV (empty first header cell, like ``write.csv`` / ``DataFrame.to_csv``).  int32 /
    int64-that-fit and floating matrices go through the parallel native writer (csrc/host/csv_matrix.cpp: doubles in their
    shortest round-trip form, so the file reads back to exactly the same values); ids containing the separator, quotes or
    line breaks, and other dtypes, fall back to pandas."""
    m = matrix.detach().cpu()
    plain = all(not any(ch in s for ch in ',"\n\r') for s in list(row_ids) + list(col_ids))
    if native and plain and m.dim() == 2 and m.shape[0] == len(row_ids) and m.shape[1] == len(col_ids) and m.numel() > 0:
        lib = _native.host()
        arr = None
        if m.dtype in (torch.float64, torch.float32, torch.bfloat16, torch.float16):
            arr, fn = np.ascontiguousarray(m.to(torch.float64).numpy()), lib.cml_host_csv_write_f64
        elif m.dtype in (torch.int32, torch.int16, torch.uint8, torch.int8) or (m.dtype == torch.int64 and int(m.abs().max()) < 2 ** 31):
            arr, fn = np.ascontiguousarray(m.to(torch.int32).numpy()), lib.cml_host_csv_write_i32
        if arr is not None:
            fn.restype = ctypes.c_long
            fn.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_long, ctypes.POINTER(ctypes.c_char_p),
                           ctypes.POINTER(ctypes.c_char_p), ctypes.c_char, ctypes.c_int]
            rid = (ctypes.c_char_p * len(row_ids))(*[str(s).encode() for s in row_ids])
            cid = (ctypes.c_char_p * len(col_ids))(*[str(s).encode() for s in col_ids])
            rc = fn(os.fspath(path).encode(), arr.ctypes.data, arr.shape[0], arr.shape[1], rid, cid, b",",
                    threads or min(16, os.cpu_count() or 1))
            if rc != 0:
                raise OSError(f"could not write {path} (code {rc})")
            return
    pd.DataFrame(m.numpy() if m.dtype != torch.bfloat16 else m.float().numpy(), index=list(row_ids), columns=list(col_ids)).to_csv(path)


def read_manifest(path: str) -> pd.DataFrame:
    """GDC manifest / id map (Manifest_Data/GCD_TARGET_Data_Manifest_AML_NBL_WT_RT.csv; GDC_Data_Download.Rmd:219-229)."""
    return pd.read_csv(path)
