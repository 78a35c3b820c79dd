"""``Cohort`` - the framework's data container.

Plays the role of the reference's ``SummarizedExperiment`` objects (assay matrix + per-gene
``rowData`` + per-sample ``colData`` + metadata; built in
composite_code/rnotebook/seobjects/make_seobj_targetaml.R:165-229 and consumed by every model
wrapper, e.g. cml_targetaml_seanalysis.Rmd:80-88).  Expression lives in a torch tensor laid out
genes x samples (the reference's orientation); row/column annotations are pandas frames.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Sequence

import numpy as np
import pandas as pd
import torch


@dataclass
class Cohort:
    assay: torch.Tensor                 # genes x samples
    gene_ids: Sequence[str]
    sample_ids: Sequence[str]
    row_data: pd.DataFrame = field(default_factory=pd.DataFrame)   # per gene (DE stats, symbols, coordinates)
    col_data: pd.DataFrame = field(default_factory=pd.DataFrame)   # per sample (clinical, split, labels)
    metadata: Dict[str, Any] = field(default_factory=dict)
    assay_name: str = "tmm_log2cpm"

    def __post_init__(self):
        self.gene_ids = list(self.gene_ids)
        self.sample_ids = list(self.sample_ids)
        g, s = self.assay.shape
        if g != len(self.gene_ids) or s != len(self.sample_ids):
            raise ValueError(f"assay is {tuple(self.assay.shape)} but {len(self.gene_ids)} genes / "
                             f"{len(self.sample_ids)} samples were named")
        if len(self.row_data) and len(self.row_data) != g:
            raise ValueError("row_data must have one row per gene")
        if len(self.col_data) and len(self.col_data) != s:
            raise ValueError("col_data must have one row per sample")

    # ---- shape / selection --------------------------------------------------------
    @property
    def n_genes(self) -> int:
        return self.assay.shape[0]

    @property
    def n_samples(self) -> int:
        return self.assay.shape[1]

    def _take(self, gi: Optional[np.ndarray], si: Optional[np.ndarray]) -> "Cohort":
        a = self.assay
        genes, samples, rd, cd = self.gene_ids, self.sample_ids, self.row_data, self.col_data
        if gi is not None:
            gi = np.asarray(gi)
            a = a[torch.as_tensor(gi, device=a.device)]
            genes = [genes[i] for i in gi]
            rd = rd.iloc[gi].reset_index(drop=True) if len(rd) else rd
        if si is not None:
            si = np.asarray(si)
            a = a[:, torch.as_tensor(si, device=a.device)]
            samples = [samples[i] for i in si]
            cd = cd.iloc[si].reset_index(drop=True) if len(cd) else cd
        return Cohort(a.contiguous(), genes, samples, rd, cd, dict(self.metadata), self.assay_name)

    def select_samples(self, mask_or_idx) -> "Cohort":
        idx = np.asarray(mask_or_idx)
        if idx.dtype == bool:
            idx = np.nonzero(idx)[0]
        return self._take(None, idx)

    def select_genes(self, mask_or_idx) -> "Cohort":
        idx = np.asarray(mask_or_idx)
        if idx.dtype == bool:
            idx = np.nonzero(idx)[0]
        elif idx.dtype.kind in "US" or idx.dtype == object:
            pos = {g: i for i, g in enumerate(self.gene_ids)}
            idx = np.array([pos[g] for g in idx])
        return self._take(idx, None)

    def exclude_genes(self, gene_ids) -> "Cohort":
        drop = set(gene_ids)
        return self.select_genes(np.array([g not in drop for g in self.gene_ids]))

    # ---- the views model wrappers want -----------------------------------------------
    def split(self, column: str = "exptset.seahack"):
        """(train, test) cohorts from a ``train``/``test`` flag column
        (the reference's ``exptset.seahack``, make_seobj_targetaml.R:81-82)."""
        flag = self.col_data[column].astype(str).to_numpy()
        return self.select_samples(flag == "train"), self.select_samples(flag == "test")

    def labels(self, column: str = "deg.risk") -> torch.Tensor:
        y = np.array(pd.to_numeric(self.col_data[column]).to_numpy(), copy=True)
        return torch.as_tensor(y, dtype=torch.float32, device=self.assay.device)

    def design(self, dtype=torch.float32) -> torch.Tensor:
        """samples x genes matrix (``t(assay(se))`` in the reference, cml_targetaml_seanalysis.Rmd:82)."""
        return self.assay.t().contiguous().to(dtype)

    def to(self, device) -> "Cohort":
        return Cohort(self.assay.to(device), self.gene_ids, self.sample_ids, self.row_data, self.col_data,
                      dict(self.metadata), self.assay_name)

    def fingerprint(self, label_col: str = "deg.risk", split_col: str = "exptset.seahack") -> str:
        """Content hash of what a model fit depends on: gene and sample ids, the label and split columns and the assay
        values.  Checkpoints are keyed by it, so two cohorts of the same shape never share a cached result."""
        import hashlib
        cached = getattr(self, "_fingerprint", None)
        if cached is not None and cached[0] == (label_col, split_col, self.assay.data_ptr(), self.assay._version):
            return cached[1]                      # same object, assay not written since: skip the device read-back and the hash
        h = hashlib.sha256()
        h.update("\x1f".join(self.gene_ids).encode())
        h.update(b"\x1e")
        h.update("\x1f".join(self.sample_ids).encode())
        for col in (label_col, split_col):
            h.update(b"\x1e")
            if col in self.col_data:
                h.update("\x1f".join(map(str, self.col_data[col].tolist())).encode())
        a = self.assay.detach().to("cpu", torch.float64).contiguous()
        h.update(a.numpy().tobytes())
        fp = h.hexdigest()[:24]
        self._fingerprint = ((label_col, split_col, self.assay.data_ptr(), self.assay._version), fp)
        return fp

    # ---- persistence --------------------------------------------------------------------
    def save(self, path: str) -> None:
        torch.save({"assay": self.assay.cpu(), "gene_ids": self.gene_ids, "sample_ids": self.sample_ids,
                    "row_data": self.row_data.to_dict(orient="list"), "col_data": self.col_data.to_dict(orient="list"),
                    "metadata": self.metadata, "assay_name": self.assay_name}, path)

    @staticmethod
    def load(path: str) -> "Cohort":
        d = torch.load(path, weights_only=False)
        return Cohort(d["assay"], d["gene_ids"], d["sample_ids"], pd.DataFrame(d["row_data"]),
                      pd.DataFrame(d["col_data"]), d["metadata"], d.get("assay_name", "assay"))

    @staticmethod
    def from_summarized_experiment(se) -> "Cohort":
        """Build from a ``RangedSummarizedExperiment`` parsed by :mod:`consensusml_b200.io.rdx`."""
        from ..io.rdx import unwrap
        dat = se["assays"][".xData"][".->data"]
        mats = unwrap(dat["listData"])
        m = np.asarray(unwrap(mats[0]), dtype=np.float64)
        cd = se["colData"]
        sample_ids = list(unwrap(cd["rownames"]))
        ld = cd["listData"]
        cols = {}
        for name, v in zip(ld.names, ld.value):
            vv = unwrap(v)
            lev = v.attrs.get("levels") if hasattr(v, "attrs") else None
            if lev is not None:  # factor -> strings
                lv = list(unwrap(lev))
                vv = [lv[i - 1] if i > 0 else None for i in np.asarray(vv).tolist()]
            cols[name] = list(vv) if not isinstance(vv, list) else vv
        rr = se["rowRanges"]
        gene_ids = list(unwrap(rr["ranges"]["NAMES"]))
        em = rr["elementMetadata"]["listData"]
        rows = {n: list(unwrap(v)) for n, v in zip(em.names, em.value)}
        return Cohort(torch.from_numpy(m), gene_ids, sample_ids, pd.DataFrame(rows), pd.DataFrame(cols),
                      {"source": "SummarizedExperiment"})


class MultiCohort(dict):
    """Named assays over the same samples - the role of the reference's MultiAssayExperiment
    (seobjects/make_seobj_targetaml.R:234-252: gene counts, TMM log-CPM and the DEG subset keyed by experiment name).
    A plain mapping ``name -> Cohort`` that checks sample alignment and subsets all members together."""

    def add(self, name: str, cohort: Cohort) -> "MultiCohort":
        if self and list(next(iter(self.values())).sample_ids) != list(cohort.sample_ids):
            raise ValueError(f"assay '{name}' has different samples than the existing assays")
        self[name] = cohort
        return self

    def select_samples(self, mask_or_idx) -> "MultiCohort":
        out = MultiCohort()
        for k, v in self.items():
            out[k] = v.select_samples(mask_or_idx)
        return out
