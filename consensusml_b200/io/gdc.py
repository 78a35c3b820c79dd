"""GDC acquisition: query the HTSeq count files of the TARGET projects, build the file <-> sample id map and the
download manifest, fetch the files and check them (L0).

Restates JSmith_code/GDC_Data_Download.Rmd: service status check (:56), the ``files()`` query for gene-expression
``HTSeq - Counts`` files of TARGET-AML / NBL / WT / CCSK / OS / RT with the associated-entity fields (:177-186, 473 files),
dropping files that map to two sample ids (:208), the id map written as ``GCD_TARGET_Data_Manifest_AML_NBL_WT_RT.csv``
(:219-229), the download manifest ``TARGET_Manifest_RNAseq_Counts`` (:239-251), the download through the API (:272), the
MD5 check the notebook leaves empty (:280-284), un-gzipping (:295) and the renaming of count-matrix columns from file names
to TARGET barcodes (:321-337).  The reference uses the R package GenomicDataCommons; this is a plain client of the same REST
API (``/status``, ``/files``, ``/data/<uuid>``) on the standard library.

The endpoint is a parameter: the build and test machines have no network, so the tests run the whole flow against a local
stand-in server (tests/test_gdc.py); against ``https://api.gdc.cancer.gov`` the same calls perform the real download."""
from __future__ import annotations

import gzip
import hashlib
import json
import os
import shutil
import urllib.request
from typing import Dict, Iterable, List, Optional, Sequence

import pandas as pd

GDC_API = "https://api.gdc.cancer.gov"
TARGET_PROJECTS = ("TARGET-AML", "TARGET-NBL", "TARGET-WT", "TARGET-CCSK", "TARGET-OS", "TARGET-RT")     # GDC:183-185

# columns of the reference's id-map CSV, in its order (Manifest_Data/GCD_TARGET_Data_Manifest_AML_NBL_WT_RT.csv)
ID_MAP_COLUMNS = ["data_type", "updated_datetime", "file_name", "submitter_id", "file_id", "file_size", "id", "created_datetime",
                  "md5sum", "data_format", "access", "state", "data_category", "type", "experimental_strategy", "project.project_id",
                  "entity_id", "case_id", "entity_submitter_id", "entity_type"]
FILE_FIELDS = ["data_type", "updated_datetime", "file_name", "submitter_id", "file_id", "file_size", "id", "created_datetime", "md5sum",
               "data_format", "access", "state", "data_category", "type", "experimental_strategy", "cases.project.project_id",
               "associated_entities.entity_id", "associated_entities.case_id", "associated_entities.entity_submitter_id",
               "associated_entities.entity_type", "analysis.analysis_type", "analysis.workflow_type", "analysis.workflow_version"]


def _get_json(url: str, payload: Optional[dict] = None, timeout: float = 60.0) -> dict:
    data = json.dumps(payload).encode() if payload is not None else None
    req = urllib.request.Request(url, data=data, headers={"Content-Type": "application/json"} if data else {})
    with urllib.request.urlopen(req, timeout=timeout) as fh:
        return json.loads(fh.read().decode())


def status_ok(endpoint: str = GDC_API) -> bool:
    """``stopifnot(GenomicDataCommons::status()$status == "OK")`` (GDC:56)."""
    return _get_json(endpoint.rstrip("/") + "/status").get("status") == "OK"


def count_files_filter(projects: Sequence[str] = TARGET_PROJECTS, workflow: str = "HTSeq - Counts") -> dict:
    """The filter of the reference's query: ``type == 'gene_expression' & analysis.workflow_type == workflow &
    cases.project.project_id in projects`` in the GDC filter grammar."""
    eq = lambda f, v: {"op": "=", "content": {"field": f, "value": v}}
    return {"op": "and", "content": [eq("type", "gene_expression"), eq("analysis.workflow_type", workflow),
                                     {"op": "in", "content": {"field": "cases.project.project_id", "value": list(projects)}}]}


def query_files(filters: dict, endpoint: str = GDC_API, fields: Sequence[str] = FILE_FIELDS, page_size: int = 500) -> List[dict]:
    """All hits of a ``/files`` query (``results_all()``): pages through ``from`` / ``size`` until ``pagination.total``."""
    return _query_all("files", filters, endpoint, fields=fields, page_size=page_size)


def id_map(hits: Iterable[dict]) -> pd.DataFrame:
    """File <-> sample id map in the layout of the reference's CSV.  A file with other than one associated entity is dropped
    (``ID.map[! sapply(ID.map$associated_entities, nrow) == 2]`` GDC:208 removes the two-sample files; a file with none
    has no barcode to rename its column to)."""
    rows = []
    for h in hits:
        ents = h.get("associated_entities") or []
        if len(ents) != 1:
            continue
        cases = h.get("cases") or [{}]
        row = {c: h.get(c) for c in ID_MAP_COLUMNS[:15]}
        row["project.project_id"] = (cases[0].get("project") or {}).get("project_id")
        for k in ("entity_id", "case_id", "entity_submitter_id", "entity_type"):
            row[k] = ents[0].get(k)
        rows.append(row)
    return pd.DataFrame(rows, columns=ID_MAP_COLUMNS)


def download_manifest(idmap: pd.DataFrame) -> pd.DataFrame:
    """The gdc-client style manifest (``id filename md5 size state``, GDC:239-251) of the files in the id map."""
    return pd.DataFrame({"id": idmap["id"], "filename": idmap["file_name"], "md5": idmap["md5sum"],
                         "size": idmap["file_size"], "state": idmap["state"]})


def _md5(path: str) -> str:
    h = hashlib.md5()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def download(manifest: pd.DataFrame, out_dir: str, endpoint: str = GDC_API, *, gunzip: bool = True, retries: int = 2,
             timeout: float = 120.0) -> Dict[str, str]:
    """Fetch every file of the manifest from ``/data/<id>`` into ``out_dir`` (GDC:272), verify size and MD5 against the
    manifest (the check the notebook leaves as an empty chunk, GDC:280-284), and un-gzip ``*.gz`` (GDC:295).  A file that is
    already present with the right checksum is not fetched again, so an interrupted download resumes.  Returns
    ``{file id: path of the usable file}``; a file that still fails its checksum after ``retries`` re-fetches raises."""
    os.makedirs(out_dir, exist_ok=True)
    out: Dict[str, str] = {}
    for rec in manifest.itertuples(index=False):
        raw = os.path.join(out_dir, rec.filename)
        final = raw[:-3] if (gunzip and raw.endswith(".gz")) else raw
        stamp = final + ".md5"                        # checksum of the archive the final file came from
        if os.path.exists(final) and os.path.exists(stamp) and open(stamp).read().strip() == rec.md5:
            out[rec.id] = final
            continue
        for attempt in range(retries + 1):
            with urllib.request.urlopen(f"{endpoint.rstrip('/')}/data/{rec.id}", timeout=timeout) as src, open(raw, "wb") as dst:
                shutil.copyfileobj(src, dst)
            if _md5(raw) == rec.md5 and (pd.isna(rec.size) or os.path.getsize(raw) == int(rec.size)):
                break
            if attempt == retries:
                os.remove(raw)
                raise IOError(f"{rec.filename} ({rec.id}): checksum mismatch after {retries + 1} downloads")
        if final != raw:
            with gzip.open(raw, "rb") as src, open(final, "wb") as dst:
                shutil.copyfileobj(src, dst)
            os.remove(raw)
        with open(stamp, "w") as fh:
            fh.write(rec.md5)
        out[rec.id] = final
    return out


def barcode_columns(columns: Sequence[str], idmap: pd.DataFrame, *, dots: bool = True) -> List[str]:
    """Rename count-matrix columns from HTSeq file names to TARGET barcodes (GDC:321-337): the ``.gz`` suffix is ignored on
    both sides, a column without an entry keeps its name, and ``dots`` writes the barcode the way R's ``data.frame``
    does (``-`` -> ``.``), which is the form the expression matrices of the repository carry."""
    strip = lambda s: s[:-3] if s.endswith(".gz") else s
    look = {strip(f): b for f, b in zip(idmap["file_name"], idmap["entity_submitter_id"])}
    out = []
    for c in columns:
        b = look.get(strip(str(c)))
        out.append(str(c) if b is None else (b.replace("-", ".") if dots else b))
    return out


def read_htseq(path: str) -> pd.Series:
    """One HTSeq count file (``gene<TAB>count``, no header); the ``__no_feature`` ... summary rows are dropped."""
    s = pd.read_csv(path, sep="\t", header=None, index_col=0).iloc[:, 0]
    return s[~s.index.astype(str).str.startswith("__")]


def assemble_counts(paths: Dict[str, str], idmap: pd.DataFrame) -> pd.DataFrame:
    """Genes x samples raw count matrix from the downloaded files, columns named by barcode (the concatenation the reference
    does with ``catRbind`` / ``cated_RNAseq`` before :321)."""
    by_id = dict(zip(idmap["id"], idmap["file_name"]))
    cols = {by_id[i]: read_htseq(p) for i, p in paths.items()}
    mat = pd.DataFrame(cols)
    mat.columns = barcode_columns(list(mat.columns), idmap)
    return mat.fillna(0).astype("int64")


def acquire(out_dir: str, endpoint: str = GDC_API, projects: Sequence[str] = TARGET_PROJECTS, workflow: str = "HTSeq - Counts"):
    """The whole notebook: status check, query, id map + manifests written next to the data, download, count matrix.
    Returns ``(counts, idmap)``."""
    if not status_ok(endpoint):
        raise ConnectionError(f"GDC at {endpoint} does not report status OK")
    idmap = id_map(query_files(count_files_filter(projects, workflow), endpoint))
    os.makedirs(out_dir, exist_ok=True)
    idmap.to_csv(os.path.join(out_dir, "GDC_Data_Manifest.csv"), index=False)
    man = download_manifest(idmap)
    man.to_csv(os.path.join(out_dir, "Manifest_RNAseq_Counts.txt"), sep="\t", index=False)
    paths = download(man, os.path.join(out_dir, "Expn_Data"), endpoint)
    return assemble_counts(paths, idmap), idmap


# ---- clinical data (GDC:352-430) ---------------------------------------------------------------------------------------

def _query_all(kind: str, filters: dict, endpoint: str, fields: Sequence[str] = (), expand: Sequence[str] = (), page_size: int = 500) -> List[dict]:
    hits: List[dict] = []
    start = 0
    while True:
        body = {"filters": filters, "format": "JSON", "size": page_size, "from": start}
        if fields:
            body["fields"] = ",".join(fields)
        if expand:
            body["expand"] = ",".join(expand)
        res = _get_json(f"{endpoint.rstrip('/')}/{kind}", body)["data"]
        hits.extend(res["hits"])
        total = int(res.get("pagination", {}).get("total", len(hits)))
        start += page_size
        if len(hits) >= total or not res["hits"]:
            return hits


def case_ids(project: str = "TARGET-AML", endpoint: str = GDC_API) -> List[str]:
    """``cases() %>% filter(~ project.project_id == project) %>% ids()`` (GDC:360-366; 988 patients for TARGET-AML)."""
    f = {"op": "=", "content": {"field": "project.project_id", "value": project}}
    return [h["id"] for h in _query_all("cases", f, endpoint, fields=["case_id"])]


def clinical(ids: Sequence[str], endpoint: str = GDC_API) -> pd.DataFrame:
    """``gdc_clinical(case_ids)`` followed by the left joins of GDC:386-391: one row per (case, diagnosis), the case's
    own fields first, then its demographic, exposure and diagnosis records, all keyed by ``case_id``."""
    f = {"op": "in", "content": {"field": "case_id", "value": list(ids)}}
    hits = _query_all("cases", f, endpoint, expand=["demographic", "exposures", "diagnoses"])
    scalar = lambda d: {k: v for k, v in d.items() if not isinstance(v, (list, dict))}
    main = pd.DataFrame([{**scalar(h), "case_id": h.get("case_id", h.get("id"))} for h in hits])
    parts = {"demographic": [], "exposures": [], "diagnoses": []}
    for h in hits:
        cid = h.get("case_id", h.get("id"))
        if isinstance(h.get("demographic"), dict):
            parts["demographic"].append({**scalar(h["demographic"]), "case_id": cid})
        for key in ("exposures", "diagnoses"):
            for rec in h.get(key) or []:
                parts[key].append({**scalar(rec), "case_id": cid})
    full = main
    for key in ("demographic", "exposures", "diagnoses"):
        if parts[key]:
            tab = pd.DataFrame(parts[key])
            full = full.merge(tab, on="case_id", how="left", suffixes=("", f".{key}"))
    return full


def clinical_supplements(out_dir: str, endpoint: str = GDC_API, projects: Sequence[str] = ("TARGET-AML", "TARGET-NBL", "TARGET-WT", "TARGET-RT")) -> Dict[str, str]:
    """The study authors' precompiled clinical spreadsheets (GDC:399-422): query ``type == 'clinical_supplement'`` for the
    four projects, then download.  The files land directly in ``out_dir`` (the notebook has to move them out of the per-UUID
    directories the R client creates, GDC:430-437)."""
    f = {"op": "and", "content": [{"op": "=", "content": {"field": "type", "value": "clinical_supplement"}},
                                  {"op": "in", "content": {"field": "cases.project.project_id", "value": list(projects)}}]}
    hits = _query_all("files", f, endpoint, fields=["id", "file_name", "md5sum", "file_size", "state"])
    man = pd.DataFrame({"id": [h["id"] for h in hits], "filename": [h["file_name"] for h in hits], "md5": [h["md5sum"] for h in hits],
                        "size": [h.get("file_size") for h in hits], "state": [h.get("state") for h in hits]})
    return download(man, out_dir, endpoint, gunzip=False)
