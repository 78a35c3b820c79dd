"""GDC acquisition against a local stand-in of the REST API (no network here): query paging, id map layout, two-entity
files dropped, checksum verification with one corrupt first delivery, gunzip, resume, barcode renaming, count assembly."""
import gzip
import hashlib
import json
import threading
from http.server import BaseHTTPRequestHandler, HTTPServer

import pandas as pd
import pytest

from consensusml_b200.io import gdc


def _htseq(seed: int) -> bytes:
    genes = [f"ENSG{i:011d}.1" for i in range(6)]
    body = "".join(f"{g}\t{(seed * 7 + i * 3) % 50}\n" for i, g in enumerate(genes))
    body += "__no_feature\t123\n__ambiguous\t4\n"
    return gzip.compress(body.encode(), mtime=0)


class _FakeGDC:
    def __init__(self, n_files=5, corrupt_first=("f1",)):
        self.files = {}
        self.hits = []
        self.queries = []
        self.served = {}
        self.corrupt_first = set(corrupt_first)
        for i in range(n_files):
            fid = f"f{i}"
            blob = _htseq(i)
            self.files[fid] = blob
            ents = [{"entity_id": f"e{i}", "case_id": f"c{i}", "entity_submitter_id": f"TARGET-20-PA{i:04d}-09A-01R", "entity_type": "aliquot"}]
            if i == 3:                                   # a file that belongs to two samples: must be dropped
                ents.append({"entity_id": "e3b", "case_id": "c3b", "entity_submitter_id": "TARGET-20-PAXXXX-09A-01R", "entity_type": "aliquot"})
            self.hits.append({"data_type": "Gene Expression Quantification", "updated_datetime": "2019-01-01", "file_name": f"{fid}.htseq.counts.gz",
                              "submitter_id": f"s{i}", "file_id": fid, "file_size": len(blob), "id": fid, "created_datetime": "2018-01-01",
                              "md5sum": hashlib.md5(blob).hexdigest(), "data_format": "TXT", "access": "open", "state": "released",
                              "data_category": "Transcriptome Profiling", "type": "gene_expression", "experimental_strategy": "RNA-Seq",
                              "cases": [{"project": {"project_id": "TARGET-AML"}}], "associated_entities": ents,
                              "analysis": {"workflow_type": "HTSeq - Counts"}})

        self.cases = [{"id": f"c{i}", "case_id": f"c{i}", "submitter_id": f"TARGET-20-PA{i:04d}", "primary_site": "Blood",
                       "demographic": {"gender": "female" if i % 2 else "male", "vital_status": "Alive", "demographic_id": f"d{i}"},
                       "exposures": [], "diagnoses": [{"diagnosis_id": f"g{i}", "age_at_diagnosis": 1000 + i, "primary_diagnosis": "AML"}]}
                      for i in range(3)]
        xl = b"PK-not-really-a-spreadsheet"
        self.files["sup0"] = xl
        self.supplements = [{"id": "sup0", "file_name": "TARGET_AML_ClinicalData.xlsx", "md5sum": hashlib.md5(xl).hexdigest(),
                             "file_size": len(xl), "state": "released"}]

    def handler(self):
        fake = self

        class H(BaseHTTPRequestHandler):
            def log_message(self, *a):
                pass

            def _send(self, blob, ctype="application/json"):
                self.send_response(200)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(blob)))
                self.end_headers()
                self.wfile.write(blob)

            def do_GET(self):
                if self.path == "/status":
                    return self._send(json.dumps({"status": "OK", "tag": "fake"}).encode())
                if self.path.startswith("/data/"):
                    fid = self.path.rsplit("/", 1)[1]
                    n = fake.served[fid] = fake.served.get(fid, 0) + 1
                    blob = fake.files[fid]
                    if fid in fake.corrupt_first and n == 1:
                        blob = blob[:-1] + bytes([blob[-1] ^ 0xFF])
                    return self._send(blob, "application/octet-stream")
                self.send_error(404)

            def do_POST(self):
                body = json.loads(self.rfile.read(int(self.headers["Content-Length"])))
                lo, size = body["from"], body["size"]
                if self.path == "/cases":
                    pool = fake.cases
                elif "clinical_supplement" in json.dumps(body["filters"]):
                    pool = fake.supplements
                else:
                    fake.queries.append(body)
                    pool = fake.hits
                if self.path == "/cases" and "expand" not in body:
                    pool = [{"id": c["id"], "case_id": c["case_id"]} for c in pool]
                out = {"data": {"hits": pool[lo:lo + size], "pagination": {"total": len(pool), "from": lo, "size": size}}}
                self._send(json.dumps(out).encode())
        return H


@pytest.fixture()
def fake_gdc():
    fake = _FakeGDC()
    srv = HTTPServer(("127.0.0.1", 0), fake.handler())
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    yield fake, f"http://127.0.0.1:{srv.server_address[1]}"
    srv.shutdown()
    srv.server_close()


def test_filter_matches_reference_query():
    f = gdc.count_files_filter()
    assert f["op"] == "and"
    fields = {c["content"]["field"]: c["content"]["value"] for c in f["content"]}
    assert fields["type"] == "gene_expression"
    assert fields["analysis.workflow_type"] == "HTSeq - Counts"
    assert set(fields["cases.project.project_id"]) == {"TARGET-AML", "TARGET-NBL", "TARGET-WT", "TARGET-CCSK", "TARGET-OS", "TARGET-RT"}


def test_query_pages_and_id_map_layout(fake_gdc):
    fake, url = fake_gdc
    assert gdc.status_ok(url)
    hits = gdc.query_files(gdc.count_files_filter(), url, page_size=2)
    assert len(hits) == 5 and [q["from"] for q in fake.queries] == [0, 2, 4]
    assert "associated_entities.entity_submitter_id" in fake.queries[0]["fields"]
    idmap = gdc.id_map(hits)
    assert list(idmap.columns) == gdc.ID_MAP_COLUMNS
    assert list(idmap["id"]) == ["f0", "f1", "f2", "f4"]             # f3 has two entities
    assert idmap["project.project_id"].eq("TARGET-AML").all()
    man = gdc.download_manifest(idmap)
    assert list(man.columns) == ["id", "filename", "md5", "size", "state"]


def test_reference_manifest_has_the_same_columns():
    ref = pd.read_csv("/root/reference/Manifest_Data/GCD_TARGET_Data_Manifest_AML_NBL_WT_RT.csv", nrows=3) \
        if __import__("os").path.exists("/root/reference/Manifest_Data/GCD_TARGET_Data_Manifest_AML_NBL_WT_RT.csv") else None
    if ref is None:
        pytest.skip("reference checkout not present")
    cols = [c for c in ref.columns if not c.startswith("Unnamed")]
    assert cols == gdc.ID_MAP_COLUMNS


def test_acquire_end_to_end(fake_gdc, tmp_path):
    fake, url = fake_gdc
    counts, idmap = gdc.acquire(str(tmp_path), url)
    assert fake.served == {"f0": 1, "f1": 2, "f2": 1, "f4": 1}        # f1 corrupt once, fetched again; f3 never
    assert list(counts.columns) == [f"TARGET.20.PA{i:04d}.09A.01R" for i in (0, 1, 2, 4)]
    assert counts.shape == (6, 4) and not counts.index.str.startswith("__").any()
    assert counts.iloc[2, 1] == (1 * 7 + 2 * 3) % 50
    assert (tmp_path / "GDC_Data_Manifest.csv").exists() and (tmp_path / "Manifest_RNAseq_Counts.txt").exists()
    assert (tmp_path / "Expn_Data" / "f0.htseq.counts").exists() and not (tmp_path / "Expn_Data" / "f0.htseq.counts.gz").exists()
    # resume: nothing is fetched a second time
    gdc.download(gdc.download_manifest(idmap), str(tmp_path / "Expn_Data"), url)
    assert fake.served == {"f0": 1, "f1": 2, "f2": 1, "f4": 1}


def test_persistent_corruption_raises(tmp_path):
    fake = _FakeGDC(n_files=1, corrupt_first=())
    fake.hits[0]["md5sum"] = "0" * 32
    srv = HTTPServer(("127.0.0.1", 0), fake.handler())
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        url = f"http://127.0.0.1:{srv.server_address[1]}"
        man = gdc.download_manifest(gdc.id_map(fake.hits))
        with pytest.raises(IOError, match="checksum"):
            gdc.download(man, str(tmp_path), url, retries=1)
        assert fake.served["f0"] == 2 and not list(tmp_path.glob("*.gz"))
    finally:
        srv.shutdown()
        srv.server_close()


def test_barcode_columns_keeps_unknown():
    idmap = pd.DataFrame({"file_name": ["a.counts.gz"], "entity_submitter_id": ["TARGET-20-X-1"]})
    assert gdc.barcode_columns(["a.counts", "other"], idmap) == ["TARGET.20.X.1", "other"]
    assert gdc.barcode_columns(["a.counts.gz"], idmap, dots=False) == ["TARGET-20-X-1"]


def test_clinical_tables_and_supplements(fake_gdc, tmp_path):
    fake, url = fake_gdc
    ids = gdc.case_ids("TARGET-AML", url)
    assert ids == ["c0", "c1", "c2"]
    full = gdc.clinical(ids, url)
    assert len(full) == 3 and {"case_id", "submitter_id", "gender", "age_at_diagnosis", "primary_diagnosis"} <= set(full.columns)
    assert list(full["age_at_diagnosis"]) == [1000, 1001, 1002] and list(full["gender"]) == ["male", "female", "male"]
    got = gdc.clinical_supplements(str(tmp_path), url)
    assert list(got) == ["sup0"] and (tmp_path / "TARGET_AML_ClinicalData.xlsx").read_bytes().startswith(b"PK")


def test_cli_acquire(fake_gdc, tmp_path, capsys):
    from consensusml_b200.cli.main import main
    fake, url = fake_gdc
    assert main(["acquire", "--endpoint", url, "--clinical", f"io.out_dir={tmp_path}"]) == 0
    out = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert out["files"] == 4 and out["samples"] == 4 and out["clinical_rows"] == 3 and out["clinical_supplements"] == 1
    back = pd.read_csv(out["csv"], index_col=0)
    assert back.shape == (6, 4)
