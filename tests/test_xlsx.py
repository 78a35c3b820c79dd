"""The workbook reader (io/xlsx.py) on a workbook written here part by part, and - when the reference tree is mounted - on
its clinical spreadsheets: manifest + spreadsheet through api.archives.clinical_frames must reproduce the reference's committed
Clinical_Data/{AML,NBL,WT}_dataframe.csv cell by cell (david_clean.ipynb)."""
import os
import zipfile

import numpy as np
import pandas as pd
import pytest

from consensusml_b200.api import archives as A
from consensusml_b200.io import xlsx as X
from conftest import REFERENCE_DIR, needs_reference


def _write_workbook(path):
    ns = 'xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" xmlns:r="http://schemas.openxmlformats.org/officeDocument/2006/relationships"'
    strings = ["id", "group", "value", "when", "flag", "a", "N/A", "line_x000D_\nbreak", "b"]
    sst = "<sst %s>%s</sst>" % (ns, "".join(f"<si><t>{s}</t></si>" for s in strings))
    styles = f'<styleSheet {ns}><numFmts count="1"><numFmt numFmtId="164" formatCode="yyyy\\-mm\\-dd"/></numFmts><cellXfs count="2"><xf numFmtId="0"/><xf numFmtId="164"/></cellXfs></styleSheet>'
    rows = ['<row r="1"><c r="A1" t="s"><v>0</v></c><c r="B1" t="s"><v>1</v></c><c r="C1" t="s"><v>2</v></c><c r="D1" t="s"><v>3</v></c><c r="E1" t="s"><v>4</v></c></row>',
            '<row r="2"><c r="A2"><v>1</v></c><c r="B2" t="s"><v>5</v></c><c r="C2"><v>2.5</v></c><c r="D2" s="1"><v>43831</v></c><c r="E2" t="b"><v>1</v></c></row>',
            '<row r="3"><c r="A3"><v>2</v></c><c r="B3" t="s"><v>6</v></c><c r="D3" s="1"><v>43832.5</v></c><c r="E3" t="b"><v>0</v></c></row>',
            '<row r="4"><c r="A4"><v>3</v></c><c r="B4" t="s"><v>7</v></c><c r="C4"><v>1e3</v></c><c r="E4" t="inlineStr"><is><t>x</t></is></c></row>']
    sheet = f'<worksheet {ns}><sheetData>{"".join(rows)}</sheetData></worksheet>'
    wb = f'<workbook {ns}><sheets><sheet name="Data" sheetId="1" r:id="rId1"/><sheet name="Other" sheetId="2" r:id="rId2"/></sheets></workbook>'
    rels = ('<Relationships xmlns="http://schemas.openxmlformats.org/package/2006/relationships">'
            '<Relationship Id="rId1" Type="t" Target="worksheets/sheet1.xml"/><Relationship Id="rId2" Type="t" Target="worksheets/sheet2.xml"/></Relationships>')
    with zipfile.ZipFile(path, "w") as z:
        z.writestr("xl/workbook.xml", wb)
        z.writestr("xl/_rels/workbook.xml.rels", rels)
        z.writestr("xl/sharedStrings.xml", sst)
        z.writestr("xl/styles.xml", styles)
        z.writestr("xl/worksheets/sheet1.xml", sheet)
        z.writestr("xl/worksheets/sheet2.xml", f'<worksheet {ns}><sheetData><row r="1"><c r="A1" t="s"><v>8</v></c></row><row r="2"><c r="A2"><v>7</v></c></row></sheetData></worksheet>')


def test_reader_on_a_hand_written_workbook(tmp_path):
    p = str(tmp_path / "w.xlsx")
    _write_workbook(p)
    assert X.sheet_names(p) == ["Data", "Other"]
    df = X.read_xlsx(p)
    assert list(df.columns) == ["id", "group", "value", "when", "flag"] and len(df) == 3
    assert list(df["id"]) == [1, 2, 3] and df["id"].dtype.kind == "i"
    assert df.loc[0, "group"] == "a" and pd.isna(df.loc[1, "group"]) and df.loc[2, "group"] == "line\r\nbreak"
    assert df.loc[0, "value"] == 2.5 and pd.isna(df.loc[1, "value"]) and df.loc[2, "value"] == 1000.0
    assert df.loc[0, "when"] == pd.Timestamp("2020-01-01") and df.loc[1, "when"] == pd.Timestamp("2020-01-02 12:00")
    assert df.loc[0, "flag"] is True or df.loc[0, "flag"] == True                      # noqa: E712
    assert df.loc[2, "flag"] == "x"
    raw = X.read_xlsx(p, keep_default_na=False)
    assert raw.loc[1, "group"] == "N/A"
    other = X.read_xlsx(p, sheet="Other")
    assert list(other.columns) == ["b"] and list(other["b"]) == [7]


@needs_reference
@pytest.mark.parametrize("key,book,table", [("AML", "TARGET_AML_ClinicalData_20160714.xlsx", "AML_dataframe.csv"),
                                            ("NBL", "TARGET_NBL_ClinicalData_20151124.xlsx", "NBL_dataframe.csv"),
                                            ("WT", "TARGET_WT_ClinicalData_Discovery_20160714_public.xlsx", "WT_dataframe.csv")])
def test_clinical_frames_from_the_original_inputs_match_the_reference_tables(key, book, table):
    man = pd.read_csv(os.path.join(REFERENCE_DIR, "Manifest_Data", "GCD_TARGET_Data_Manifest_AML_NBL_WT_RT.csv"))
    clin = X.read_xlsx(os.path.join(REFERENCE_DIR, "Clinical_Data", book))
    got = A.clinical_frames(man, {key: clin})[key].reset_index(drop=True)
    want = pd.read_csv(os.path.join(REFERENCE_DIR, "Clinical_Data", table), index_col=0).reset_index(drop=True)
    assert list(got.columns) == list(want.columns) and got.shape == want.shape
    for c in want.columns:
        for x, y in zip(got[c], want[c]):
            if pd.isna(y):
                assert pd.isna(x), (c, x, y)
            elif isinstance(y, (int, float, np.integer, np.floating)) and not isinstance(y, bool):
                assert abs(float(x) - float(y)) <= 1e-9 * max(1.0, abs(float(y))), (c, x, y)
            else:
                assert str(x) == str(y), (c, x, y)
