"""Minimal reader for Excel workbooks (``.xlsx`` = zip of XML parts), enough for the TARGET clinical spreadsheets the reference
loads with ``pd.read_excel`` (scripts/model_walkthrough.ipynb and the archived notebooks: ``TARGET_AML_ClinicalData_20160714.xlsx``,
``TARGET_NBL_...``, ``TARGET_WT_...``) - neither openpyxl nor xlrd is a dependency of this package.

Handled: shared and inline strings, numbers (integers stay integers when a column holds nothing else), booleans, formula
results, date-formatted cells (-> ``datetime64``), empty cells (-> NaN), several sheets.  Not handled: merged cells, rich
formatting, formulas without a cached value."""
from __future__ import annotations

import re
import zipfile
from typing import Dict, List, Optional, Union
from xml.etree import ElementTree as ET

import numpy as np
import pandas as pd

_NS = {"m": "http://schemas.openxmlformats.org/spreadsheetml/2006/main",
       "r": "http://schemas.openxmlformats.org/officeDocument/2006/relationships",
       "pr": "http://schemas.openxmlformats.org/package/2006/relationships"}
_BUILTIN_DATE_FORMATS = set(range(14, 23)) | set(range(45, 48))
# the strings pandas' readers turn into missing values by default
NA_STRINGS = {"", "#N/A", "#N/A N/A", "#NA", "-1.#IND", "-1.#QNAN", "-NaN", "-nan", "1.#IND", "1.#QNAN", "<NA>", "N/A", "NA", "NULL", "NaN",
              "None", "n/a", "nan", "null"}
_ESCAPE = re.compile(r"_x([0-9A-Fa-f]{4})_")


def _text(s: str) -> str:
    """OOXML escapes control characters in strings as ``_xHHHH_`` (``_x000D_`` = carriage return)."""
    return _ESCAPE.sub(lambda m: chr(int(m.group(1), 16)), s) if "_x" in s else s


def _col_index(ref: str) -> int:
    n = 0
    for ch in ref:
        if ch.isalpha():
            n = n * 26 + (ord(ch.upper()) - 64)
        else:
            break
    return n - 1


def sheet_names(path: str) -> List[str]:
    with zipfile.ZipFile(path) as z:
        wb = ET.fromstring(z.read("xl/workbook.xml"))
        return [s.get("name") for s in wb.find("m:sheets", _NS)]


def read_xlsx(path: str, sheet: Union[int, str] = 0, header: bool = True, keep_default_na: bool = True) -> pd.DataFrame:
    """The sheet as a DataFrame (first row = column names), like ``pd.read_excel(path, sheet_name=sheet)``; with
    ``keep_default_na`` the strings pandas treats as missing (``N/A``, ``NA``, ``None``, ...) become NaN as they do there."""
    with zipfile.ZipFile(path) as z:
        wb = ET.fromstring(z.read("xl/workbook.xml"))
        sheets = list(wb.find("m:sheets", _NS))
        entry = sheets[sheet] if isinstance(sheet, int) else next(s for s in sheets if s.get("name") == sheet)
        rid = entry.get(f"{{{_NS['r']}}}id")
        rels = ET.fromstring(z.read("xl/_rels/workbook.xml.rels"))
        target = next(r.get("Target") for r in rels if r.get("Id") == rid)
        target = target.lstrip("/")
        part = target if target.startswith("xl/") else "xl/" + target
        shared: List[str] = []
        if "xl/sharedStrings.xml" in z.namelist():
            for si in ET.fromstring(z.read("xl/sharedStrings.xml")).findall("m:si", _NS):
                shared.append(_text("".join(t.text or "" for t in si.iter(f"{{{_NS['m']}}}t"))))
        date_styles = set()
        if "xl/styles.xml" in z.namelist():
            st = ET.fromstring(z.read("xl/styles.xml"))
            custom = {}
            nf = st.find("m:numFmts", _NS)
            if nf is not None:
                for f in nf:
                    code = re.sub(r'"[^"]*"|\[[^\]]*\]', "", f.get("formatCode") or "")
                    custom[int(f.get("numFmtId"))] = bool(re.search(r"[ymdhs]", code, re.I)) and not re.search(r"[#0]", code.replace("0.", ""))
            xfs = st.find("m:cellXfs", _NS)
            if xfs is not None:
                for i, xf in enumerate(xfs):
                    fid = int(xf.get("numFmtId") or 0)
                    if fid in _BUILTIN_DATE_FORMATS or custom.get(fid, False):
                        date_styles.add(i)
        rows: List[Dict[int, object]] = []
        ws = ET.fromstring(z.read(part))
        for row in ws.find("m:sheetData", _NS):
            cells: Dict[int, object] = {}
            nxt = 0
            for c in row.findall("m:c", _NS):
                ci = _col_index(c.get("r")) if c.get("r") else nxt
                nxt = ci + 1
                t = c.get("t")
                v = c.find("m:v", _NS)
                if t == "inlineStr":
                    val: object = _text("".join(x.text or "" for x in c.iter(f"{{{_NS['m']}}}t")))
                elif v is None or v.text is None:
                    continue
                elif t == "s":
                    val = shared[int(v.text)]
                elif t == "b":
                    val = v.text == "1"
                elif t in ("str", "e"):
                    val = v.text
                else:
                    f = float(v.text)
                    if c.get("s") is not None and int(c.get("s")) in date_styles:
                        val = pd.Timestamp("1899-12-30") + pd.to_timedelta(f, unit="D")
                    else:
                        val = int(f) if f.is_integer() and "." not in v.text and "e" not in v.text.lower() else f
                if keep_default_na and isinstance(val, str) and val in NA_STRINGS and rows:      # (never in the header row)
                    continue
                cells[ci] = val
            rows.append(cells)
    while rows and not rows[-1]:
        rows.pop()
    if not rows:
        return pd.DataFrame()
    width = max(max(r) for r in rows if r) + 1
    grid = [[r.get(j, np.nan) for j in range(width)] for r in rows]
    if not header:
        return pd.DataFrame(grid)
    names = [str(x) if not (isinstance(x, float) and np.isnan(x)) else f"Unnamed: {j}" for j, x in enumerate(grid[0])]
    df = pd.DataFrame(grid[1:], columns=names)
    return df.infer_objects()


def read_table(path: str, **kw) -> pd.DataFrame:
    """A clinical / annotation table from ``.xlsx``, ``.csv`` or tab-separated text, by extension."""
    low = path.lower()
    if low.endswith((".xlsx", ".xlsm")):
        return read_xlsx(path, **kw)
    if low.endswith((".tsv", ".txt")):
        return pd.read_csv(path, sep="\t", **kw)
    return pd.read_csv(path, **kw)
