"""The oracle against every pin we have (no GPU needed): cache.db goldens, the committed outputs
of the reference's own FileParser, and — when /root/reference is present — the reference live."""
import asyncio
import base64
import json
import sys
import types
from pathlib import Path

import pytest

from oracle import file_parser_oracle as O
from oracle import xlsx_reader as X
from tests import corpus

G = Path(__file__).parent / "golden"
REF = Path("/root/reference")


def _load(name):
    return json.loads((G / name).read_text())


def test_cache_row12_xlsx_asset():
    """Config C1: `.cache/42/cache.db` row 12 pins raw_text of tests/assets/attachment.xlsx."""
    parts = _load("attachment_parts.json")
    gold = _load("cache_rows.json")["12"]
    dates, tds = X.date_style_ids(base64.b64decode(parts["styles"]))
    got = O.parse_excel_parts({
        "sheets": [(n, base64.b64decode(s)) for n, s in zip(parts["sheet_names"], parts["sheets"])],
        "sst": base64.b64decode(parts["sst"]), "date_styles": dates, "td_styles": tds})
    assert got == gold
    assert len(got) == 696


def row14_table():
    """Reconstruction of the 23-row sheet behind cache row 14."""
    rows = [["Part Number", "Description", "Quantity", "Target Price (INR)"]]
    for i in range(23):
        rows.append([f"MP-{1000 + i}", f"Dummy Part {i + 1} - Industrial component",
                     100 * (i + 1), 20 + 5.75 * i])
    return rows


def test_cache_row14_float_formatting():
    gold = _load("cache_rows.json")["14"]
    df = X.table_to_frame(row14_table(), 1000)
    assert O.sheets_to_text([("Sheet1", df)], 1000) == gold


def test_csv_goldens():
    for case in _load("ref_csv.json"):
        data = base64.b64decode(case["data"])
        try:
            got = {"text": O.parse_csv_bytes(data, case["nrows"])}
        except O.OracleParsingError as e:
            got = {"error": str(e)}
        want = {k: case[k] for k in ("text", "error") if k in case}
        assert got == want, case["name"]


def test_txt_goldens():
    for case in _load("ref_txt.json"):
        assert O.parse_txt_bytes(base64.b64decode(case["data"])) == case["text"], case["name"]


def test_golden_inputs_are_reproducible():
    """The committed fixtures are exactly what tests/corpus.py generates (seeded)."""
    want = {c["name"]: c["data"] for c in _load("ref_csv.json")}
    for name, data, _ in corpus.csv_golden_cases():
        assert base64.b64encode(data).decode() == want[name], name


@pytest.mark.skipif(not (REF / "app/file_parser.py").exists(), reason="reference not mounted")
def test_live_reference_differential(tmp_path):
    sys.path.insert(0, str(REF))
    sys.modules.setdefault("docx", types.ModuleType("docx"))
    import logging
    logging.disable(logging.CRITICAL)
    from app.file_parser import FileParser, FileParsingError
    fp = FileParser(max_file_size_mb=4096, max_excel_rows=1000)
    for seed in range(1000, 1060):
        data = corpus.fuzz_csv(seed)
        p = tmp_path / "x.csv"
        p.write_bytes(data)
        try:
            want = fp._parse_csv_sync(p)
        except FileParsingError as e:
            want = "ERR " + str(e)
        try:
            got = O.parse_csv_bytes(data, 1000)
        except O.OracleParsingError as e:
            got = "ERR " + str(e)
        assert got == want, seed
        t = corpus.text_body(seed)
        p2 = tmp_path / "x.txt"
        p2.write_bytes(t)
        assert O.parse_txt_bytes(t) == asyncio.run(fp._parse_txt_async(p2)), seed
    logging.disable(logging.NOTSET)


def test_host_date_styles_match_oracle():
    """The product's host-side styles.xml reader (date / timedelta number formats, date1904) agrees
    with the oracle's restatement of openpyxl's is_date_format / is_timedelta_format."""
    from replisense_rfq_b200 import xlsx_host as Hh
    from oracle import xlsx_reader as X
    codes = ["General", "0.00", "yyyy\\-mm\\-dd", "dd/mm/yyyy;@", "[h]:mm:ss", "[mm]:ss.0", "[$-409]d\\-mmm\\-yy;@",
             '#,##0.00\\ "days"', "[Red]0.00", "h:mm AM/PM", '"day" 0', "0.0E+00", "mmm\\ yy", "[ss].00",
             "_-* #,##0.00\\ _d_-", "\\d0", "@", "[hh]:mm", "[Blue]yy"]
    fmts = "".join('<numFmt numFmtId="%d" formatCode="%s"/>' % (170 + i, c.replace("&", "&amp;").replace('"', "&quot;"))
                   for i, c in enumerate(codes))
    xfs = "".join('<xf numFmtId="%d"/>' % k for k in list(range(0, 50)) + [170 + i for i in range(len(codes))])
    styles = ('<styleSheet xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main"><numFmts>%s</numFmts>'
              '<cellXfs>%s</cellXfs></styleSheet>' % (fmts, xfs)).encode()
    dates, tds = X.date_style_ids(styles)
    got = Hh._date_style_ids(styles)
    assert {i & 0x7FFFFFFF for i in got} == dates and len(dates) > 12
    assert {i & 0x7FFFFFFF for i in got if i & 0x80000000} == tds and len(tds) >= 4
