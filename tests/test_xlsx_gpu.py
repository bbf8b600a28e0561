"""GPU parity: sheet XML + sharedStrings path vs the cache.db goldens and the oracle."""
import base64
import io
import json
import warnings
import zipfile
from pathlib import Path

import pytest

from oracle import file_parser_oracle as O
from oracle import xlsx_reader as X
from tests import corpus

pytestmark = pytest.mark.gpu
G = Path(__file__).parent / "golden"


@pytest.fixture(scope="module")
def eng():
    from replisense_rfq_b200 import Engine, build
    build.build()
    return Engine(0)


def _oracle(sx, ss, nrows=1000):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        df = X.table_to_frame(X.sheet_table(sx, X.read_string_table(ss), nrows=nrows), nrows)
        return None if df.empty else O.frame_to_text(df)


def _parts():
    p = json.loads((G / "attachment_parts.json").read_text())
    return {k: ([base64.b64decode(x) for x in v] if isinstance(v, list) and k == "sheets" else v)
            for k, v in p.items()}


def test_c1_golden_sheet(eng):
    """Config C1: the reference's own xlsx asset, pinned by .cache/42/cache.db row 12."""
    p = _parts()
    gold = json.loads((G / "cache_rows.json").read_text())["12"]
    r = eng.sheets_batch([base64.b64decode(p["sst"]), p["sheets"][0]], [1, 0], [-1, 0], 1000)
    assert int(r.status[1]) == 0
    assert "=== Sheet: Sheet1 ===\n" + r.doc_str(1) == gold


def test_c1_through_file_parser(eng, tmp_path):
    """The same asset as an .xlsx file through FileParser.parse_file_async (zip rebuilt from the
    committed parts)."""
    import asyncio
    from replisense_rfq_b200 import FileParser
    p = _parts()
    path = tmp_path / "email_4 attachment.xlsx"
    with zipfile.ZipFile(path, "w", zipfile.ZIP_DEFLATED) as z:
        z.writestr("xl/workbook.xml", base64.b64decode(p["workbook"]))
        z.writestr("xl/_rels/workbook.xml.rels", base64.b64decode(p["rels"]))
        z.writestr("xl/styles.xml", base64.b64decode(p["styles"]))
        z.writestr("xl/sharedStrings.xml", base64.b64decode(p["sst"]))
        z.writestr("xl/worksheets/sheet1.xml", p["sheets"][0])
    res = asyncio.run(FileParser(engine=eng).parse_file_async(str(path)))
    assert res["raw_text"] == json.loads((G / "cache_rows.json").read_text())["12"]
    assert res["parsing_method"] == "async_xlsx" and res["source_file"] == "email_4 attachment.xlsx"
    assert res["raw_text"] == O.parse_excel_path(path, 1000)


def test_row14_golden(eng):
    """cache.db row 14: 23-row sheet with a float column (20.00 ... 146.50)."""
    gold = json.loads((G / "cache_rows.json").read_text())["14"]
    strings, rows = [], []

    def sid(s):
        strings.append(s)
        return len(strings) - 1
    hdr = ["Part Number", "Description", "Quantity", "Target Price (INR)"]
    rows.append("".join('<c r="%s1" t="s"><v>%d</v></c>' % (corpus.col_letters(c), sid(h)) for c, h in enumerate(hdr)))
    for i in range(23):
        r = i + 2
        rows.append('<c r="A%d" t="s"><v>%d</v></c><c r="B%d" t="s"><v>%d</v></c><c r="C%d"><v>%d</v></c>'
                    '<c r="D%d"><v>%s</v></c>' % (r, sid(f"MP-{1000 + i}"), r,
                                                   sid(f"Dummy Part {i + 1} - Industrial component"), r,
                                                   100 * (i + 1), r, repr(20 + 5.75 * i)))
    sx = ('<worksheet xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main"><sheetData>' +
          "".join('<row r="%d">%s</row>' % (i + 1, x) for i, x in enumerate(rows)) + "</sheetData></worksheet>").encode()
    ss = ('<sst xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main">' +
          "".join("<si><t>%s</t></si>" % corpus.xml_escape(s) for s in strings) + "</sst>").encode()
    r = eng.sheets_batch([ss, sx], [1, 0], [-1, 0], 1000)
    assert int(r.status[1]) == 0 and "=== Sheet: Sheet1 ===\n" + r.doc_str(1) == gold


def _batch(eng, pairs, nrows=1000):
    docs, kind, sst_of = [], [], []
    for sx, ss in pairs:
        docs += [ss, sx]; kind += [1, 0]; sst_of += [-1, len(docs) - 2]
    r = eng.sheets_batch(docs, kind, sst_of, nrows)
    out = []
    for i in range(len(pairs)):
        st = int(r.status[2 * i + 1])
        out.append((st, r.doc_str(2 * i + 1) if st == 0 else None))
    return out


def test_sheet_corpus_c3_style(eng):
    pairs = [corpus.sheet_xml_doc(s, 5, 300) for s in range(120)]
    got = _batch(eng, pairs)
    bad = [i for i, ((st, g), (sx, ss)) in enumerate(zip(got, pairs)) if (None if st == 3 else g) != _oracle(sx, ss) or st == 2]
    assert not bad, bad[:10]


def test_sheet_fuzz(eng):
    pairs = [corpus.sheet_xml_fuzz(s) for s in range(400)]
    got = _batch(eng, pairs)
    bad, unsup = [], 0
    for i, ((st, g), (sx, ss)) in enumerate(zip(got, pairs)):
        if st == 2:
            unsup += 1
            continue
        if (None if st == 3 else g) != _oracle(sx, ss):
            bad.append(i)
    assert not bad, bad[:10]
    assert unsup <= 0.02 * len(pairs)


def test_sheet_nrows_cap(eng):
    sx, ss = corpus.sheet_xml_doc(7, 200, 300)
    for cap in (1, 10, 57):
        (st, g), = _batch(eng, [(sx, ss)], cap)
        assert (None if st == 3 else g) == _oracle(sx, ss, cap), cap


def _oracle_dates(sx, ss, ids, tds, d1904, nrows=1000):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        df = X.table_to_frame(X.sheet_table(sx, X.read_string_table(ss), ids, tds, d1904, nrows), nrows)
        return None if df.empty else O.frame_to_text(df)


def test_sheet_dates(eng):
    """Date-styled numbers: datetime64 columns (date only / seconds / milliseconds, NaT), datetimes
    and times inside object columns, the 1900 leap-year shift, the 1904 epoch, out-of-range serials."""
    cases = [corpus.sheet_xml_dates(s) for s in range(500)]
    docs, kind, sst_of, ids = [], [], [], []
    for sx, ss, dates, tds, d1904 in cases:
        docs += [ss, sx]; kind += [1, 2 if d1904 else 0]; sst_of += [-1, len(docs) - 2]
        ids += [(), [i | (0x80000000 if i in tds else 0) for i in dates]]
    r = eng.sheets_batch(docs, kind, sst_of, 1000, ids)
    bad, unsup = [], 0
    for i, (sx, ss, dates, tds, d1904) in enumerate(cases):
        st = int(r.status[2 * i + 1])
        if st == 2:
            unsup += 1
            continue
        got = None if st == 3 else r.doc_str(2 * i + 1)
        if got != _oracle_dates(sx, ss, dates, tds, d1904):
            bad.append(i)
    assert not bad, bad[:10]
    assert unsup <= 0.02 * len(cases)


def test_sheet_timedeltas(eng):
    """Timedelta-styled numbers ([h]:mm:ss): openpyxl's from_excel(value, timedelta=True) over CPython's
    timedelta arithmetic; a timedelta64 column prints "N days" / "N days HH:MM:SS[.ffffff]" with NaT for
    blanks, an object column and a label print str(timedelta); more than 999 999 999 days is an error cell."""
    cases = [corpus.sheet_xml_dates(s, p_td=0.6) for s in range(400)]
    docs, kind, sst_of, ids = [], [], [], []
    for sx, ss, dates, tds, d1904 in cases:
        docs += [ss, sx]; kind += [1, 2 if d1904 else 0]; sst_of += [-1, len(docs) - 2]
        ids += [(), [i | (0x80000000 if i in tds else 0) for i in dates]]
    r = eng.sheets_batch(docs, kind, sst_of, 1000, ids)
    bad, unsup, n_td = [], 0, 0
    for i, (sx, ss, dates, tds, d1904) in enumerate(cases):
        st = int(r.status[2 * i + 1])
        if st == 2:
            unsup += 1
            continue
        got = None if st == 3 else r.doc_str(2 * i + 1)
        want = _oracle_dates(sx, ss, dates, tds, d1904)
        n_td += bool(want) and " days" in want
        if got != want:
            bad.append(i)
    assert not bad, bad[:10]
    assert unsup <= 0.02 * len(cases) and n_td > 15


_STYLES = ('<?xml version="1.0" encoding="UTF-8"?><styleSheet xmlns="http://schemas.openxmlformats.org/'
           'spreadsheetml/2006/main"><numFmts count="2"><numFmt numFmtId="164" formatCode="yyyy\\-mm\\-dd"/>'
           '<numFmt numFmtId="165" formatCode="#,##0.00\\ &quot;days&quot;"/></numFmts><cellXfs count="5">'
           '<xf numFmtId="0"/><xf numFmtId="165"/><xf numFmtId="2"/><xf numFmtId="164"/><xf numFmtId="22"/>'
           '</cellXfs></styleSheet>')


def test_xlsx_with_dates_through_file_parser(eng, tmp_path):
    """A workbook whose styles.xml declares a custom date format (cellXfs 3), a builtin one
    (cellXfs 4 = numFmtId 22) and a non-date custom format with a quoted 'days' (cellXfs 1)."""
    import asyncio
    from replisense_rfq_b200 import FileParser
    ss = ('<sst xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main"><si><t>Part</t></si>'
          '<si><t>Delivery</t></si><si><t>Stamp</t></si><si><t>Lead</t></si><si><t>RJF544</t></si></sst>')
    rows = ['<row r="1"><c r="A1" t="s"><v>0</v></c><c r="B1" t="s"><v>1</v></c><c r="C1" t="s"><v>2</v></c>'
            '<c r="D1" t="s"><v>3</v></c></row>']
    for i in range(6):
        r = i + 2
        rows.append('<row r="%d"><c r="A%d" t="s"><v>4</v></c><c r="B%d" s="3"><v>%d</v></c>'
                    '<c r="C%d" s="4"><v>%s</v></c><c r="D%d" s="1"><v>%d</v></c></row>'
                    % (r, r, r, 45900 + 7 * i, r, repr(45900 + i / 7), r, 10 + i))
    sx = ('<worksheet xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main"><sheetData>'
          + "".join(rows) + "</sheetData></worksheet>")
    wb = ('<workbook xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" xmlns:r="http://'
          'schemas.openxmlformats.org/officeDocument/2006/relationships"><sheets><sheet name="RFQ" '
          'sheetId="1" r:id="rId1"/></sheets></workbook>')
    rels = ('<Relationships xmlns="http://schemas.openxmlformats.org/package/2006/relationships">'
            '<Relationship Id="rId1" Type="http://schemas.openxmlformats.org/officeDocument/2006/'
            'relationships/worksheet" Target="worksheets/sheet1.xml"/></Relationships>')
    path = tmp_path / "dates.xlsx"
    with zipfile.ZipFile(path, "w", zipfile.ZIP_DEFLATED) as z:
        z.writestr("xl/workbook.xml", wb)
        z.writestr("xl/_rels/workbook.xml.rels", rels)
        z.writestr("xl/styles.xml", _STYLES)
        z.writestr("xl/sharedStrings.xml", ss)
        z.writestr("xl/worksheets/sheet1.xml", sx)
    want = O.parse_excel_path(path, 1000)
    assert "2025-08-31" in want and "Lead" in want
    res = asyncio.run(FileParser(engine=eng).parse_file_async(str(path)))
    assert res["raw_text"] == want


def test_sheet_without_refs(eng):
    """Writers that omit r= on <row> and/or <c>: rows and columns are numbered by position
    (openpyxl worksheet/_reader.py row_counter / col_counter)."""
    pairs = []
    for seed in range(300):
        sx, ss = corpus.sheet_xml_fuzz(seed) if seed % 2 else corpus.sheet_xml_doc(seed, 5, 120)
        pairs.append((corpus.strip_refs(sx, rows=seed % 3 != 1, cells=seed % 3 != 2), ss))
    got = _batch(eng, pairs)
    bad, unsup = [], 0
    for i, ((st, g), (sx, ss)) in enumerate(zip(got, pairs)):
        if st == 2:
            unsup += 1
            continue
        if (None if st == 3 else g) != _oracle(sx, ss):
            bad.append(i)
    assert not bad, bad[:10]
    assert unsup <= 0.02 * len(pairs)


def test_sheet_with_partial_refs(eng):
    """Writers that state r= only on SOME rows / cells: an element without it continues the count of the
    element before it (openpyxl worksheet/_reader.py parse_row / parse_cell: row_counter, col_counter)."""
    pairs = []
    for seed in range(300):
        sx, ss = corpus.sheet_xml_fuzz(seed) if seed % 3 == 0 else corpus.sheet_xml_doc(seed, 5, 60)
        pairs.append((corpus.strip_refs_partial(sx, seed, 0.0 if seed % 5 == 1 else 0.4, 0.0 if seed % 5 == 2 else 0.4), ss))
    got = _batch(eng, pairs)
    bad, unsup = [], 0
    for i, ((st, g), (sx, ss)) in enumerate(zip(got, pairs)):
        if st == 2:
            unsup += 1
            continue
        if (None if st == 3 else g) != _oracle(sx, ss):
            bad.append(i)
    assert not bad, bad[:10]
    assert unsup <= 0.02 * len(pairs)


def test_sheet_float_labels(eng):
    """Float column labels next to string labels (an object Index): named by repr(label), printed as
    f"{x: .6f}" with the zeros trimmed (pandas format.py _GenericArrayFormatter), a negative one keeps the
    other labels' blank; int + float labels (a Float64 Index) stay declined."""
    S = lambda i: '<c r="REF" t="s"><v>%d</v></c>' % i
    F = lambda v: '<c r="REF"><v>%s</v></c>' % v
    sst = ('<sst xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main"><si><t>Part</t></si>'
           '<si><t>  lead</t></si><si><t>4.5</t></si></sst>').encode()

    def sheet(labels):
        out = [corpus._XML_HEAD, '<row r="1">']
        out += [c.replace("REF", "%s1" % corpus.col_letters(k)) for k, c in enumerate(labels)]
        out.append("</row>")
        for i in range(3):
            out.append('<row r="%d">' % (i + 2))
            out += ['<c r="%s%d"><v>%d</v></c>' % (corpus.col_letters(k), i + 2, 10 * i + k) for k in range(len(labels))]
            out.append("</row>")
        out.append(corpus._XML_TAIL)
        return "".join(out).encode()
    cases = [[S(0), F("4.5")], [S(0), F("4.5"), F("4.5")], [S(0), F("-4.5"), F("0.1234567")], [S(1), F("2024.0")],
             [S(0), F("1e-7"), F("123456.7890123")], [S(0), F("4.5"), S(2)], [S(0), F("7.25"), '<c r="REF"><v>7</v></c>'],
             [S(0), F("1e5"), '<c r="REF" t="b"><v>1</v></c>'], [F("4.5"), '<c r="REF"><v>7</v></c>']]
    pairs = [(sheet(c), sst) for c in cases]
    got = _batch(eng, pairs)
    for i, ((st, g), (sx, ss)) in enumerate(zip(got, pairs)):
        if i < 5:
            assert st == 0, i
        if i == len(cases) - 1:
            assert st == 2
        if st != 2:
            assert g == _oracle(sx, ss), i
