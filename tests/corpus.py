"""Seeded synthetic corpora for the parity tests, the golden fixtures and bench.py.

Shapes follow SURVEY.md §8d: C2 RFQ CSV (8 columns, quoted description with `,` and `""`,
80 % empty notes), C3 Google-Sheets-style sheet XML + sharedStrings, C4 e-mail / PDF-extracted
text bodies (30 % CRLF, 1 % invalid UTF-8, blank lines around), C5 mixed.
"""
from __future__ import annotations

import random

import numpy as np

CSV_HEADER = "part_number,description,quantity,target_price,currency,delivery_date,supplier,notes"
_CURR = ["USD", "EUR", "INR", "GBP"]
_SUPP = ["Amphenol", "TE Connectivity", "Molex", "Hirose", "JAE", "Samtec", "Würth", "3M"]
_DESC = ["Circular MIL Spec Connector, 12 Pin", "RJ Field Ethernet Connector, IP67",
         "Flexible Coax Cable Assembly, 3 Meters", "D-Sub Connector, 25 Pin, Female",
         "Backshell, \"\"Straight\"\" type", "Heat shrink boot, size 14", "SMA Male to SMA Male",
         "Panel Mount, \"\"IP68\"\", black"]
_NOTE = ["urgent", "RoHS required", "quote MOQ", "alt part ok", "ship to Chennai"]


def rfq_csv(n_rows: int, seed: int = 0, newline: str = "\n") -> bytes:
    """Config C2: `n_rows` x 8 RFQ CSV."""
    rs = np.random.RandomState(seed)
    qty = rs.randint(1, 5001, size=n_rows)
    cents = rs.randint(1, 100000, size=n_rows)
    cur = rs.randint(0, len(_CURR), size=n_rows)
    day = rs.randint(1, 31, size=n_rows)
    sup = rs.randint(0, len(_SUPP), size=n_rows)
    dsc = rs.randint(0, len(_DESC), size=n_rows)
    dnum = rs.randint(0, 100, size=n_rows)
    note = rs.randint(0, 5 * len(_NOTE), size=n_rows)  # 80 % empty
    lines = [CSV_HEADER]
    ap = lines.append
    for i in range(n_rows):
        n = note[i]
        ap('PRT-%07d-A,"%s %d",%d,%d.%02d,%s,2025-09-%02d,%s,%s' % (
            i, _DESC[dsc[i]], dnum[i], qty[i], cents[i] // 100, cents[i] % 100, _CURR[cur[i]],
            day[i], _SUPP[sup[i]], _NOTE[n] if n < len(_NOTE) else ""))
    return (newline.join(lines) + newline).encode("utf-8")


# ------------------------------------------------------------------ CSV fuzz
_NA = ["", "#N/A", "#N/A N/A", "#NA", "-1.#IND", "-1.#QNAN", "-NaN", "-nan", "1.#IND", "1.#QNAN",
       "<NA>", "N/A", "NA", "NULL", "NaN", "None", "n/a", "nan", "null"]
_WORDS = ["abc", "x y", "PRT-001", "₹500", "日本語", "Würth", "a\"b", "-", "N.A.", "TRUE", "true",
          "False", "1,200", "0x10", "1_000", "inf", "-inf", "Infinity", "+5", "007", "-0", ".5",
          "5.", "1e3", "1E-2", " 8", "9 ", "1.50", "2.50", "12:30", "2025-09-01", "é", "naïve",
          "tab\there", "NAN", "nan ", "1e", "e5", "--1", "+-1", "1.2.3", "1e400", "1e-400",
          "123456789012345678", "9223372036854775807", "0.000001", "0.0000001", "1234567.891",
          "99999999.999999", "1e6", "1000001", "0.1", "0.25", "3.14159265358979"]


def _rand_cell(r: random.Random, kind: str) -> str:
    if kind == "int":
        return str(r.randint(-10000, 100000))
    if kind == "float2":
        return "%.2f" % (r.randint(0, 10 ** 7) / 100)
    if kind == "float":
        return repr(r.uniform(-1e4, 1e4))
    if kind == "floatwild":
        return r.choice(["%g", "%e", "%.10f", "%.3f"]) % (r.uniform(-1, 1) * 10 ** r.randint(-9, 12))
    if kind == "bool":
        return r.choice(["True", "False", "true", "false", "TRUE", "FALSE"])
    if kind == "word":
        return r.choice(_WORDS)
    if kind == "na":
        return r.choice(_NA)
    if kind == "text":
        n = r.randint(1, 4)
        return " ".join(r.choice(["conn", "pin", "M12", "IP67", "Ω", "µm", "ok", "10", "x"]) for _ in range(n))
    return ""


def bare_cr_blank_first(data: bytes) -> bool:
    """Necessary condition for the one tokenizer case the GPU path declines (RFQ_DOC_UNSUPPORTED)
    instead of reproducing: a line that starts with a blank right after a bare CR (or after the
    delimiter pandas swallows behind a blank line's CR).  pandas walks back to the previous LF there
    and re-reads earlier rows; what it prints depends on its read-buffer positions."""
    import re
    return re.search(rb"\r[,;\t]?[ \t]", data) is not None


def fuzz_csv(seed: int, max_rows: int = 40) -> bytes:
    """A small random CSV exercising quoting, NA strings, ragged rows, mixed dtypes, CRLF,
    blank lines, non-ASCII, separators other than ','."""
    r = random.Random(seed)
    sep = r.choice([",", ",", ",", ";", "\t"])
    nl = r.choice(["\n", "\n", "\r\n", "\r"]) if r.random() < 0.3 else "\n"
    ncols = r.randint(2, 7)
    nrows = r.randint(0, max_rows)
    kinds = [r.choice(["int", "float2", "float", "floatwild", "bool", "word", "text", "na", "mix"])
             for _ in range(ncols)]
    na_rate = [r.choice([0, 0, 0.1, 0.5]) for _ in range(ncols)]
    header = []
    for c in range(ncols):
        h = r.choice(["part", "qty", "price", "desc", "a", "b", "Sl No", "Ünit", "", "qty"])
        header.append(h)

    def quote(s: str, force: bool = False) -> str:
        if force or sep in s or '"' in s or "\n" in s or "\r" in s:
            return '"' + s.replace('"', '""') + '"'
        return s

    lines = [sep.join(quote(h) for h in header)]
    for _ in range(nrows):
        row = []
        for c in range(ncols):
            k = kinds[c]
            if k == "mix":
                k = r.choice(["int", "float2", "word", "na", "bool", "text"])
            if r.random() < na_rate[c]:
                k = "na"
            s = _rand_cell(r, k)
            if r.random() < 0.03:
                s = s + r.choice([sep, "\n", '"', " "])
            row.append(quote(s, force=r.random() < 0.1))
        x = r.random()
        if x < 0.04 and ncols > 2:
            row = row[: r.randint(1, ncols - 1)]  # short row -> NaN fill
        elif x < 0.05:
            row.append(_rand_cell(r, "word"))  # long row -> ParserError (unless first data row)
        lines.append(sep.join(row))
        if r.random() < 0.03:
            lines.append(r.choice(["", " ", "  \t" if sep != "\t" else "  "]))
    body = nl.join(lines)
    if r.random() < 0.8:
        body += nl
    data = body.encode("utf-8")
    if r.random() < 0.03:
        data = data.replace(b"a", b"\xe9", 1)  # latin-1 byte -> utf-8 attempt fails
    if r.random() < 0.03:
        data = b"\xef\xbb\xbf" + data
    return data


def csv_golden_cases():
    """(name, bytes, nrows) — hand-written edge cases + seeded fuzz."""
    E = []

    def add(name, s, nrows=1000):
        E.append((name, s if isinstance(s, bytes) else s.encode("utf-8"), nrows))

    add("basic", "a,b,c\n1,2,3\n4,5,6\n")
    add("no_trailing_newline", "a,b\n1,2\n3,4")
    add("crlf", "a,b\r\n1,x\r\n2,y\r\n")
    add("bare_cr", "a,b\r1,x\r2,y\r")
    add("blank_lines", "a,b\n\n1,2\n   \n\n3,4\n\n")
    add("quoted", 'a,b\n"x,1","he said ""hi"""\n"multi\nline",2\n')
    add("quote_mid_field", 'a,b\nx"y"z,1\n"x"y,2\n')
    add("crlf_in_quotes", 'a,b\n"l1\r\nl2",1\n')
    add("tab_in_cell", 'a,b\n"x\ty",1\n')
    add("floats_trim", "p,q\n20.00,a\n25.75,b\n146.50,c\n")
    add("floats_3dec", "p,q\n0.1,a\n0.25,b\n1234567.891,c\n")
    add("floats_sci_small", "p,q\n1.5,a\n2,b\n1e-7,c\n")
    add("floats_sci_large", "p,q\n1.5,a\n2,b\n12345678.1234,c\n")
    add("floats_large_ok", "p,q\n1.5,a\n2,b\n1234567.5,c\n")
    add("float_obj_holes", "p,q\n1.5,a\n,b\n1e-7,c\n0.1,d\n-2.25,e\n")
    add("int_na_float", "p,q\n1,a\n,b\n3,c\n")
    add("bool_col", "p,q\nTrue,1\nFalse,2\n")
    add("bool_na", "p,q\nTrue,1\n,2\nfalse,3\n")
    add("bool_mixed_case", "p,q\nTRUE,1\nfalse,2\nTrue,3\n")
    add("na_strings", "p,q\nNA,1\nN/A,2\nnull,3\nx,4\nNaN,5\n-,6\n")
    add("all_na_col", "p,q\n,1\n,2\n")
    add("all_na_row", "p,q\n1,2\n,\n3,4\n")
    add("numeric_forms", "p,q\n 8,+5\n9 ,007\n-0,12\n")
    add("float_forms", "p,q\n.5,1e3\n5.,1E-2\ninf,-inf\n")
    add("not_numbers", "p,q\n0x10,1_000\n\"1,200\",1e\n")
    add("uint64", "p,q\n9223372036854775808,1\n1,2\n")
    add("too_big", "p,q\n18446744073709551616,1\n1,2\n")
    add("int64_edge", "p,q\n9223372036854775807,-9223372036854775808\n1,2\n")
    add("mixed_keep_text", "p,q\n2.50,1\nabc,2\n1.50,3\n")
    add("short_rows", "a,b,c\n1,2\n3\n4,5,6\n")
    add("long_row_error", "a,b\n1,2\n3,4,5\n")
    add("implicit_index", "a,b\nx,1,2\ny,3,4\n")
    add("trailing_delim_rows", "a,b\n1,2,\n3,4,\n")
    add("first_row_longer_only", "a,b\nx,1,2\ny,3\n")
    add("header_only", "a,b,c\n")
    add("header_only_nonl", "a,b,c")
    add("empty", "")
    add("only_blank", "\n\n  \n")
    add("one_column", "a\n1\n2\n")
    add("semicolon", "a;b\n1,5;x\n2,5;y\n")
    add("tab", "a\tb\n1\tx\n2\ty\n")
    add("comma_first", "a,x;b,y\n1,2;3,4\n")
    add("dup_names", "a,a,b,a\n1,2,3,4\n")
    add("unnamed", "a,,b,\n1,2,3,4\n")
    add("bom", b"\xef\xbb\xbfa,b\n1,2\n")
    add("latin1", b"a,b\ncaf\xe9,1\n")
    add("latin1_semicolon", b"a;b\ncaf\xe9;1\n")
    add("unicode", "näme,₹\n日本語,₹500\nWürth,1\n")
    add("unterminated_quote", 'a,b\n"x,1\n2,3\n')
    add("nul_byte", b"a,b\nx\x00y,1\n")
    add("nrows_cap", "a,b\n" + "".join(f"{i},{i * 1.5}\n" for i in range(50)), 10)
    add("nrows_hides_bad_row", "a,b\n1,2\n3,4\n5,6,7,8\n", 2)
    add("nrows_hides_bad_utf8", b"a,b\n1,x\n2,y\n3,\xff\n", 2)
    add("ws_line_then_data", "a,b\n  x,1\n \n  ,2\n")
    add("neg_floats", "a,b\n-1.5,x\n2.25,y\n-0.0,z\n")
    add("big_ints_as_float", "a,b\n123456789012345678,x\n,y\n")
    add("exp_forms", "a,b\n1e5,x\n2.5E+3,y\n-3e-2,z\n")
    add("long_digits", "a,b\n0.12345678901234567890123,x\n123456789012345678901234567890,y\n")
    add("inf_col", "a,b\ninf,x\n1.5,y\n")
    add("str_numeric_header", "1,2\n3,4\n")
    add("spaces_in_cells", "a,b\n x ,1\n  ,2\n")
    add("quoted_empty", 'a,b\n"",1\n"NA",2\nx,3\n')
    add("escapes", 'a,b\n"l1\nl2\tq\rz",1\n')
    # column labels as to_string prints them: escapes and trim_front (pandas Index._format_flat)
    add("hdr_all_lead_blank", " a, b\n1,2\n")
    add("hdr_padded", " a , b \nx,y\n")
    add("hdr_some_lead_blank", " a,b\n1,2\n")
    add("hdr_newline", '"a\nb",c\n1,2\n')
    add("hdr_tab", "a\tx,b\n1,2\n")
    add("hdr_cr", '"a\rx",b\n1,2\n')
    add("hdr_nbsp", b"\xc2\xa0a,\xc2\xa0b\n1,2\n")
    add("hdr_tab_lead", '"\ta"," b"\n1,2\n')
    add("hdr_blank_empty_frame", " a, b\n")
    add("hdr_newline_empty_frame", '"a\nb", c\n')
    add("hdr_only_blanks", "   ,  x\n1,2\n")
    add("hdr_vt_ff", b"\x0ba,\x0cb\n1,2\n")
    add("hdr_ideographic_space", b"\xe3\x80\x80a,\xe2\x80\x83b\n1,2\n")
    add("hdr_dup_lead_blank", "  a,  a,  b\n1,2,3\n")
    add("hdr_semicolon_lead_blank", "  x;  y\n1;2\n")
    # integer overflow followed by something that is not a digit: an invalid token, not an overflow
    big = "1" * 21
    for i, vals in enumerate([[big + ".5", "2"], [big, "1.5"], [big + ".5", big], [big, big + ".5"], ["2", big + ".5"],
                              [big + " ", "2"], [big + "x", "2"], ["-" + big + ".5", "2"], ["-" + big, "2.5"],
                              ["9223372036854775808 ", "2"], [" 9223372036854775808", "2"],
                              ["18446744073709551616 ", "2"], ["18446744073709551616", "2 "],
                              ["99999999999999999999 ", "2"], ["9223372036854775808 ", "-2"], [big + "e0", "2"],
                              ["1" * 40 + ".25", "3"], ["9223372036854775808.5", "-1"]]):
        add(f"overflow_shape{i}", "a,b\n" + "".join(v + ",k\n" for v in vals))
    # bare CR: fine as long as no line starts with a blank ...
    add("bare_cr_blank_line", "a,b\r1,x\r  \r2,y\r")
    add("bare_cr_quoted_nl", 'a,b\r"l1\rl2",x\r2,y\r')
    add("crlf_blank_first", "a,b\r\n 1,x\r\n  2,y\r\n")
    add("lf_blank_first", "a,b\n 1,x\n\t2,y\n")
    # ... and the case the GPU path declines (RFQ_DOC_UNSUPPORTED): pandas walks back to the previous LF
    add("bare_cr_blank_first", "a,b\r 1,x\r2,y\r")
    add("bare_cr_blank_first_mixed", "a,b\n1,x\r 2,y\r\n3,z\n")
    add("bare_cr_swallowed_delim", "a,b\r1,x\r\r, 2,y\r")
    for seed in range(120):
        add(f"fuzz{seed}", fuzz_csv(seed), 1000)
    for seed in range(120, 140):
        add(f"fuzz{seed}_n5", fuzz_csv(seed), 5)
    add("rfq200", rfq_csv(200, seed=1), 1000)
    add("rfq2000_cap", rfq_csv(2000, seed=2), 1000)
    return E


# ------------------------------------------------------------------ text bodies (C4)
_BODY_SEEDS = [
    "hello amphenol,\n\npls send ur best prices for below. shipment to chennai by 10-sep is needed.\n\n"
    "- part no: 62GB-56T-16-8S  \n  qty: 750  \n  desc: Circular connector\n",
    "Dear Team,\n\nWe are looking to receive a quotation for the following components for our\n"
    "July production batch.\n\nCompany: ABC Manufacturing Ltd.\nPhone: +91 98765 43210\n",
    "Request for Quotation (RFQ)\nTo: sales@example.com\nSubject: RFQ - RF Components\n",
]
_WS_EDGE = ["", " ", "\n", "\r\n", "\t", "\x0b\x0c", "\x1c\x1d\x1e\x1f", "", " ",
            " ", "   ", "  ", " ", " ", "　",
            "​", "﻿", "\x00"]


def text_body(seed: int) -> bytes:
    r = random.Random(seed)
    parts = [r.choice(_WS_EDGE) for _ in range(r.randint(0, 3))]
    body = r.choice(_BODY_SEEDS)
    for i in range(r.randint(0, 30)):
        body += "%d) part %s qty: %d @ ₹%d/pc, total %s\n" % (
            i + 1, "MP-%04d" % r.randint(0, 9999), r.randint(1, 5000), r.randint(1, 999),
            format(r.randint(1000, 99999), ","))
    parts.append(body)
    parts += [r.choice(_WS_EDGE) for _ in range(r.randint(0, 3))]
    s = "".join(parts)
    if r.random() < 0.3:
        s = s.replace("\n", "\r\n")
    elif r.random() < 0.1:
        s = s.replace("\n", "\r")
    data = s.encode("utf-8")
    if r.random() < 0.2 and data:
        # sprinkle invalid UTF-8
        bad = [b"\xff", b"\xc0\xaf", b"\xed\xa0\x80", b"\xf4\x90\x80\x80", b"\xe2\x82", b"\x80",
               b"\xf0\x9f", b"\xc3"]
        ba = bytearray(data)
        for _ in range(r.randint(1, 4)):
            p = r.randint(0, len(ba))
            ba[p:p] = r.choice(bad)
        data = bytes(ba)
    return data


def txt_golden_cases():
    E = []

    def add(name, b):
        E.append((name, b if isinstance(b, bytes) else b.encode("utf-8")))

    add("empty", b"")
    add("only_ws", " \n\t\r\n ")
    add("plain", "hello\nworld")
    add("crlf", "  a\r\nb\rc\r\n\r\n")
    add("cr_at_end", "abc\r")
    add("bom_kept", b"\xef\xbb\xbfhello ")
    add("invalid", b"a\xc0\xafb\xed\xa0\x80c\xf4\x90\x80\x80d")
    add("truncated_tail", b"abc\xe2\x82")
    add("truncated_then_ascii", b"abc\xe2\x82def\xf0\x9f\x98x")
    add("nul_not_stripped", b"\x00 a \x00")
    add("zwsp_not_stripped", "​ a ​")
    add("unicode_ws", "　  x     ")
    add("fs_gs", "\x1c\x1d\x1e\x1fx\x1f")
    add("vt_ff", "\x0b\x0cx\x0c\x0b")
    add("interior_kept", "a  \n\n  b\t　c")
    add("all_invalid", b"\xff\xfe\x80")
    add("ws_after_invalid", b" \xff x \xff ")
    add("four_byte", "😀 ok 😀")
    add("cr_split_invalid", b"a\r\xff\nb")
    add("e0_overlong", b"x\xe0\x80\x80y\xe0\xa0\x80z")
    add("f0_overlong", b"x\xf0\x80\x80\x80y\xf0\x90\x80\x80z")
    add("lead_then_lead", b"\xe2\xe2\x82\xacx")
    for seed in range(60):
        add(f"body{seed}", text_body(seed))
    return E


# ------------------------------------------------------------------ sheet XML (C3)
_XML_HEAD = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>\n'
             '<worksheet xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" '
             'xmlns:r="http://schemas.openxmlformats.org/officeDocument/2006/relationships">'
             '<sheetPr><pageSetUpPr/></sheetPr><sheetViews><sheetView workbookViewId="0"/>'
             '</sheetViews><sheetFormatPr customHeight="1" defaultColWidth="14.43" '
             'defaultRowHeight="15.0"/><sheetData>')
_XML_TAIL = '</sheetData><drawing r:id="rId1"/></worksheet>'


def col_letters(c: int) -> str:
    s = ""
    c += 1
    while c:
        c, rem = divmod(c - 1, 26)
        s = chr(65 + rem) + s
    return s


def xml_escape(s: str) -> str:
    return s.replace("&", "&amp;").replace("<", "&lt;").replace(">", "&gt;")


def sheet_xml_doc(seed: int, min_rows: int = 5, max_rows: int = 60):
    """One Google-Sheets-style worksheet + its sharedStrings. Returns (sheet_xml, sst_xml)."""
    r = random.Random(seed)
    ncols = r.randint(4, 12)
    nrows = r.randint(min_rows, max_rows)
    strings, index = [], {}

    def sid(s):
        if s not in index:
            index[s] = len(strings)
            strings.append(s)
        return index[s]

    kinds = [r.choice(["str", "int", "price", "float", "str", "mixed"]) for _ in range(ncols)]
    names = ["Sl No", "Part Number", "Description", "Quantity", "Target Price (INR)",
             "Required Delivery Date", "Supplier & Co", "Notes <x>", "Ünit", "Lead", "MOQ", "UOM"]
    out = [_XML_HEAD, '<row r="1">']
    for c in range(ncols):
        out.append('<c r="%s1" s="1" t="s"><v>%d</v></c>' % (col_letters(c), sid(names[c])))
    out.append("</row>")
    rownum = 1
    for i in range(nrows):
        rownum += 1
        if r.random() < 0.05:
            rownum += r.randint(1, 3)  # sparse: missing rows
        if r.random() < 0.03:
            out.append('<row r="%d" ht="15.75" customHeight="1"/>' % rownum)
            continue
        out.append('<row r="%d">' % rownum)
        for c in range(ncols):
            if r.random() < 0.08:
                if r.random() < 0.5:
                    continue  # missing cell
                out.append('<c r="%s%d" s="2"/>' % (col_letters(c), rownum))
                continue
            k = kinds[c]
            if k == "mixed":
                k = r.choice(["str", "int", "price"])
            ref = "%s%d" % (col_letters(c), rownum)
            if k == "str":
                s = r.choice(["RJF544", "62IN-56T12-8S", "Circular MIL Spec Connector, 12 Pin",
                              "R&D sample", "a<b", "Würth Elektronik", "日本語", "25-Sep-2025",
                              "D-Sub Connector, 25 Pin, Female, Solder Cup", "x %d" % r.randint(0, 50)])
                out.append('<c r="%s" s="2" t="s"><v>%d</v></c>' % (ref, sid(s)))
            elif k == "int":
                out.append('<c r="%s" s="2"><v>%d.0</v></c>' % (ref, r.randint(1, 5000)))
            elif k == "price":
                out.append('<c r="%s" s="2"><v>%s</v></c>' % (ref, repr(r.randint(1, 10 ** 6) / 100)))
            else:
                out.append('<c r="%s" s="2"><v>%s</v></c>' % (ref, repr(r.uniform(0, 1000))))
        out.append("</row>")
    for rr in range(rownum + 1, rownum + 1 + r.randint(0, 30)):
        out.append('<row r="%d" ht="15.75" customHeight="1"/>' % rr)
    out.append(_XML_TAIL)
    sst = ['<?xml version="1.0" encoding="UTF-8" standalone="yes"?>\n'
           '<sst xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" '
           'count="%d" uniqueCount="%d">' % (len(strings), len(strings))]
    for s in strings:
        sst.append("<si><t>%s</t></si>" % xml_escape(s))
    sst.append("</sst>")
    return "".join(out).encode("utf-8"), "".join(sst).encode("utf-8")


def sheet_xml_fuzz(seed: int):
    """A small adversarial worksheet: every cell type openpyxl distinguishes (n, s, str,
    inlineStr, b, e), entities, numeric-looking and NA strings, duplicate / blank / numeric
    headers, sparse rows and columns, 17-digit numbers. Returns (sheet_xml, sst_xml)."""
    r = random.Random(seed)
    ncols = r.randint(2, 7)
    nrows = r.randint(0, 25)
    strings, index = [], {}

    def sid(s):
        if s not in index:
            index[s] = len(strings)
            strings.append(s)
        return index[s]

    def cell(ref, kind):
        if kind == "empty":
            return '<c r="%s" s="2"/>' % ref
        if kind == "emptyv":
            return '<c r="%s" t="s"><v></v></c>' % ref
        if kind == "int":
            return '<c r="%s"><v>%d</v></c>' % (ref, r.randint(-1000, 100000))
        if kind == "intf":
            return '<c r="%s" s="2"><v>%d.0</v></c>' % (ref, r.randint(0, 5000))
        if kind == "price":
            return '<c r="%s"><v>%s</v></c>' % (ref, repr(r.randint(1, 10 ** 7) / 100))
        if kind == "float":
            return '<c r="%s"><v>%s</v></c>' % (ref, repr(r.uniform(-1, 1) * 10 ** r.randint(-8, 9)))
        if kind == "float17":
            return '<c r="%s"><v>%.17g</v></c>' % (ref, r.uniform(0, 1000))
        if kind == "sci":
            return '<c r="%s"><v>%sE%d</v></c>' % (ref, r.randint(1, 999) / 100, r.randint(-10, 10))
        if kind == "bool":
            return '<c r="%s" t="b"><v>%d</v></c>' % (ref, r.randint(0, 1))
        if kind == "err":
            return '<c r="%s" t="e"><v>#DIV/0!</v></c>' % ref
        words = ["RJF544", "a & b", "x<y>z", "Würth", "日本語", "12", "4.50", "NA", "N/A", "true", "False",
                 "TRUE", "inf", "", "  pad ", "line1\nline2", "tab\there", "1e5", "nan", "-", "0x1F",
                 "quote\"s", "it's", "₹500", "null"]
        w = r.choice(words)
        if kind == "sst":
            return '<c r="%s" t="s"><v>%d</v></c>' % (ref, sid(w))
        if kind == "str":
            return '<c r="%s" t="str"><f>A1&amp;"x"</f><v>%s</v></c>' % (ref, xml_escape(w).replace("\n", "&#10;"))
        if kind == "inline":
            return '<c r="%s" t="inlineStr"><is><t>%s</t></is></c>' % (ref, xml_escape(w))
        raise ValueError(kind)

    col_kinds = [r.choice(["int", "intf", "price", "float", "bool", "sst", "mixed", "numstr", "boolish",
                           "float17", "sci"]) for _ in range(ncols)]
    out = [_XML_HEAD]
    rownum = 0
    hdr_words = ["Sl No", "Part", "Part", "Qty", "", "a&b", "Ünit", "Qty", "2024"]
    int_header = r.random() < 0.12
    for i in range(nrows + 1):
        rownum += 1
        if r.random() < 0.08:
            rownum += r.randint(1, 2)
        if i and r.random() < 0.05:
            out.append('<row r="%d" ht="15.75" customHeight="1"/>' % rownum)
            continue
        out.append('<row r="%d" spans="1:%d">' % (rownum, ncols))
        for c in range(ncols):
            ref = "%s%d" % (col_letters(c), rownum)
            if r.random() < 0.1:
                continue
            if i == 0 and rownum == 1:
                k = r.random()
                if int_header:
                    out.append('<c r="%s"><v>%d</v></c>' % (ref, r.choice([2024, 2025, 2026, 7, 12345, 2024])))
                elif k < 0.95:
                    out.append('<c r="%s" t="s"><v>%d</v></c>' % (ref, sid(r.choice(hdr_words))))
                elif k < 0.98:
                    out.append(cell(ref, "int"))
                elif k < 0.99:
                    out.append(cell(ref, "price"))
                else:
                    out.append(cell(ref, "bool"))
                continue
            k = col_kinds[c]
            if k == "mixed":
                k = r.choice(["int", "price", "sst", "bool", "err", "inline", "str", "empty", "emptyv"])
            elif k == "numstr":
                k = r.choice(["sst", "int", "price"])
            elif k == "boolish":
                k = r.choice(["bool", "bool", "empty"])
            if r.random() < 0.05:
                k = r.choice(["empty", "err", "emptyv"])
            out.append(cell(ref, k))
        out.append("</row>")
    out.append(_XML_TAIL)
    sst = ['<?xml version="1.0" encoding="UTF-8" standalone="yes"?>\r\n'
           '<sst xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" '
           'count="%d" uniqueCount="%d">' % (len(strings), len(strings))]
    for i, s in enumerate(strings):
        if i % 7 == 3 and s:
            h = len(s) // 2
            sst.append('<si><r><rPr><b/></rPr><t xml:space="preserve">%s</t></r><r><t>%s</t></r>'
                       '<rPh sb="0" eb="1"><t>PHON</t></rPh></si>' % (xml_escape(s[:h]), xml_escape(s[h:])))
        else:
            sst.append("<si><t>%s</t></si>" % xml_escape(s))
    sst.append("</sst>")
    return "".join(out).encode("utf-8"), "".join(sst).encode("utf-8")


def sheet_xml_dates(seed: int, p_td: float = 0.04):
    """A worksheet with date-styled numbers (style 3 = a date format, style 4 = a timedelta format):
    whole days, day fractions, serials below 1 (times) and below 60 (the 1900 leap-year bug),
    out-of-range serials, date columns with blanks / error cells / strings / plain numbers.
    Returns (sheet_xml, sst_xml, date_ids, td_ids, date1904)."""
    r = random.Random(1000003 * seed + 17)
    ncols = r.randint(2, 6)
    nrows = r.randint(1, 30)
    date1904 = r.random() < 0.15
    strings, index = [], {}

    def sid(s):
        if s not in index:
            index[s] = len(strings)
            strings.append(s)
        return index[s]

    def serial(kind):
        if kind == "day":
            return "%d" % r.randint(20000, 60000)
        if kind == "dayf":
            return "%d.0" % r.randint(20000, 60000)
        if kind == "sec":
            return repr(r.randint(30000, 50000) + r.randint(0, 86399) / 86400)
        if kind == "ms":
            return repr(r.uniform(30000, 50000))
        if kind == "time":
            return repr(r.choice([0.5, 0.25, 0.75001, r.random(), 0.999999999, 0.0]))
        if kind == "early":
            return r.choice(["0", "1", "59", "60", "61", "1.5", "59.999", "31.25", "0.99999999999"])
        if kind == "edge":
            return r.choice(["2958465", "2958465.9999", "2958465.99999999", "2958466", "1e7", "3e6",
                             "4000001", "99999999999", "2958464.5"])
        raise ValueError(kind)

    def iso():
        y, mo, da = r.randint(1990, 2040), r.randint(1, 12), r.randint(1, 28)
        form = "bad" if r.random() < 0.004 else r.choice(["d", "d", "T", "T", " ", "hm", "ms", "us0"])
        base = "%04d-%02d-%02d" % (y, mo, da)
        hms = "%02d:%02d:%02d" % (r.randint(0, 23), r.randint(0, 59), r.randint(0, 59))
        return {"d": base, "T": base + "T" + hms, " ": base + " " + hms, "hm": base + "T" + hms[:5],
                "ms": base + "T" + hms + ".%03d" % r.randint(0, 999), "us0": base + "T" + hms + ".250000",
                "bad": base + "T" + hms + "Z"}[form]

    col_kinds = [r.choice(["day", "day", "dayf", "sec", "ms", "time", "early", "edge", "anydate", "iso",
                           "date+blank", "date+str", "date+num", "int", "price", "sst"])
                 for _ in range(ncols)]
    use_td = r.random() < p_td
    out = [_XML_HEAD]
    for i in range(nrows + 1):
        rownum = i + 1
        out.append('<row r="%d">' % rownum)
        for c in range(ncols):
            ref = "%s%d" % (col_letters(c), rownum)
            if i == 0:
                h = r.random()
                if c and h < 0.06:      # a date / time / bool as a column label, next to string labels
                    out.append('<c r="%s" s="3"><v>%s</v></c>' % (ref, r.choice(["45000", "45000", "45007.5", "0.25"])))
                elif c and h < 0.08:
                    out.append('<c r="%s" t="b"><v>%d</v></c>' % (ref, r.randint(0, 1)))
                else:
                    out.append('<c r="%s" t="s"><v>%d</v></c>' % (ref, sid(r.choice(["Date", "Delivery", "ETA", "Qty", "When"]))))
                continue
            k = col_kinds[c]
            if k == "int":
                out.append('<c r="%s"><v>%d</v></c>' % (ref, r.randint(1, 5000)))
            elif k == "price":
                out.append('<c r="%s" s="2"><v>%s</v></c>' % (ref, repr(r.randint(1, 10 ** 6) / 100)))
            elif k == "sst":
                out.append('<c r="%s" t="s"><v>%d</v></c>' % (ref, sid(r.choice(["RJF544", "tbd", "NA", "asap"]))))
            elif k == "iso":
                out.append('<c r="%s" t="d"><v>%s</v></c>' % (ref, iso()))
            else:
                if k == "anydate":
                    kk = r.choice(["day", "dayf", "sec", "ms", "time", "early", "edge"])
                elif k == "date+blank":
                    kk = r.choice(["day", "day", "blank", "err", "nastr"])
                elif k == "date+str":
                    kk = r.choice(["day", "sec", "str", "blank"])
                elif k == "date+num":
                    kk = r.choice(["day", "ms", "num", "bool"])
                else:
                    kk = k if r.random() > 0.06 else r.choice(["blank", "err"])
                if kk == "blank":
                    if r.random() < 0.5:
                        out.append('<c r="%s" s="3"/>' % ref)
                elif kk == "err":
                    out.append('<c r="%s" s="3" t="e"><v>#N/A</v></c>' % ref)
                elif kk == "nastr":
                    out.append('<c r="%s" t="s"><v>%d</v></c>' % (ref, sid(r.choice(["NA", "", "null"]))))
                elif kk == "str":
                    out.append('<c r="%s" t="s"><v>%d</v></c>' % (ref, sid(r.choice(["tbd", "2024-01-05", "12"]))))
                elif kk == "num":
                    out.append('<c r="%s"><v>%s</v></c>' % (ref, r.choice(["5", "45000", "2.5"])))
                elif kk == "bool":
                    out.append('<c r="%s" t="b"><v>%d</v></c>' % (ref, r.randint(0, 1)))
                else:
                    style = 4 if use_td and r.random() < 0.3 else 3
                    v = serial(kk)
                    if style == 4 and kk == "edge":      # the edges of a duration, not of a calendar date
                        v = r.choice(["0", "0.000001", "149999.5", "-3.25", "1e9", "1e10", "-0.0000005", "0.99999999999",
                                      "1.0000000001", "-149999.999", "0.5000005"])
                    out.append('<c r="%s" s="%d"><v>%s</v></c>' % (ref, style, v))
        out.append("</row>")
    out.append(_XML_TAIL)
    sst = ['<?xml version="1.0" encoding="UTF-8" standalone="yes"?>\r\n'
           '<sst xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" '
           'count="%d" uniqueCount="%d">' % (len(strings), len(strings))]
    for s in strings:
        sst.append("<si><t>%s</t></si>" % xml_escape(s))
    sst.append("</sst>")
    return "".join(out).encode("utf-8"), "".join(sst).encode("utf-8"), {3, 4}, {4}, date1904


def strip_refs(sheet_xml: bytes, rows: bool = True, cells: bool = True) -> bytes:
    """The same worksheet as a writer that omits r= would emit it (rows and cells are then numbered
    by position: openpyxl worksheet/_reader.py row_counter / col_counter).  Only meaningful for
    dense sheets (no skipped rows or columns)."""
    import re
    s = sheet_xml
    if rows:
        s = re.sub(rb'<row r="\d+"', b"<row", s)
    if cells:
        s = re.sub(rb'<c r="[A-Z]+\d+"', b"<c", s)
    return s


def strip_refs_partial(sheet_xml: bytes, seed: int, p_row: float = 0.4, p_cell: float = 0.4) -> bytes:
    """The same worksheet with r= removed from a random subset of its rows and cells (a writer that
    only states a reference where the position jumps): an element without r= continues the count of
    the element before it (openpyxl worksheet/_reader.py parse_row / parse_cell)."""
    import re
    r = random.Random(seed)
    s = re.sub(rb'<row r="\d+"', lambda m: b"<row" if r.random() < p_row else m.group(0), sheet_xml)
    return re.sub(rb'<c r="[A-Z]+\d+"', lambda m: b"<c" if r.random() < p_cell else m.group(0), s)


# ------------------------------------------------------------------ .xlsx containers
_WB_XML = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>\n'
           '<workbook xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" '
           'xmlns:r="http://schemas.openxmlformats.org/officeDocument/2006/relationships"><sheets>%s</sheets></workbook>')
_RELS_XML = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>\n'
             '<Relationships xmlns="http://schemas.openxmlformats.org/package/2006/relationships">%s</Relationships>')
_REL_WS = "http://schemas.openxmlformats.org/officeDocument/2006/relationships/worksheet"
_REL_SST = "http://schemas.openxmlformats.org/officeDocument/2006/relationships/sharedStrings"


def make_xlsx(sheets, sst_xml, names=None, level: int = 6) -> bytes:
    """An .xlsx container (zip, DEFLATE) around worksheet parts + one sharedStrings part."""
    import io
    import zipfile
    names = names or ["Sheet%d" % (i + 1) for i in range(len(sheets))]
    sh = "".join('<sheet name="%s" sheetId="%d" r:id="rId%d"/>' % (xml_escape(nm), i + 1, i + 1) for i, nm in enumerate(names))
    rel = "".join('<Relationship Id="rId%d" Type="%s" Target="worksheets/sheet%d.xml"/>' % (i + 1, _REL_WS, i + 1)
                  for i in range(len(sheets)))
    rel += '<Relationship Id="rId%d" Type="%s" Target="sharedStrings.xml"/>' % (len(sheets) + 1, _REL_SST)
    buf = io.BytesIO()
    with zipfile.ZipFile(buf, "w", zipfile.ZIP_DEFLATED if level else zipfile.ZIP_STORED, compresslevel=level or None) as z:
        z.writestr("xl/workbook.xml", _WB_XML % sh)
        z.writestr("xl/_rels/workbook.xml.rels", _RELS_XML % rel)
        z.writestr("xl/sharedStrings.xml", sst_xml)
        for i, x in enumerate(sheets):
            z.writestr("xl/worksheets/sheet%d.xml" % (i + 1), x)
    return buf.getvalue()


def mixed_batch(n: int, seed: int = 0):
    """Config C5 in miniature: (documents, kinds) — 40 % CSV (10-2000 rows), 30 % .xlsx, 30 % text bodies."""
    r = random.Random(seed)
    docs, kinds = [], []
    for i in range(n):
        x = r.random()
        if x < 0.4:
            docs.append(rfq_csv(r.choice([10, 25, 60, 150, 400, 2000]), seed=seed * 100003 + i)); kinds.append("csv")
        elif x < 0.7:
            sx, ss = sheet_xml_doc(seed * 100003 + i)
            docs.append(make_xlsx([sx], ss)); kinds.append("xlsx")
        else:
            docs.append(text_body(seed * 100003 + i)); kinds.append("txt")
    return docs, kinds


# ------------------------------------------------------------------ .json documents
def _json_value(r, depth=0):
    k = r.random()
    if depth > 5 or k < 0.45:
        c = r.random()
        if c < 0.25:
            return r.choice([0, 1, -1, 7, 1200, -45, 10 ** 12, 2 ** 63, -2 ** 70, 12345678901234567890123])
        if c < 0.5:
            return r.choice([0.0, -0.0, 1.5, 4.25, 1250.75, 1e15, 1e16, 1e-4, 9.5e-5, 1e22, 2.5e-7, 0.1, 1 / 3, 1e300,
                             r.uniform(-1e6, 1e6), round(r.uniform(0, 1e5), 2), r.uniform(0, 1) * 10 ** r.randint(-30, 30)])
        if c < 0.6:
            return r.choice([True, False, None])
        return r.choice(["", "RJF544", "a & b", "Würth", "日本語", "₹500", "line1\nline2", "tab\there", "q\"uote", "back\\slash",
                         "\u0001ctl", "emoji 😀", "/slash", "x" * r.randint(0, 40), " lead", "null", "1e5", "\x7f"])
    if k < 0.75:
        return {r.choice(["part", "qty", "price", "desc", "ünit", "a b", "", "k%d" % r.randint(0, 50), "n\"q", "items"]) + str(i):
                _json_value(r, depth + 1) for i in range(r.randint(0, 5))}
    return [_json_value(r, depth + 1) for _ in range(r.randint(0, 5))]


def json_doc(seed: int) -> bytes:
    """A valid .json document: nested objects / arrays, every scalar type, escapes written both raw and as
    \\uXXXX, assorted whitespace and newline conventions, NaN / Infinity, numbers in odd spellings."""
    import json
    r = random.Random(seed * 7919 + 5)
    obj = _json_value(r)
    style = r.randint(0, 4)
    if style == 0:
        t = json.dumps(obj)                                   # ensure_ascii: non-ASCII as \uXXXX (surrogate pairs too)
    elif style == 1:
        t = json.dumps(obj, ensure_ascii=False, separators=(",", ":"))
    elif style == 2:
        t = json.dumps(obj, indent=r.choice([1, 2, 4, "\t"]), ensure_ascii=False)
    elif style == 3:
        t = json.dumps(obj, indent=2).replace("\n", "\r\n")
    else:
        t = "  " + json.dumps(obj, ensure_ascii=False, indent=3).replace("\n", "\r") + " \n"
    if r.random() < 0.2:   # spellings json.dumps never writes
        t = t.replace("1.5", r.choice(["1.50", "15e-1", "0.15E+1", "1.5000000000000000", "1.4999999999999999999"]))
        t = t.replace("1200", r.choice(["1200", "1.2e3", "12E2", "1200.0"]))
    if r.random() < 0.1:
        t = t.replace("null", r.choice(["NaN", "Infinity", "-Infinity", "null"]))
    return t.encode("utf-8")


def json_doc_broken(seed: int) -> bytes:
    """A .json document with one defect: a truncation, a swapped / dropped / inserted character."""
    r = random.Random(seed * 104729 + 11)
    d = bytearray(json_doc(seed))
    if not d:
        return b""
    k = r.randint(0, 5)
    pos = r.randrange(len(d))
    if k == 0:
        del d[pos:]
    elif k == 1:
        del d[pos]
    elif k == 2:
        d[pos] = r.choice(b'{}[],:"\\ xe.-0\n\t\x01')
    elif k == 3:
        d.insert(pos, r.choice(b'{}[],:"\\ x.-0\n'))
    elif k == 4:
        d += r.choice([b" x", b",", b"]", b"}", b" 1", b"\xef\xbb\xbf"])
    else:
        d = bytearray(r.choice([b"\xef\xbb\xbf", b"", b" "])) + d
    return bytes(d)

