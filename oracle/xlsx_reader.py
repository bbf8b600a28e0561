"""CPU ORACLE (test infrastructure): restatement of the reader `pd.read_excel` uses for .xlsx.

The reference calls `pd.read_excel(path, sheet_name=None, nrows=...)` (app/file_parser.py:225).
pandas delegates to openpyxl 3.1.5 (pinned in the reference's uv.lock:577-578), which is NOT
installed in this image and cannot be installed (no network), so its published algorithm is
restated here with xml.etree (openpyxl itself iterparses with ElementTree/expat) and fed to the
REAL pandas `TextParser`, exactly as `pandas/io/excel/_base.py:913-937` does.

Follows (openpyxl 3.1.5, source not in this container; behaviour restated from its public code):
  openpyxl/reader/excel.py        ExcelReader.read: workbook.xml sheets + rels, shared strings,
                                  styles -> date formats
  openpyxl/reader/strings.py      read_string_table: per <si> Text.content, .replace('x005F_','')
  openpyxl/cell/text.py           Text.content: <t> + every <r><t>, phonetic runs excluded
  openpyxl/worksheet/_reader.py   WorkSheetParser.parse_row / parse_cell, _cast_number
  openpyxl/worksheet/_read_only.py  _cells_by_row (row gaps -> empty rows, out-of-order rows
                                  dropped), _get_row (width = last cell's column)
  openpyxl/styles/numbers.py      is_date_format, builtin formats
  openpyxl/utils/datetime.py      from_excel
and (installed pandas 3.0.2, read directly):
  pandas/io/excel/_openpyxl.py:596-646   _convert_cell, get_sheet_data
  pandas/io/excel/_base.py:636-697,700-937  _calc_rows, parse, _parse_sheet

Pinned by the golden vectors of `.cache/42/cache.db` rows 12 and 14 (tests/golden/): shared
strings, numeric / string cells, row gaps, header inference, float formatting.  Beyond those two
sheets the READER is a restatement of openpyxl 3.1.5's published code (openpyxl is not installed
here and cannot be), while everything downstream of it — TextParser, dtype inference, to_string —
is the real pandas.  What the restatement is checked against:
  * date serials: ECMA-376 part 1 §18.17.4 (serial 1 = 1900-01-01, the 1900 leap-year gap at 60, the
    1904 date system) and the values openpyxl's own tests use (openpyxl/utils/tests/test_datetime.py,
    3.1.5: from_excel(39385) = 2007-10-30, 0.5 -> 12:00, 40372.27616898148 -> 2010-07-13 06:37:41) in
    tests/test_hostcore.py::test_excel_serial_known_answers;
  * timedelta-styled numbers: CPython's own datetime.timedelta does the arithmetic here, and the device
    restatement is compared with it microsecond for microsecond on 12 000 values
    (tests/test_hostcore.py::test_timedelta_conversion_matches_python);
  * t="d" cells: datetime.fromisoformat of the running CPython.
PARITY UNPINNED only for what has no external answer in this container: the order in which
openpyxl's reader numbers rows and cells without r= (row_counter / col_counter, restated from
worksheet/_reader.py) and its treatment of inline strings, boolean and error cells.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
from __future__ import annotations

import datetime
import io
import posixpath
import re
import zipfile
import xml.etree.ElementTree as ET

import numpy as np
import pandas as pd
from pandas.io.parsers import TextParser

NS_MAIN = "http://schemas.openxmlformats.org/spreadsheetml/2006/main"
NS_REL = "http://schemas.openxmlformats.org/officeDocument/2006/relationships"
NS_PKG_REL = "http://schemas.openxmlformats.org/package/2006/relationships"


def _t(name: str) -> str:
    return "{%s}%s" % (NS_MAIN, name)


# ---------------------------------------------------------------- styles -> date formats
_BUILTIN_FORMATS = {
    0: "General", 1: "0", 2: "0.00", 3: "#,##0", 4: "#,##0.00", 5: '"$"#,##0_);("$"#,##0)',
    6: '"$"#,##0_);[Red]("$"#,##0)', 7: '"$"#,##0.00_);("$"#,##0.00)',
    8: '"$"#,##0.00_);[Red]("$"#,##0.00)', 9: "0%", 10: "0.00%", 11: "0.00E+00", 12: "# ?/?",
    13: "# ??/??", 14: "mm-dd-yy", 15: "d-mmm-yy", 16: "d-mmm", 17: "mmm-yy", 18: "h:mm AM/PM",
    19: "h:mm:ss AM/PM", 20: "h:mm", 21: "h:mm:ss", 22: "m/d/yy h:mm", 37: "#,##0_);(#,##0)",
    38: "#,##0_);[Red](#,##0)", 39: "#,##0.00_);(#,##0.00)", 40: "#,##0.00_);[Red](#,##0.00)",
    41: r'_(* #,##0_);_(* \(#,##0\);_(* "-"_);_(@_)',
    42: r'_("$"* #,##0_);_("$"* \(#,##0\);_("$"* "-"_);_(@_)',
    43: r'_(* #,##0.00_);_(* \(#,##0.00\);_(* "-"??_);_(@_)',
    44: r'_("$"* #,##0.00_)_("$"* \(#,##0.00\)_("$"* "-"??_)_(@_)', 45: "mm:ss",
    46: "[h]:mm:ss", 47: "mmss.0", 48: "##0.0E+0", 49: "@",
}
_STRIP_RE = re.compile(r"\"[^\"]*\"|\[(?!hh?\]|mm?\]|ss?\])[^\]]*\]")
_TIMEDELTA_RE = re.compile(r"\[hh?\](:mm(:ss(\.0*)?)?)?|\[mm?\](:ss(\.0*)?)?|\[ss?\](\.0*)?", re.I)


def is_date_format(fmt) -> bool:
    if fmt is None:
        return False
    fmt = fmt.split(";")[0]
    fmt = _STRIP_RE.sub("", fmt)
    return re.search(r"(?<![_\\])[dmhysDMHYS]", fmt) is not None


def is_timedelta_format(fmt) -> bool:
    if fmt is None:
        return False
    fmt = fmt.split(";")[0]
    return _TIMEDELTA_RE.search(fmt) is not None


def date_style_ids(styles_xml: bytes | None) -> "tuple[set[int], set[int]]":
    """Indices into cellXfs whose number format is a date (and timedelta) format."""
    if not styles_xml:
        return set(), set()
    root = ET.fromstring(styles_xml)
    custom = {}
    nf = root.find(_t("numFmts"))
    if nf is not None:
        for el in nf.findall(_t("numFmt")):
            custom[int(el.get("numFmtId"))] = el.get("formatCode")
    dates, tds = set(), set()
    xfs = root.find(_t("cellXfs"))
    if xfs is not None:
        for idx, xf in enumerate(xfs.findall(_t("xf"))):
            fid = int(xf.get("numFmtId", "0"))
            code = custom.get(fid, _BUILTIN_FORMATS.get(fid))
            if is_date_format(code):
                dates.add(idx)
                if is_timedelta_format(code):
                    tds.add(idx)
    return dates, tds


# ---------------------------------------------------------------- workbook plumbing
def open_workbook(data: bytes) -> dict:
    """zip -> {'sheets': [(name, xml bytes)], 'sst': bytes|None, 'date_styles', 'td_styles',
    'date1904'} in workbook order (openpyxl/reader/excel.py)."""
    z = zipfile.ZipFile(io.BytesIO(data))
    names = set(z.namelist())
    wb_path = "xl/workbook.xml"
    rels = {}
    rels_path = "xl/_rels/workbook.xml.rels"
    if rels_path in names:
        for r in ET.fromstring(z.read(rels_path)).findall("{%s}Relationship" % NS_PKG_REL):
            target = r.get("Target")
            if target.startswith("/"):
                target = target[1:]
            else:
                target = posixpath.normpath(posixpath.join("xl", target))
            rels[r.get("Id")] = (r.get("Type"), target)
    root = ET.fromstring(z.read(wb_path))
    pr = root.find(_t("workbookPr"))
    date1904 = pr is not None and pr.get("date1904") in ("1", "true")
    sheets = []
    for sh in root.find(_t("sheets")).findall(_t("sheet")):
        rid = sh.get("{%s}id" % NS_REL)
        typ, target = rels[rid]
        if not typ.endswith("/worksheet"):
            continue
        sheets.append((sh.get("name"), z.read(target)))
    sst = None
    for typ, target in rels.values():
        if typ.endswith("/sharedStrings") and target in names:
            sst = z.read(target)
    if sst is None and "xl/sharedStrings.xml" in names:
        sst = z.read("xl/sharedStrings.xml")
    styles = z.read("xl/styles.xml") if "xl/styles.xml" in names else None
    dates, tds = date_style_ids(styles)
    return {"sheets": sheets, "sst": sst, "date_styles": dates, "td_styles": tds,
            "date1904": date1904}


# ---------------------------------------------------------------- shared strings
def _text_content(node) -> str:
    """openpyxl/cell/text.py Text.content for an <si> or <is> element."""
    snippets = []
    t = node.find(_t("t"))
    if t is not None and t.text is not None:
        snippets.append(t.text)
    for r in node.findall(_t("r")):
        rt = r.find(_t("t"))
        if rt is not None and rt.text is not None:
            snippets.append(rt.text)
    return "".join(snippets)


def read_string_table(sst_xml: bytes | None) -> "list[str]":
    strings = []
    if not sst_xml:
        return strings
    for _, node in ET.iterparse(io.BytesIO(sst_xml)):
        if node.tag == _t("si"):
            strings.append(_text_content(node).replace("x005F_", ""))
            node.clear()
    return strings


# ---------------------------------------------------------------- worksheet
_COORD_RE = re.compile(r"^[$]?([A-Za-z]{1,3})[$]?(\d+)$")
WINDOWS_EPOCH = datetime.datetime(1899, 12, 30)
MAC_EPOCH = datetime.datetime(1904, 1, 1)


def _col_index(letters: str) -> int:
    n = 0
    for ch in letters.upper():
        n = n * 26 + (ord(ch) - 64)
    return n


def _cast_number(value: str):
    if "." in value or "E" in value or "e" in value:
        return float(value)
    return int(value)


def from_excel(value, epoch=WINDOWS_EPOCH, timedelta=False):
    if value is None:
        return None
    if timedelta:
        td = datetime.timedelta(days=value)
        if td.microseconds:
            td = datetime.timedelta(seconds=td.total_seconds() // 1,
                                    microseconds=round(td.microseconds, -3))
        return td
    day, fraction = divmod(value, 1)
    diff = datetime.timedelta(milliseconds=round(fraction * 86400 * 1000))
    if 0 <= value < 1 and diff.days == 0:
        mins, seconds = divmod(diff.seconds, 60)
        hours, mins = divmod(mins, 60)
        return datetime.time(hours, mins, seconds, diff.microseconds)
    if 0 < value < 60 and epoch == WINDOWS_EPOCH:
        day += 1
    return epoch + datetime.timedelta(days=day) + diff


class _Cell:
    __slots__ = ("column", "value", "data_type")

    def __init__(self, column, value, data_type):
        self.column, self.value, self.data_type = column, value, data_type


def _parse_rows(sheet_xml: bytes, shared, date_styles, td_styles, epoch):
    """WorkSheetParser.parse/parse_row/parse_cell with data_only=True."""
    row_counter = 0
    for _, el in ET.iterparse(io.BytesIO(sheet_xml)):
        if el.tag != _t("row"):
            continue
        r = el.get("r")
        if r is not None:
            try:
                row_counter = int(r)
            except ValueError:
                val = float(r)
                if not val.is_integer():
                    raise ValueError(f"{r} is not a valid row number")
                row_counter = int(val)
        else:
            row_counter += 1
        col_counter = 0
        cells = []
        for c in el:
            if c.tag != _t("c"):
                # openpyxl runs parse_cell on every child of <row>; non-<c> children are not
                # produced by any writer we know of; ignore them.
                continue
            data_type = c.get("t", "n")
            coordinate = c.get("r")
            style_id = c.get("s", 0)
            if style_id:
                style_id = int(style_id)
            if data_type == "inlineStr":
                value = None
            else:
                value = c.findtext(_t("v"), None) or None
            if coordinate:
                m = _COORD_RE.match(coordinate)
                if not m:
                    raise ValueError(f"{coordinate} is not a valid coordinate or range")
                col_counter = _col_index(m.group(1))
            else:
                col_counter += 1
            column = col_counter
            if value is not None:
                if data_type == "n":
                    value = _cast_number(value)
                    if style_id in date_styles:
                        data_type = "d"
                        try:
                            value = from_excel(value, epoch, timedelta=style_id in td_styles)
                        except (OverflowError, ValueError):
                            data_type = "e"
                            value = "#VALUE!"
                elif data_type == "s":
                    value = shared[int(value)]
                elif data_type == "b":
                    value = bool(int(value))
                elif data_type == "str":
                    data_type = "s"
                elif data_type == "d":
                    value = datetime.datetime.fromisoformat(value)
            elif data_type == "inlineStr":
                child = c.find(_t("is"))
                if child is not None:
                    data_type = "s"
                    value = _text_content(child)
            cells.append(_Cell(column, value, data_type))
        el.clear()
        yield row_counter, cells


def _cells_by_row(sheet_xml, shared, date_styles, td_styles, epoch):
    """ReadOnlyWorksheet._cells_by_row after reset_dimensions(): max_row/max_col unknown."""
    counter = 1
    for idx, row in _parse_rows(sheet_xml, shared, date_styles, td_styles, epoch):
        for _ in range(counter, idx):  # some rows are missing
            counter += 1
            yield ()
        if counter <= idx:
            counter += 1
            yield _get_row(row)


_EMPTY = _Cell(0, None, "n")


def _get_row(row):
    if not row:
        return ()
    max_col = row[-1].column
    new_row = [_EMPTY] * max_col
    for cell in row:
        if 1 <= cell.column <= max_col:
            new_row[cell.column - 1] = cell
    return tuple(new_row)


def _convert_cell(cell):
    """pandas/io/excel/_openpyxl.py:596-612."""
    if cell.value is None:
        return ""
    elif cell.data_type == "e":
        return np.nan
    elif cell.data_type == "n":
        val = int(cell.value)
        if val == cell.value:
            return val
        return float(cell.value)
    return cell.value


def get_sheet_data(rows, file_rows_needed):
    """pandas/io/excel/_openpyxl.py:614-646."""
    data = []
    last_row_with_data = -1
    for row_number, row in enumerate(rows):
        converted_row = [_convert_cell(cell) for cell in row]
        while converted_row and converted_row[-1] == "":
            converted_row.pop()
        if converted_row:
            last_row_with_data = row_number
        data.append(converted_row)
        if file_rows_needed is not None and len(data) >= file_rows_needed:
            break
    data = data[: last_row_with_data + 1]
    if len(data) > 0:
        max_width = max(len(r) for r in data)
        if min(len(r) for r in data) < max_width:
            data = [r + (max_width - len(r)) * [""] for r in data]
    return data


def sheet_table(sheet_xml, shared, date_styles=(), td_styles=(), date1904=False, nrows=None):
    epoch = MAC_EPOCH if date1904 else WINDOWS_EPOCH
    file_rows_needed = None if nrows is None else 1 + nrows  # _calc_rows, header=0
    return get_sheet_data(
        _cells_by_row(sheet_xml, shared, set(date_styles), set(td_styles), epoch),
        file_rows_needed)


def table_to_frame(data, nrows):
    """pandas/io/excel/_base.py:764-766 and :913-937 (header=0, everything else default)."""
    if not data:
        return pd.DataFrame()
    try:
        parser = TextParser(data, header=0, skip_blank_lines=False, nrows=nrows,
                            names=None, index_col=None, has_index_names=False, dtype=None,
                            true_values=None, false_values=None, skiprows=None, na_values=None,
                            parse_dates=False, date_format=None, thousands=None, decimal=".",
                            comment=None, skipfooter=0, usecols=None)
        return parser.read(nrows=nrows)
    except pd.errors.EmptyDataError:
        return pd.DataFrame()


def read_parts(parts: dict, nrows: int):
    shared = read_string_table(parts.get("sst"))
    out = []
    for name, xml in parts["sheets"]:
        data = sheet_table(xml, shared, parts.get("date_styles", ()), parts.get("td_styles", ()),
                           parts.get("date1904", False), nrows)
        out.append((name, table_to_frame(data, nrows)))
    return out


def read_excel_all_sheets(data: bytes, nrows: int):
    return read_parts(open_workbook(data), nrows)
