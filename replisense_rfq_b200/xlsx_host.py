"""Host side of the .xlsx path: zip container + workbook plumbing (SURVEY.md §8 a7.1 — stays on
the host), then batches of inflated worksheet / sharedStrings parts to the GPU, then the
reference's own assembly of the per-sheet texts (app/file_parser.py:227-252).

What stays here mirrors openpyxl's ExcelReader (openpyxl/reader/excel.py): workbook.xml sheet
names + order, workbook.xml.rels targets, the sharedStrings part, and which cellXfs of styles.xml
carry a date number format (openpyxl/styles/numbers.py is_date_format) — O(parts) work on small
XML files.  Cell decoding, type inference and formatting all happen on the device.
"""
from __future__ import annotations

import io
import posixpath
import re
import zipfile
import xml.etree.ElementTree as ET

from . import _lib

_NS = "http://schemas.openxmlformats.org/spreadsheetml/2006/main"
_NS_REL = "http://schemas.openxmlformats.org/officeDocument/2006/relationships"
_NS_PKG = "http://schemas.openxmlformats.org/package/2006/relationships"
_EMPTY = "Excel file appears to be empty or unreadable"  # app/file_parser.py:250

_BUILTIN_DATE_IDS = set(range(14, 23)) | {45, 46, 47}
_STRIP = re.compile(r"\"[^\"]*\"|\[(?!hh?\]|mm?\]|ss?\])[^\]]*\]")


def _is_date_format(code) -> bool:
    if code is None:
        return False
    code = _STRIP.sub("", code.split(";")[0])
    return re.search(r"(?<![_\\])[dmhysDMHYS]", code) is not None


_TIMEDELTA = re.compile(r"\[hh?\](:mm(:ss(\.0*)?)?)?|\[mm?\](:ss(\.0*)?)?|\[ss?\](\.0*)?", re.I)
_TD_BIT = 0x80000000  # rfq_sheets_run: a date style whose format is a timedelta format


def _is_timedelta_format(code) -> bool:
    return code is not None and _TIMEDELTA.search(code.split(";")[0]) is not None


def _date_style_ids(styles_xml) -> "set[int]":
    """cellXfs indices with a date number format (openpyxl styles/stylesheet.py _normalise_numbers);
    timedelta formats carry _TD_BIT."""
    if not styles_xml:
        return set()
    root = ET.fromstring(styles_xml)
    custom = {}
    nf = root.find("{%s}numFmts" % _NS)
    if nf is not None:
        for el in nf.findall("{%s}numFmt" % _NS):
            custom[int(el.get("numFmtId"))] = el.get("formatCode")
    out = set()
    xfs = root.find("{%s}cellXfs" % _NS)
    if xfs is not None:
        for idx, xf in enumerate(xfs.findall("{%s}xf" % _NS)):
            fid = int(xf.get("numFmtId", "0"))
            if fid in custom:
                if _is_date_format(custom[fid]):
                    out.add(idx | (_TD_BIT if _is_timedelta_format(custom[fid]) else 0))
            elif fid in _BUILTIN_DATE_IDS:
                out.add(idx | (_TD_BIT if fid == 46 else 0))  # 46 = "[h]:mm:ss"
    return out


class Workbook:
    __slots__ = ("sheet_names", "sheets", "sst", "date_styles", "date1904")

    def __init__(self, sheet_names, sheets, sst, date_styles, date1904=False):
        self.sheet_names, self.sheets, self.sst, self.date_styles = sheet_names, sheets, sst, date_styles
        self.date1904 = date1904


class Member:
    """One zip member as it lies in the container: raw (still compressed) bytes + what the zip directory
    says about it.  The GPU inflates it (rfq_inflate_run)."""
    __slots__ = ("raw", "method", "size", "crc")

    def __init__(self, raw, method, size, crc):
        self.raw, self.method, self.size, self.crc = raw, method, size, crc


def _member(z: zipfile.ZipFile, data: bytes, name: str) -> Member:
    zi = z.getinfo(name)
    if zi.flag_bits & 0x1:
        raise zipfile.BadZipFile("encrypted member %r" % name)
    ho = zi.header_offset
    if data[ho:ho + 4] != b"PK\x03\x04":
        raise zipfile.BadZipFile("bad local header of %r" % name)
    fnlen = int.from_bytes(data[ho + 26:ho + 28], "little")
    exlen = int.from_bytes(data[ho + 28:ho + 30], "little")
    start = ho + 30 + fnlen + exlen
    return Member(memoryview(data)[start:start + zi.compress_size], zi.compress_type, zi.file_size, zi.CRC)


def scan_workbook(data: bytes) -> Workbook:
    """open_workbook without inflating the big parts: worksheet and sharedStrings parts come back as
    `Member`s (raw zip data for the GPU); only workbook.xml, its rels and styles.xml — a few KB that
    decide names, order and date styles — are inflated here."""
    z = zipfile.ZipFile(io.BytesIO(data))
    names = set(z.namelist())
    rels = {}
    if "xl/_rels/workbook.xml.rels" in names:
        for r in ET.fromstring(z.read("xl/_rels/workbook.xml.rels")).findall("{%s}Relationship" % _NS_PKG):
            t = r.get("Target")
            t = t[1:] if t.startswith("/") else posixpath.normpath(posixpath.join("xl", t))
            rels[r.get("Id")] = (r.get("Type"), t)
    root = ET.fromstring(z.read("xl/workbook.xml"))
    pr = root.find("{%s}workbookPr" % _NS)
    date1904 = pr is not None and pr.get("date1904") in ("1", "true")
    sheet_names, sheets = [], []
    for sh in root.find("{%s}sheets" % _NS).findall("{%s}sheet" % _NS):
        typ, target = rels[sh.get("{%s}id" % _NS_REL)]
        if not typ.endswith("/worksheet"):
            continue
        sheet_names.append(sh.get("name"))
        sheets.append(_member(z, data, target))
    sst = None
    for typ, target in rels.values():
        if typ.endswith("/sharedStrings") and target in names:
            sst = _member(z, data, target)
    if sst is None and "xl/sharedStrings.xml" in names:
        sst = _member(z, data, "xl/sharedStrings.xml")
    styles = z.read("xl/styles.xml") if "xl/styles.xml" in names else None
    return Workbook(sheet_names, sheets, sst, _date_style_ids(styles), date1904)


def open_workbook(data: bytes) -> Workbook:
    z = zipfile.ZipFile(io.BytesIO(data))
    names = set(z.namelist())
    rels = {}
    if "xl/_rels/workbook.xml.rels" in names:
        for r in ET.fromstring(z.read("xl/_rels/workbook.xml.rels")).findall("{%s}Relationship" % _NS_PKG):
            t = r.get("Target")
            t = t[1:] if t.startswith("/") else posixpath.normpath(posixpath.join("xl", t))
            rels[r.get("Id")] = (r.get("Type"), t)
    root = ET.fromstring(z.read("xl/workbook.xml"))
    pr = root.find("{%s}workbookPr" % _NS)
    date1904 = pr is not None and pr.get("date1904") in ("1", "true")   # openpyxl WorkbookParser.epoch
    sheet_names, sheets = [], []
    for sh in root.find("{%s}sheets" % _NS).findall("{%s}sheet" % _NS):
        typ, target = rels[sh.get("{%s}id" % _NS_REL)]
        if not typ.endswith("/worksheet"):
            continue
        sheet_names.append(sh.get("name"))
        sheets.append(z.read(target))
    sst = None
    for typ, target in rels.values():
        if typ.endswith("/sharedStrings") and target in names:
            sst = z.read(target)
    if sst is None and "xl/sharedStrings.xml" in names:
        sst = z.read("xl/sharedStrings.xml")
    styles = z.read("xl/styles.xml") if "xl/styles.xml" in names else None
    return Workbook(sheet_names, sheets, sst, _date_style_ids(styles), date1904)


def parse_workbooks(engine, books, max_rows: int = 1000) -> "list":
    """raw_text of every workbook (one GPU batch for all their parts); an EngineError instance in
    place of the text where a sheet uses a construct that is not decoded on the GPU path.  Workbooks
    from scan_workbook carry their parts still compressed: the batch is inflated on the device."""
    docs, kind, sst_of, dates, owner = [], [], [], [], []
    deflated = any(isinstance(wb.sst, Member) or any(isinstance(x, Member) for x in wb.sheets) for wb in books)
    for b, wb in enumerate(books):
        sst_idx = -1
        if wb.sst is not None:
            sst_idx = len(docs)
            docs.append(wb.sst); kind.append(1); sst_of.append(-1); dates.append(()); owner.append((b, None))
        for k, xml in enumerate(wb.sheets):
            docs.append(xml); kind.append(2 if wb.date1904 else 0); sst_of.append(sst_idx); dates.append(wb.date_styles)
            owner.append((b, k))
    from .engine import EngineError
    parts = [[] for _ in books]
    failed = {}
    with engine.lock:   # the pinned output is ours until the texts have been built
        if docs and deflated:
            res, inflate_status = engine.sheets_batch_deflated(docs, kind, sst_of, max_rows, dates)
            for d, (b, k) in enumerate(owner):
                if int(inflate_status[d]) != _lib.DOC_OK:
                    failed[b] = EngineError("zip member is corrupt (bad DEFLATE stream, size or CRC-32)")
        else:
            res = engine.sheets_batch(docs, kind, sst_of, max_rows, dates) if docs else None
        for d, (b, k) in enumerate(owner):
            if k is None or b in failed:
                continue
            st = int(res.status[d])
            if st == _lib.DOC_EMPTY_SHEET:
                continue  # `if df.empty: continue` (app/file_parser.py:228-229)
            if st != _lib.DOC_OK:
                failed[b] = EngineError(f"sheet {books[b].sheet_names[k]!r}: construct not decoded on the GPU path")
                continue
            parts[b].append(f"=== Sheet: {books[b].sheet_names[k]} ===\n" + res.doc_str(d))
    return [failed[b] if b in failed else ("\n\n".join(p) if p else _EMPTY) for b, p in enumerate(parts)]
