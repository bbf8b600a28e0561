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
import struct
import zlib
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


class _ZipDir:
    """The central directory of a zip container, read with `struct` (the stdlib `zipfile` builds a ZipInfo object
    per member and a file object per read: most of the host time of a batch of small workbooks).  Same verdicts
    as zipfile for what an .xlsx is made of; zip64, multi-disk and encrypted archives are refused."""
    __slots__ = ("data", "entries")
    _EOCD = struct.Struct("<4s4H2LH")
    _CEN = struct.Struct("<4s6H3L5H2L")

    def __init__(self, data: bytes):
        self.data = data
        n = len(data)
        pos = data.rfind(b"PK\x05\x06", max(0, n - 65557))
        if pos < 0 or pos + 22 > n:
            raise zipfile.BadZipFile("File is not a zip file")
        _, disk, cd_disk, n_here, n_total, cd_size, cd_off, _ = self._EOCD.unpack_from(data, pos)
        if disk or cd_disk or n_here != n_total or n_total == 0xFFFF or cd_off == 0xFFFFFFFF:
            raise zipfile.BadZipFile("zip64 / multi-disk archives are not supported")
        if cd_off + cd_size > pos:
            raise zipfile.BadZipFile("Bad central directory")
        self.entries = ent = {}
        p = cd_off
        for _ in range(n_total):
            if data[p:p + 4] != b"PK\x01\x02":
                raise zipfile.BadZipFile("Bad magic number for central directory")
            (_, _, _, flags, method, _, _, crc, csize, usize, fn, ex, cm, _, _, _, hoff) = self._CEN.unpack_from(data, p)
            raw = data[p + 46:p + 46 + fn]
            name = raw.decode("utf-8" if flags & 0x800 else "cp437")
            ent[name] = (flags, method, crc, csize, usize, hoff)
            p += 46 + fn + ex + cm

    def member(self, name: str) -> Member:
        flags, method, crc, csize, usize, ho = self.entries[name]
        if flags & 0x1:
            raise zipfile.BadZipFile("encrypted member %r" % name)
        data = self.data
        if data[ho:ho + 4] != b"PK\x03\x04":
            raise zipfile.BadZipFile("bad local header of %r" % name)
        fnlen = int.from_bytes(data[ho + 26:ho + 28], "little")
        exlen = int.from_bytes(data[ho + 28:ho + 30], "little")
        start = ho + 30 + fnlen + exlen
        return Member(memoryview(data)[start:start + csize], method, usize, crc)

    def read(self, name: str) -> bytes:
        """A small member, inflated here (workbook.xml, its rels, styles.xml)."""
        m = self.member(name)
        if m.method == 0:
            out = bytes(m.raw)
        elif m.method == 8:
            out = zlib.decompressobj(-15).decompress(m.raw)
        else:
            raise NotImplementedError("That compression method is not supported")
        if len(out) != m.size or (zlib.crc32(out) & 0xFFFFFFFF) != m.crc:
            raise zipfile.BadZipFile("Bad CRC-32 for file %r" % name)
        return out


# workbooks written by one program share their workbook.xml / rels / styles.xml byte for byte: what those parts
# say is kept per distinct content (bounded), so a batch parses each of them once
_plan_cache: dict = {}
_style_cache: dict = {}
_CACHE_MAX = 512


def _workbook_plan(rels_xml, workbook_xml):
    """(date1904, [(sheet name, part path)], [sharedStrings part paths named by the rels])."""
    key = (rels_xml, workbook_xml)
    hit = _plan_cache.get(key)
    if hit is not None:
        return hit
    rels = {}
    if rels_xml is not None:
        for r in ET.fromstring(rels_xml).findall("{%s}Relationship" % _NS_PKG):
            t = r.get("Target")
            t = t[1:] if t.startswith("/") else posixpath.normpath(posixpath.join("xl", t))
            rels[r.get("Id")] = (r.get("Type"), t)
    root = ET.fromstring(workbook_xml)
    pr = root.find("{%s}workbookPr" % _NS)
    date1904 = pr is not None and pr.get("date1904") in ("1", "true")
    sheets = []
    for sh in root.find("{%s}sheets" % _NS).findall("{%s}sheet" % _NS):
        typ, target = rels[sh.get("{%s}id" % _NS_REL)]
        if typ.endswith("/worksheet"):
            sheets.append((sh.get("name"), target))
    sst_targets = [target for typ, target in rels.values() if typ.endswith("/sharedStrings")]
    plan = (date1904, sheets, sst_targets)
    if len(_plan_cache) >= _CACHE_MAX:
        _plan_cache.clear()
    _plan_cache[key] = plan
    return plan


def _date_styles_cached(styles_xml):
    hit = _style_cache.get(styles_xml)
    if hit is None:
        hit = _date_style_ids(styles_xml)
        if len(_style_cache) >= _CACHE_MAX:
            _style_cache.clear()
        _style_cache[styles_xml] = hit
    return hit


def scan_workbook(data: bytes) -> Workbook:
    """open_workbook without inflating the big parts: worksheet and sharedStrings parts come back as
    `Member`s (raw zip data for the GPU); only workbook.xml, its rels and styles.xml — a few KB that
    decide names, order and date styles — are inflated here."""
    z = _ZipDir(data)
    names = z.entries
    rels_xml = z.read("xl/_rels/workbook.xml.rels") if "xl/_rels/workbook.xml.rels" in names else None
    date1904, plan, sst_targets = _workbook_plan(rels_xml, z.read("xl/workbook.xml"))
    sheet_names = [nm for nm, _ in plan]
    sheets = [z.member(target) for _, target in plan]
    sst = None
    for target in sst_targets:
        if target in names:
            sst = z.member(target)
    if sst is None and "xl/sharedStrings.xml" in names:
        sst = z.member("xl/sharedStrings.xml")
    styles = z.read("xl/styles.xml") if "xl/styles.xml" in names else None
    return Workbook(sheet_names, sheets, sst, _date_styles_cached(styles), date1904)


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
        want = []
        for d, (b, k) in enumerate(owner):
            if k is None or b in failed:
                continue
            st = int(res.status[d])
            if st == _lib.DOC_EMPTY_SHEET:
                continue  # `if df.empty: continue` (app/file_parser.py:228-229)
            if st != _lib.DOC_OK:
                failed[b] = EngineError(f"sheet {books[b].sheet_names[k]!r}: construct not decoded on the GPU path")
                continue
            want.append(d)
        for d, text in zip(want, res.doc_strs(want) if want else []):   # all sheet texts built in one go (library threads)
            b, k = owner[d]
            if b not in failed:
                parts[b].append(f"=== Sheet: {books[b].sheet_names[k]} ===\n" + text)
    return [failed[b] if b in failed else ("\n\n".join(p) if p else _EMPTY) for b, p in enumerate(parts)]
