"""Generate the committed golden fixtures under tests/golden/.

Run in the AUTHORING container only (needs /root/reference):
    python tests/golden/make_goldens.py

Produces
  cache_rows.json     raw_text strings pinned by the reference's own `.cache/42/cache.db`
                      (rows 12, 14 = xlsx goldens; 13 = PDF framing; 2,3,5,6,7 = e-mail bodies).
                      Extracted WITHOUT unpickling (pickletools.genops over the key blob).
  attachment_parts.json  config C1: the inflated sheet XML / sharedStrings / names of
                      tests/assets/attachment.xlsx (== uploads/'email_4 attachment.xlsx'), base64.
  ref_csv.json / ref_txt.json  inputs (base64) + outputs of the reference's UNMODIFIED
                      FileParser._parse_csv_sync / _parse_txt_async, imported from /root/reference
                      with a stub `docx` module.
Nothing under /root/reference is copied: only its outputs and the binary test asset's inflated
XML parts (test data, not source).
"""
from __future__ import annotations

import asyncio
import base64
import json
import os
import pickletools
import sqlite3
import sys
import tempfile
import types
import zipfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path("/root/reference")
sys.path.insert(0, str(HERE.parent.parent))


def b64(b: bytes) -> str:
    return base64.b64encode(b).decode("ascii")


def cache_rows():
    con = sqlite3.connect(f"file:{REF}/.cache/42/cache.db?mode=ro&immutable=1", uri=True)
    out = {}
    for rid in (2, 3, 5, 6, 7, 12, 13, 14):
        key = con.execute("select key from Cache where rowid=?", (rid,)).fetchone()[0]
        prompt = max((a for _, a, _ in pickletools.genops(bytes(key)) if isinstance(a, str)),
                     key=len)
        out[str(rid)] = prompt.split('\n"""\n', 1)[1].rsplit('\n"""', 1)[0]
    return out


def attachment_parts():
    z = zipfile.ZipFile(REF / "tests/assets/attachment.xlsx")
    return {
        "sheet_names": ["Sheet1"],
        "sheets": [b64(z.read("xl/worksheets/sheet1.xml"))],
        "sst": b64(z.read("xl/sharedStrings.xml")),
        "styles": b64(z.read("xl/styles.xml")),
        "workbook": b64(z.read("xl/workbook.xml")),
        "rels": b64(z.read("xl/_rels/workbook.xml.rels")),
    }


def import_reference():
    sys.path.insert(0, str(REF))
    sys.modules.setdefault("docx", types.ModuleType("docx"))
    import logging
    logging.disable(logging.CRITICAL)
    from app.file_parser import FileParser, FileParsingError  # noqa: E402
    return FileParser, FileParsingError


def main():
    from tests import corpus  # seeded generators shared with the parity tests

    FileParser, FileParsingError = import_reference()
    tmp = Path(tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None))

    (HERE / "cache_rows.json").write_text(json.dumps(cache_rows(), ensure_ascii=False, indent=0))
    (HERE / "attachment_parts.json").write_text(json.dumps(attachment_parts()))

    # --- CSV: hand-written edge cases + seeded fuzz, through the reference itself
    csv_cases = []
    for name, data, nrows in corpus.csv_golden_cases():
        p = tmp / "case.csv"
        p.write_bytes(data)
        fp = FileParser(max_file_size_mb=4096, max_excel_rows=nrows)
        try:
            out = {"text": fp._parse_csv_sync(p)}
        except FileParsingError as e:
            out = {"error": str(e)}
        csv_cases.append({"name": name, "nrows": nrows, "data": b64(data), **out})
    (HERE / "ref_csv.json").write_text(json.dumps(csv_cases, ensure_ascii=False, indent=0))

    # --- TXT
    txt_cases = []
    fp = FileParser()
    for name, data in corpus.txt_golden_cases():
        p = tmp / "case.txt"
        p.write_bytes(data)
        txt_cases.append({"name": name, "data": b64(data),
                          "text": asyncio.run(fp._parse_txt_async(p))})
    (HERE / "ref_txt.json").write_text(json.dumps(txt_cases, ensure_ascii=False, indent=0))
    print("csv cases", len(csv_cases), "txt cases", len(txt_cases))


if __name__ == "__main__":
    main()
