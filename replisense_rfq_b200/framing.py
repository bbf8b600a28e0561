"""Text framing on the GPU: what `FileParser._parse_pdf_sync` does with the pages the
PDF engine extracted (app/file_parser.py:166-212) — page headers around the stripped page text, table
blocks (" | "-joined rows), image notes, "\\n\\n" between the parts, a fixed text for a document with
nothing extractable.  The extraction itself (MuPDF: layout analysis, table detection) is out of scope;
this takes its output for many documents and assembles every raw_text in one batch (rfq_frame_run)."""
from __future__ import annotations

from typing import Any, Optional, Sequence

import numpy as np

from . import _lib

PDF_EMPTY = "PDF appears to be empty or contains no extractable text"   # app/file_parser.py:210


class Page:
    """One page as the PDF engine reports it: get_text(), find_tables() -> extract() rows, get_images()."""
    __slots__ = ("text", "tables", "n_images")

    def __init__(self, text: str, tables: Optional[Sequence[Sequence[Sequence[Any]]]] = None, n_images: int = 0):
        self.text, self.tables, self.n_images = text, list(tables or []), n_images


def _table_text(table_num: int, page_num: int, table_data) -> str:
    """app/file_parser.py:188-193, the part that is pure formatting of already-extracted cells."""
    t = f"\n=== Table {table_num} on Page {page_num} ===\n"
    for row in table_data:
        if row and any(cell for cell in row if cell):
            t += " | ".join(str(cell) if cell else "" for cell in row) + "\n"
    return t


def frame_pdf_texts(engine, docs: "Sequence[Sequence[Page]]") -> "list[str]":
    """raw_text of every document (a sequence of Pages), as _parse_pdf_sync builds it.  The page texts —
    the bulk of the bytes — are stripped, tested for emptiness and laid out on the device; headers and
    table blocks are small literal pieces."""
    pieces, strip = [], []
    part_kind, part_first = [], [0]
    doc_first = [0]
    for pages in docs:
        for k, pg in enumerate(pages):
            pieces += [f"=== Page {k + 1} ===\n".encode("utf-8"), pg.text.encode("utf-8")]
            strip += [0, 1]
            part_kind.append(_lib.PART_IF_TEXT); part_first.append(len(pieces))
            for t, table in enumerate(pg.tables, 1):
                if table:                                   # `if table_data:`
                    pieces.append(_table_text(t, k + 1, table).encode("utf-8")); strip.append(0)
                    part_kind.append(_lib.PART_ALWAYS); part_first.append(len(pieces))
            if pg.n_images:
                pieces.append(f"\n=== Images on Page {k + 1} ===\nFound {pg.n_images} image(s)".encode("utf-8")); strip.append(0)
                part_kind.append(_lib.PART_ALWAYS); part_first.append(len(pieces))
        pieces.append(PDF_EMPTY.encode("utf-8")); strip.append(0)
        part_kind.append(_lib.PART_FALLBACK); part_first.append(len(pieces))
        doc_first.append(len(part_kind))
    return engine.frame_many(pieces, strip, part_kind, part_first, doc_first)
