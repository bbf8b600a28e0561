"""Host-side mirror of the reference's `FileParser` (app/file_parser.py:25-375) for the formats
on the north-star path — .txt, .csv, .xlsx, .json — with every byte of parsing done by the CUDA
library behind include/rfq_parse.h.  Same constructor arguments, method names, result envelope
and error behaviour as the reference, plus the batched `parse_many` entry point.

Out of scope here: .pdf, .docx, .xls — they are reported as unsupported
file types rather than routed to a CPU implementation.
"""
from __future__ import annotations

import asyncio
import hashlib
import logging
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import Any, Dict, Optional, Sequence

from . import _lib
from .engine import Engine, EngineError, default_pool

logger = logging.getLogger(__name__)

HAS_PYMUPDF = False  # PDF extraction (MuPDF layout engine) is not part of this build

_SUPPORTED = {".txt", ".xlsx", ".csv", ".json"}

_CSV_FAIL = "Unable to parse CSV with any supported encoding/separator combination"
_EXCEL_EMPTY = "Excel file appears to be empty or unreadable"


class FileParsingError(Exception):
    """Same role as app/file_parser.py:25-27."""


class FileParser:
    """Drop-in for app/file_parser.py:29 (`FileParser(max_file_size_mb=10, max_excel_rows=1000)`)."""

    def __init__(self, max_file_size_mb: int = 10, max_excel_rows: int = 1000,
                 device: int = 0, engine: Optional[Engine] = None):
        self.max_file_size = max_file_size_mb * 1024 * 1024
        self.max_excel_rows = max_excel_rows
        self.thread_pool = ThreadPoolExecutor(max_workers=4)  # app/file_parser.py:33
        self.supported_extensions = set(_SUPPORTED)
        self._device = device
        self._engine = engine

    @property
    def engine(self) -> Engine:
        if self._engine is not None:
            return self._engine
        # two streams per device: concurrent requests of the shared parser (app/main.py:47) overlap
        # one call's copies with the other's kernels.  Raises without a CUDA device.
        return default_pool(self._device).pick()

    # ------------------------------------------------------------------ envelope (:44-99)
    async def parse_file_async(self, file_path: str) -> Dict[str, Any]:
        file_path = Path(file_path)
        await self._validate_file_async(file_path)
        file_hash = await self._get_file_hash_async(file_path)
        cached = self._get_cached_result(file_hash)
        if cached:
            return cached
        suffix = file_path.suffix.lower()
        try:
            if suffix == ".txt":
                text = await self._parse_txt_async(file_path)
            elif suffix == ".xlsx":
                text = await self._parse_excel_async(file_path)
            elif suffix == ".csv":
                text = await self._parse_csv_async(file_path)
            elif suffix == ".json":
                text = await self._parse_json_async(file_path)
            else:
                raise FileParsingError(f"Unsupported file format: {file_path.suffix}")
            result = {
                "raw_text": text,
                "source_file": str(file_path.name),
                "file_size": file_path.stat().st_size,
                "file_hash": file_hash,
                "parsing_method": f"async_{suffix[1:]}",
            }
            self._cache_result(file_hash, result)
            return result
        except Exception as e:  # the reference wraps everything from the dispatch (:93-95)
            raise FileParsingError(f"Failed to parse {file_path.name}: {str(e)}")

    def parse_file(self, file_path: str) -> Dict[str, Any]:
        return asyncio.run(self.parse_file_async(file_path))

    async def _validate_file_async(self, file_path: Path) -> None:
        """app/file_parser.py:101-120: FileNotFoundError escapes un-wrapped."""
        if not file_path.exists():
            raise FileNotFoundError(f"File not found: {file_path}")
        file_size = file_path.stat().st_size
        if file_size > self.max_file_size:
            raise FileParsingError(
                f"File too large: {file_size / 1024 / 1024:.1f}MB "
                f"(max: {self.max_file_size / 1024 / 1024}MB)")
        if file_path.suffix.lower() not in self.supported_extensions:
            raise FileParsingError(f"Unsupported file type: {file_path.suffix}")
        try:
            with open(file_path, "rb") as f:
                f.read(1)
        except PermissionError:
            raise FileParsingError(f"Permission denied reading file: {file_path}")

    async def _get_file_hash_async(self, file_path: Path) -> str:
        """MD5 stays on the host (app/file_parser.py:122-130)."""
        h = hashlib.md5()
        with open(file_path, "rb") as f:
            for chunk in iter(lambda: f.read(1 << 20), b""):
                h.update(chunk)
        return h.hexdigest()

    def _get_cached_result(self, file_hash: str) -> Optional[Dict[str, Any]]:
        return None  # the reference's cache hooks are no-ops (:132-141)

    def _cache_result(self, file_hash: str, result: Dict[str, Any]) -> None:
        pass

    # ------------------------------------------------------------------ text (:143-159)
    async def _parse_txt_async(self, file_path: Path) -> str:
        return self.engine.parse_text_file(file_path)

    # ------------------------------------------------------------------ csv (:288-313)
    async def _parse_csv_async(self, file_path: Path) -> str:
        return await asyncio.get_event_loop().run_in_executor(
            self.thread_pool, self._parse_csv_sync, file_path)

    def _parse_csv_sync(self, file_path: Path) -> str:
        try:
            text = self.engine.parse_csv_file(file_path, self.max_excel_rows)   # read straight into pinned memory
            if text is None:
                raise FileParsingError(_CSV_FAIL)
            return text
        except Exception as e:
            raise FileParsingError(f"CSV parsing failed: {str(e)}")

    # ------------------------------------------------------------------ json (:315-331)
    @staticmethod
    def _json_result(t):
        """A slot of Engine.parse_json_many as the reference's result: the text, or its FileParsingError."""
        if isinstance(t, EngineError):
            if getattr(t, "json_error", None) is not None:
                return FileParsingError(f"Invalid JSON format: {t.json_error}")
            return FileParsingError(f"JSON parsing failed: {t}")
        return t

    async def _parse_json_async(self, file_path: Path) -> str:
        t = self._json_result(self.engine.parse_json_many([Path(file_path).read_bytes()])[0])
        if isinstance(t, Exception):
            raise t
        return t

    # ------------------------------------------------------------------ excel (:214-252)
    async def _parse_excel_async(self, file_path: Path) -> str:
        return await asyncio.get_event_loop().run_in_executor(
            self.thread_pool, self._parse_excel_sync, file_path)

    def _parse_excel_sync(self, file_path: Path) -> str:
        try:
            from . import xlsx_host
            parts = xlsx_host.scan_workbook(Path(file_path).read_bytes())   # big parts stay compressed: inflated on the GPU
            text = xlsx_host.parse_workbooks(self.engine, [parts], self.max_excel_rows)[0]
            if isinstance(text, Exception):
                raise text
            return text
        except Exception as e:
            raise FileParsingError(f"Excel parsing failed: {str(e)}")

    # ------------------------------------------------------------------ batched entry point
    def parse_many(self, docs: Sequence[bytes], kinds: Sequence[str]) -> "list[str | FileParsingError]":
        """Many documents in one go. `kinds[i]` in {'txt', 'csv', 'xlsx', 'json'}; returns raw_text or the
        FileParsingError the reference would have raised for that document, in input order.  Text bodies
        and CSV files travel in ONE packed arena and one upload (rfq_mixed_parse).  The worksheet /
        sharedStrings parts of all .xlsx containers travel as they lie in the zip — still compressed — and
        are inflated on the device (rfq_inflate_run) before the sheet path runs over them."""
        from . import xlsx_host
        n = len(docs)
        out: "list[Any]" = [None] * n
        flat, fkind, owner = [], [], []
        books, book_idx = [], []
        jdocs, jidx = [], []
        for i, (doc, k) in enumerate(zip(docs, kinds)):
            k = k.lower().lstrip(".")
            if k == "txt":
                flat.append(doc); fkind.append(_lib.KIND_TEXT); owner.append(i)
            elif k == "csv":
                flat.append(doc); fkind.append(_lib.KIND_CSV); owner.append(i)
            elif k == "json":
                jdocs.append(doc); jidx.append(i)
            elif k == "xlsx":
                try:
                    books.append(xlsx_host.scan_workbook(doc))
                    book_idx.append(i)
                except Exception as e:  # noqa: BLE001
                    out[i] = FileParsingError(f"Excel parsing failed: {e}")
            else:
                out[i] = FileParsingError(f"Unsupported file type: .{k}")
        eng = self.engine
        if flat:
            with eng.lock:   # the pinned output is ours until every text has been built
                r = eng.mixed_batch(flat, fkind, self.max_excel_rows)
                ok = [d for d in range(len(owner)) if fkind[d] == _lib.KIND_TEXT or int(r.status[d]) == _lib.DOC_OK]
                texts = dict(zip(ok, r.doc_strs(ok)))
                for d, i in enumerate(owner):
                    st = int(r.status[d])
                    if fkind[d] == _lib.KIND_TEXT:
                        out[i] = texts[d]
                    else:
                        out[i] = (texts[d] if st == _lib.DOC_OK
                                  else FileParsingError(f"CSV parsing failed: {_CSV_FAIL}") if st == _lib.DOC_CSV_UNPARSEABLE
                                  else FileParsingError("CSV parsing failed: unsupported construct on the GPU path "
                                                        f"(status {st})"))
        if jdocs:
            for i, t in zip(jidx, eng.parse_json_many(jdocs)):
                out[i] = self._json_result(t)
        if books:
            for i, t in zip(book_idx, xlsx_host.parse_workbooks(eng, books, self.max_excel_rows)):
                out[i] = FileParsingError(f"Excel parsing failed: {t}") if isinstance(t, Exception) else t
        return out

    _GPU_HASH_MAX = 1 << 20   # MD5 is sequential per document: bigger files keep hashlib

    def parse_many_envelopes(self, docs: Sequence[bytes], names: Sequence[str]) -> "list[Any]":
        """parse_file_async's result dict (app/file_parser.py:80-86) for many in-memory files:
        raw_text from the batched parsers, file_hash from the batched MD5 kernel.  `names[i]` is the
        file name (its suffix selects the parser).  Entries are dicts, or the exception
        parse_file_async would have raised: un-wrapped size / type errors (:109,113), wrapped
        FileParsingError(f"Failed to parse {name}: ...") for everything from the dispatch (:95)."""
        n = len(docs)
        out: "list[Any]" = [None] * n
        kinds = [Path(nm).suffix.lower() for nm in names]
        ok = []
        for i, (b, k) in enumerate(zip(docs, kinds)):
            if len(b) > self.max_file_size:
                out[i] = FileParsingError(f"File too large: {len(b) / 1024 / 1024:.1f}MB "
                                          f"(max: {self.max_file_size / 1024 / 1024}MB)")
            elif k not in self.supported_extensions:
                out[i] = FileParsingError(f"Unsupported file type: {k}")
            else:
                ok.append(i)
        small = [i for i in ok if len(docs[i]) <= self._GPU_HASH_MAX]
        hashes = dict(zip(small, self.engine.md5_many([docs[i] for i in small]))) if small else {}
        for i in ok:
            if i not in hashes:
                hashes[i] = hashlib.md5(docs[i]).hexdigest()
        texts = self.parse_many([docs[i] for i in ok], [kinds[i] for i in ok])
        for i, t in zip(ok, texts):
            name = Path(names[i]).name
            if isinstance(t, Exception):
                out[i] = FileParsingError(f"Failed to parse {name}: {t}")
            else:
                out[i] = {"raw_text": t, "source_file": name, "file_size": len(docs[i]),
                          "file_hash": hashes[i], "parsing_method": f"async_{kinds[i][1:]}"}
        return out

    def __del__(self):
        if hasattr(self, "thread_pool"):
            self.thread_pool.shutdown(wait=False)


async def parse_file_async(file_path: str, max_file_size_mb: int = 10) -> Dict[str, Any]:
    """app/file_parser.py:365-368."""
    return await FileParser(max_file_size_mb=max_file_size_mb).parse_file_async(file_path)


def get_supported_extensions() -> set:
    """app/file_parser.py:370-375, restricted to the formats this build parses."""
    return set(_SUPPORTED)
