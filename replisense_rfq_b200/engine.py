"""Batched many-documents entry point over the C ABI (SURVEY.md §8b "new entry point")."""
from __future__ import annotations

import ctypes as C
import threading
from typing import Sequence

import numpy as np

from . import _lib

ENCODINGS = ["utf-8", "latin-1", "cp1252"]  # app/file_parser.py:298
SEPARATORS = [",", ";", "\t"]  # app/file_parser.py:299


class EngineError(RuntimeError):
    pass


_decode_utf8 = C.pythonapi.PyUnicode_DecodeUTF8
_decode_utf8.restype = C.py_object
_decode_utf8.argtypes = [C.c_void_p, C.c_ssize_t, C.c_char_p]


# Big texts: CPython's UTF-8 decoder is one thread writing into freshly mapped pages (60 ms for 135 MB).
# When the code-point count is known (the device counted it), the string object is allocated at its final
# width and the library fills its buffer with several threads (rfq_out_widen).  The offsets of a compact
# string's data are checked on a small string at import; on any surprise the plain decoder is used.
_py_new = C.pythonapi.PyUnicode_New
_py_new.restype = C.py_object
_py_new.argtypes = [C.c_ssize_t, C.c_uint32]


def _str_data_offsets():
    import sys
    try:
        # header sizes of CPython's compact strings, from objects whose size is header + units + terminator
        off_ascii, off_compact = sys.getsizeof("") - 1, sys.getsizeof("a\xe9") - 3
        if sys.implementation.name != "cpython" or (off_ascii, off_compact) != (40, 56):
            return None
        for text, maxchar, off, unit in (("abc", 127, off_ascii, 1), ("a\xe9c", 255, off_compact, 1),
                                         ("a\u20b9c", 0xFFFF, off_compact, 2), ("a\U0001f600c", 0x10FFFF, off_compact, 4)):
            o = _py_new(len(text), maxchar)
            raw = text.encode({1: "latin-1", 2: "utf-16-le", 4: "utf-32-le"}[unit])
            C.memmove(id(o) + off, raw, len(raw))
            if o != text or len(o) != len(text):
                return None
        return off_ascii, off_compact
    except Exception:  # noqa: BLE001
        return None


_STR_OFFSETS = _str_data_offsets()
_FAST_STR_MIN = 1 << 20

_PACK_THREADS = 8
_pack_executor = None


def _pack_pool():
    global _pack_executor
    if _pack_executor is None:
        from concurrent.futures import ThreadPoolExecutor
        _pack_executor = ThreadPoolExecutor(max_workers=_PACK_THREADS)
    return _pack_executor


class PinnedArena:
    """Pinned host memory the caller packs documents into (16-byte aligned starts)."""

    def __init__(self, engine: "Engine", capacity: int):
        self.engine = engine
        self.capacity = max(int(capacity), 16)
        self.ptr = engine.lib.rfq_host_alloc(engine.ctx, self.capacity)
        if not self.ptr:
            raise EngineError("pinned allocation failed")
        self.view = np.ctypeslib.as_array((C.c_uint8 * self.capacity).from_address(self.ptr))

    def pack(self, docs: Sequence[bytes]):
        n = len(docs)
        lengths = np.fromiter((len(d) for d in docs), dtype=np.uint64, count=n)
        padded = (lengths + np.uint64(15)) & ~np.uint64(15)
        offsets = np.zeros(n, dtype=np.uint64)
        if n > 1:
            np.cumsum(padded[:-1], out=offsets[1:])
        total = int(padded.sum())
        if total > self.capacity:
            raise EngineError("arena too small")
        v = self.view
        offs = offsets.tolist()

        def copy(lo, hi):
            for i in range(lo, hi):
                d = docs[i]
                if d:
                    v[offs[i]:offs[i] + len(d)] = np.frombuffer(d, dtype=np.uint8)

        if total < (8 << 20) or n < 8:
            copy(0, n)
        else:
            # big batch: the copies release the GIL, so a few threads fill the arena side by side
            # (one thread moves ~5 GB/s, which would otherwise be the slowest step of the call)
            k = min(_PACK_THREADS, n)
            target = total / k
            cuts, acc, nxt = [0], 0, target
            for i, p in enumerate(padded.tolist()):
                acc += p
                if acc >= nxt and len(cuts) < k:
                    cuts.append(i + 1)
                    nxt += target
            cuts.append(n)
            list(_pack_pool().map(lambda ab: copy(*ab), zip(cuts[:-1], cuts[1:])))
        return offsets, lengths, total

    def free(self):
        if self.ptr:
            self.view = None
            self.engine.lib.rfq_host_free(self.engine.ctx, self.ptr)
            self.ptr = None


class BatchResult:
    """Views over the library-owned pinned output of the last run.  The text stays valid only until
    the engine's next call: read it while holding `engine.lock` (the `parse_*_many` methods and
    `Engine.locked()` do), or from the one thread that owns the engine."""

    def __init__(self, eng: "Engine"):
        lib, ctx = eng.lib, eng.ctx
        n = lib.rfq_out_count(ctx)
        self.n = n
        self.total = lib.rfq_out_total_bytes(ctx)
        self.text_ptr = lib.rfq_out_text(ctx)

        def arr(ptr, ctype, dtype):
            if n == 0 or not ptr:
                return np.zeros(0, dtype=dtype)
            return np.ctypeslib.as_array((ctype * n).from_address(ptr)).copy()

        self.offsets = arr(lib.rfq_out_offsets(ctx), C.c_uint64, np.uint64)
        self.lengths = arr(lib.rfq_out_lengths(ctx), C.c_uint64, np.uint64)
        self.status = arr(lib.rfq_out_status(ctx), C.c_int32, np.int32)
        self.rows = arr(lib.rfq_out_rows(ctx), C.c_uint32, np.uint32)
        self.cols = arr(lib.rfq_out_cols(ctx), C.c_uint32, np.uint32)
        self.attempt = arr(lib.rfq_out_attempt(ctx), C.c_uint8, np.uint8)
        self._lib, self._ctx = lib, ctx
        cp = lib.rfq_out_chars(ctx)
        self.chars = np.ctypeslib.as_array((C.c_uint64 * n).from_address(cp)).copy() if n and cp else None

    def doc_bytes(self, d: int) -> bytes:
        ln = int(self.lengths[d])
        if ln == 0:
            return b""
        return C.string_at(self.text_ptr + int(self.offsets[d]), ln)

    def doc_str(self, d: int) -> str:
        ln = int(self.lengths[d])
        if ln == 0:
            return ""
        if ln >= _FAST_STR_MIN and _STR_OFFSETS is not None and self.chars is not None:
            nch = int(self.chars[d])
            if nch != 0xFFFFFFFFFFFFFFFF and 0 < nch <= ln:
                mc = C.c_uint32()
                if self._lib.rfq_out_maxchar(self._ctx, d, C.byref(mc)) == 0:
                    unit = 1 if mc.value <= 0xFF else 2 if mc.value <= 0xFFFF else 4
                    o = _py_new(nch, mc.value)
                    off = _STR_OFFSETS[0] if mc.value <= 0x7F else _STR_OFFSETS[1]
                    if self._lib.rfq_out_widen(self._ctx, d, unit, id(o) + off, nch) == 0:
                        return o
        # decode straight out of the pinned buffer (no intermediate bytes object)
        return _decode_utf8(self.text_ptr + int(self.offsets[d]), ln, None)


    def doc_strs(self, ds) -> "list[str]":
        """doc_str for many documents of this result: the texts are sized and filled by library threads
        (rfq_out_plan_many / rfq_out_widen_many), Python only allocates the str objects.  For a batch of sheets,
        bodies or small files the one-at-a-time decode is the largest host cost of the call."""
        ds = [int(d) for d in ds]
        lens = self.lengths
        if _STR_OFFSETS is None or len(ds) < 8 or sum(int(lens[d]) for d in ds) < (1 << 20):
            return [self.doc_str(d) for d in ds]
        live = [d for d in ds if int(lens[d]) > 0]
        n = len(live)
        idx = np.asarray(live, dtype=np.uint32)
        nch = np.zeros(n, dtype=np.uint64)
        mc = np.zeros(n, dtype=np.uint32)
        if self._lib.rfq_out_plan_many(self._ctx, n, idx.ctypes.data, nch.ctypes.data, mc.ctypes.data) != 0:
            return [self.doc_str(d) for d in ds]
        units = np.empty(n, dtype=np.uint8)
        dsts = np.empty(n, dtype=np.uint64)
        objs = []
        off_a, off_c = _STR_OFFSETS
        for i, (k, m) in enumerate(zip(nch.tolist(), mc.tolist())):
            o = _py_new(k, m)
            objs.append(o)
            units[i] = 1 if m <= 0xFF else 2 if m <= 0xFFFF else 4
            dsts[i] = id(o) + (off_a if m <= 0x7F else off_c)
        if self._lib.rfq_out_widen_many(self._ctx, n, idx.ctypes.data, units.ctypes.data, dsts.ctypes.data, nch.ctypes.data) != 0:
            return [self.doc_str(d) for d in ds]
        texts = dict(zip(live, objs))
        return [texts.get(d, "") for d in ds]


class Engine:
    """One CUDA device, one stream. Thread-safe (calls are serialised per engine)."""

    def __init__(self, device: int = 0):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.rfq_create(device, C.byref(h))
        if rc != 0 or not h:
            raise EngineError(
                f"rfq_create(device={device}) failed: no CUDA device — there is no CPU fallback")
        self.ctx = h
        self.device = device
        # re-entrant: a caller that has to read a BatchResult takes it around the *_batch call
        self.lock = threading.RLock()
        self._arena = None

    # -- helpers
    def _check(self, rc: int, what: str):
        if rc != 0:
            raise EngineError(f"{what} failed ({rc}): {self.lib.rfq_last_error(self.ctx).decode()}")

    def arena(self, nbytes: int) -> PinnedArena:
        if self._arena is None or self._arena.capacity < nbytes:
            if self._arena is not None:
                self._arena.free()
            self._arena = PinnedArena(self, int(nbytes * 1.25) + 4096)
        return self._arena

    def _pack(self, docs: Sequence[bytes]):
        need = sum((len(d) + 15) & ~15 for d in docs) + 64
        a = self.arena(need)
        offsets, lengths, total = a.pack(docs)
        return a, offsets, lengths

    def read_file(self, path) -> "tuple[PinnedArena, np.ndarray, np.ndarray]":
        """The bytes of one file read straight into this engine's pinned arena (no bytes object, no second
        copy): what a file-level parser stages.  Call with the engine lock held, like _pack."""
        import os
        with open(path, "rb", buffering=0) as f:
            n = os.fstat(f.fileno()).st_size
            a = self.arena(((n + 15) & ~15) + 64)
            got, mv = 0, memoryview(a.view)
            while got < n:
                k = f.readinto(mv[got:n])
                if not k:
                    break
                got += k
        return a, np.zeros(1, dtype=np.uint64), np.array([got], dtype=np.uint64)

    def parse_csv_file(self, path, max_rows=1000) -> "str | None":
        """raw_text of one .csv file (None = every (encoding, separator) attempt failed); raises EngineError for a
        construct the GPU path declines."""
        with self.lock:
            a, off, ln = self.read_file(path)
            mr = self._mr(1, max_rows)
            self._check(self.lib.rfq_csv_parse(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, 1, mr.ctypes.data),
                        "rfq_csv_parse")
            r = BatchResult(self)
            st = int(r.status[0])
            if st == _lib.DOC_OK:
                return r.doc_str(0)
            if st == _lib.DOC_CSV_UNPARSEABLE:
                return None
            raise EngineError(f"document 0: unsupported construct on the GPU path (status {st})")

    def stage(self, arena: PinnedArena, offsets: np.ndarray, lengths: np.ndarray):
        self._check(self.lib.rfq_stage(self.ctx, arena.ptr, offsets.ctypes.data,
                                       lengths.ctypes.data, len(offsets)), "rfq_stage")

    def set_stream(self, stream_ptr: int | None):
        self._check(self.lib.rfq_set_stream(self.ctx, stream_ptr), "rfq_set_stream")

    # -- text
    def text_batch(self, docs: Sequence[bytes]) -> BatchResult:
        with self.lock:
            a, off, ln = self._pack(docs)
            self._check(self.lib.rfq_text_parse(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data,
                                                len(docs)), "rfq_text_parse")
            return BatchResult(self)

    def parse_text_many(self, docs: Sequence[bytes]) -> "list[str]":
        with self.lock:
            a, off, ln = self._pack(docs)
            self._check(self.lib.rfq_text_parse(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data,
                                                len(docs)), "rfq_text_parse")
            r = BatchResult(self)
            return r.doc_strs(range(r.n))

    def parse_text_file(self, path) -> str:
        """raw_text of one .txt file, read straight into pinned memory."""
        with self.lock:
            a, off, ln = self.read_file(path)
            self._check(self.lib.rfq_text_parse(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, 1), "rfq_text_parse")
            return BatchResult(self).doc_str(0)

    # -- file hashes (app/file_parser.py:122-130), one batch
    def md5_many(self, docs: Sequence[bytes]) -> "list[str]":
        if not docs:
            return []
        with self.lock:
            a, off, ln = self._pack(docs)
            self._check(self.lib.rfq_stage(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, len(docs)),
                        "rfq_stage")
            self._check(self.lib.rfq_md5_run(self.ctx), "rfq_md5_run")
            self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
            r = BatchResult(self)
            return [r.doc_str(d) for d in range(r.n)]

    # -- quantity / price validators (app/rfq_agent.py:27-39)
    def numeric_batch(self, fields: Sequence[bytes], kinds) -> "tuple[np.ndarray, np.ndarray]":
        """kinds[i]: 0 = int(v.replace(',', '')), 1 = float(v.replace(',', '')).  Returns (raw 8-byte
        values as uint64, status)."""
        kinds = np.ascontiguousarray(kinds, dtype=np.uint8)
        if not len(fields):
            return np.zeros(0, np.uint64), np.zeros(0, np.int32)
        with self.lock:
            a, off, ln = self._pack(fields)
            self._check(self.lib.rfq_stage(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, len(fields)),
                        "rfq_stage")
            self._check(self.lib.rfq_numeric_run(self.ctx, kinds.ctypes.data), "rfq_numeric_run")
            self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
            r = BatchResult(self)
            raw = (np.frombuffer(C.string_at(r.text_ptr, 8 * r.n), dtype=np.uint64).copy() if r.n
                   else np.zeros(0, np.uint64))
            return raw, np.asarray(r.status).copy()

    # -- text framing (app/file_parser.py:166-212): documents assembled from pieces on the device
    def frame_many(self, pieces: Sequence[bytes], strip, part_kind, part_first_piece, doc_first_part) -> "list[str]":
        n_out = len(doc_first_part) - 1
        if n_out <= 0:
            return []
        strip = np.ascontiguousarray(strip, dtype=np.uint8)
        pk = np.ascontiguousarray(part_kind, dtype=np.uint8)
        pfp = np.ascontiguousarray(part_first_piece, dtype=np.uint32)
        dfp = np.ascontiguousarray(doc_first_part, dtype=np.uint32)
        with self.lock:
            a, off, ln = self._pack(pieces)
            self._check(self.lib.rfq_stage(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, len(pieces)), "rfq_stage")
            self._check(self.lib.rfq_frame_run(self.ctx, n_out, dfp.ctypes.data, len(pk), pk.ctypes.data, pfp.ctypes.data,
                                               strip.ctypes.data), "rfq_frame_run")
            self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
            r = BatchResult(self)
            return [r.doc_str(d) for d in range(r.n)]

    # -- where the JSON sits in LLM replies (app/rfq_agent.py:208-236)
    def json_spans(self, replies: Sequence[bytes]) -> np.ndarray:
        """[n, 4] u32: first '{', last '}' (0xFFFFFFFF = none), cleaned [begin, end) of every reply."""
        if not len(replies):
            return np.zeros((0, 4), np.uint32)
        with self.lock:
            a, off, ln = self._pack(replies)
            self._check(self.lib.rfq_stage(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, len(replies)), "rfq_stage")
            self._check(self.lib.rfq_json_spans_run(self.ctx), "rfq_json_spans_run")
            self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
            r = BatchResult(self)
            return np.frombuffer(C.string_at(r.text_ptr, 16 * r.n), dtype=np.uint32).reshape(r.n, 4).copy()

    # -- the .json file type (app/file_parser.py:315-331)
    def parse_json_many(self, docs: Sequence[bytes]) -> "list[str | EngineError]":
        """raw_text of every .json document ("=== JSON Data ===" + json.dumps(json.loads(text), indent=2,
        ensure_ascii=False)), or an EngineError in its slot: `.json_error` holds the text of the
        json.JSONDecodeError the reference would see ("Expecting value: line 1 column 1 (char 0)"); a
        document the GPU path declines has `.json_error` None."""
        if not len(docs):
            return []
        with self.lock:
            a, off, ln = self._pack(docs)
            self._check(self.lib.rfq_stage(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, len(docs)), "rfq_stage")
            self._check(self.lib.rfq_json_run(self.ctx), "rfq_json_run")
            self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
            r = BatchResult(self)
            out: "list[Any]" = []
            ok = [d for d in range(r.n) if int(r.status[d]) == _lib.DOC_OK]
            texts = dict(zip(ok, r.doc_strs(ok)))
            for d in range(r.n):
                st = int(r.status[d])
                if st == _lib.DOC_OK:
                    out.append(texts[d])
                elif st == _lib.DOC_JSON_INVALID:
                    e = EngineError("invalid JSON")
                    e.json_error = "%s: line %d column %d (char %d)" % (_lib.JSON_ERRORS[int(r.attempt[d])], int(r.rows[d]),
                                                                       int(r.cols[d]), int(r.chars[d]))
                    out.append(e)
                else:
                    e = EngineError("JSON document not decoded on the GPU path (status %d): not UTF-8, duplicate keys, a lone "
                                    "surrogate escape, an integer beyond 4300 digits or nesting beyond 96" % st)
                    e.json_error = None
                    out.append(e)
            return out

    # -- prompt assembly (the step after the parser: app/rfq_agent.py:138-151)
    def prompt_batch(self, docs: Sequence[bytes], head: bytes, tail: bytes, mark: bytes,
                     max_chars: int = 8000) -> BatchResult:
        """docs: raw_text documents (UTF-8).  status RFQ_DOC_EMPTY_TEXT where the text is blank."""
        with self.lock:
            a, off, ln = self._pack(docs)
            self._check(self.lib.rfq_prompt_build(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data,
                                                  len(docs), head, len(head), tail, len(tail), mark,
                                                  len(mark), max_chars), "rfq_prompt_build")
            return BatchResult(self)

    # -- csv
    def csv_batch(self, docs: Sequence[bytes], max_rows=1000) -> BatchResult:
        n = len(docs)
        if np.isscalar(max_rows):
            mr = np.full(n, int(max_rows), dtype=np.int64)
        else:
            mr = np.asarray(max_rows, dtype=np.int64)
        with self.lock:
            a, off, ln = self._pack(docs)
            self._check(self.lib.rfq_csv_parse(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data,
                                               n, mr.ctypes.data), "rfq_csv_parse")
            return BatchResult(self)

    def parse_csv_many(self, docs: Sequence[bytes], max_rows=1000,
                       on_unsupported: str = "raise") -> "list[str | None | EngineError]":
        """raw_text per document; None where every (encoding, separator) attempt failed (the
        caller raises the reference's FileParsingError).  A document that uses a construct the GPU
        path does not decode raises EngineError, or — with on_unsupported="return" — gets an
        EngineError instance in its slot so the rest of the batch survives.  The texts are built
        while the engine lock is held: the pinned output belongs to this call until then."""
        with self.lock:
            r = self.csv_batch(docs, max_rows)
            out = []
            ok = [d for d in range(r.n) if int(r.status[d]) == _lib.DOC_OK]
            texts = dict(zip(ok, r.doc_strs(ok)))
            for d in range(r.n):
                st = int(r.status[d])
                if st == _lib.DOC_OK:
                    out.append(texts[d])
                elif st == _lib.DOC_CSV_UNPARSEABLE:
                    out.append(None)
                else:
                    err = EngineError(f"document {d}: unsupported construct on the GPU path (status {st})")
                    if on_unsupported == "raise":
                        raise err
                    out.append(err)
            return out

    # -- csv parse fused with prompt assembly: raw_text never leaves the device
    def csv_prompt_arena(self, arena: PinnedArena, offsets, lengths, max_rows, head: bytes, tail: bytes,
                         mark: bytes, max_chars: int = 8000) -> BatchResult:
        """rfq_csv_prompt_parse over a packed pinned arena.  The BatchResult holds the prompts;
        `.raw_bytes` / `.raw_chars` give the size of each raw_text (text_length of the reference)."""
        mr = self._mr(len(offsets), max_rows)
        with self.lock:
            self._check(self.lib.rfq_csv_prompt_parse(self.ctx, arena.ptr, offsets.ctypes.data, lengths.ctypes.data,
                                                      len(offsets), mr.ctypes.data, head, len(head), tail, len(tail),
                                                      mark, len(mark), max_chars), "rfq_csv_prompt_parse")
            r = BatchResult(self)
            n = r.n
            for name, fn in (("raw_bytes", self.lib.rfq_out_raw_bytes), ("raw_chars", self.lib.rfq_out_raw_chars)):
                ptr = fn(self.ctx)
                setattr(r, name, np.ctypeslib.as_array((C.c_uint64 * n).from_address(ptr)).copy() if n and ptr
                        else np.zeros(0, np.uint64))
            return r

    def csv_prompts_many(self, docs: Sequence[bytes], max_rows, head: bytes, tail: bytes, mark: bytes,
                         max_chars: int = 8000) -> "list[tuple[str | None, int, int]]":
        """(prompt or None, text_length in code points, status) per CSV document, one fused GPU call."""
        with self.lock:
            a, off, ln = self._pack(docs)
            r = self.csv_prompt_arena(a, off, ln, max_rows, head, tail, mark, max_chars)
            return [(r.doc_str(d) if int(r.status[d]) == _lib.DOC_OK else None, int(r.raw_chars[d]), int(r.status[d]))
                    for d in range(r.n)]

    # -- sheets
    def sheets_batch(self, docs: Sequence[bytes], kind, sst_of, max_rows=1000, date_ids=None) -> BatchResult:
        """docs: worksheet XML parts (kind 0, or 2 = workbook with the 1904 epoch) and sharedStrings
        parts (kind 1); sst_of[d] = index of the sharedStrings part of sheet d or -1; date_ids[d] =
        iterable of date style ids (bit 31 = a timedelta format)."""
        with self.lock:
            a, off, ln = self._pack(docs)
            return self._sheets(a, off, ln, kind, sst_of, max_rows, date_ids)

    def sheets_batch_deflated(self, members, kind, sst_of, max_rows=1000, date_ids=None):
        """sheets_batch over zip members that are still compressed (xlsx_host.Member): upload the raw
        data, inflate on the device (rfq_inflate_run), make the result the staged batch
        (rfq_restage_output), run the sheet path.  Returns (BatchResult, inflate status per member)."""
        n = len(members)
        method = np.fromiter((m.method for m in members), dtype=np.uint8, count=n)
        size = np.fromiter((m.size for m in members), dtype=np.uint64, count=n)
        crc = np.fromiter((m.crc for m in members), dtype=np.uint32, count=n)
        with self.lock:
            a, off, ln = self._pack([m.raw for m in members])
            self._check(self.lib.rfq_stage(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, n), "rfq_stage")
            self._check(self.lib.rfq_inflate_run(self.ctx, method.ctypes.data, size.ctypes.data, crc.ctypes.data),
                        "rfq_inflate_run")
            self._check(self.lib.rfq_fetch_tables(self.ctx), "rfq_fetch_tables")
            inflate_status = BatchResult(self).status.copy()
            self._check(self.lib.rfq_restage_output(self.ctx), "rfq_restage_output")
            kind = np.ascontiguousarray(kind, dtype=np.uint8)
            so = np.ascontiguousarray(sst_of, dtype=np.int32)
            mr = self._mr(n, max_rows)
            doff = np.zeros(n + 1, dtype=np.uint32)
            flat = []
            if date_ids is not None:
                for d in range(n):
                    if date_ids[d]:
                        flat.extend(sorted(date_ids[d]))
                    doff[d + 1] = len(flat)
            dids = np.asarray(flat if flat else [0], dtype=np.uint32)
            self._check(self.lib.rfq_sheets_run(self.ctx, kind.ctypes.data, so.ctypes.data, mr.ctypes.data,
                                                doff.ctypes.data, dids.ctypes.data), "rfq_sheets_run")
            self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
            return BatchResult(self), inflate_status

    def inflate_many(self, members) -> "list[bytes | None]":
        """The inflated bytes of every member (None where the stream, its size or its CRC-32 is bad)."""
        n = len(members)
        if n == 0:
            return []
        method = np.fromiter((m.method for m in members), dtype=np.uint8, count=n)
        size = np.fromiter((m.size for m in members), dtype=np.uint64, count=n)
        crc = np.fromiter((m.crc for m in members), dtype=np.uint32, count=n)
        with self.lock:
            a, off, ln = self._pack([m.raw for m in members])
            self._check(self.lib.rfq_stage(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, n), "rfq_stage")
            self._check(self.lib.rfq_inflate_run(self.ctx, method.ctypes.data, size.ctypes.data, crc.ctypes.data),
                        "rfq_inflate_run")
            self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
            r = BatchResult(self)
            return [r.doc_bytes(d) if int(r.status[d]) == _lib.DOC_OK else None for d in range(n)]

    def sheets_parse_arena(self, arena: PinnedArena, offsets, lengths, kind, sst_of, max_rows=1000,
                           date_ids=None) -> BatchResult:
        """sheets_batch over parts the caller already packed into a pinned arena (no host copy)."""
        with self.lock:
            return self._sheets(arena, offsets, lengths, kind, sst_of, max_rows, date_ids)

    def _sheets(self, a, off, ln, kind, sst_of, max_rows, date_ids) -> BatchResult:
        n = len(off)
        kind = np.ascontiguousarray(kind, dtype=np.uint8)
        sst_of = np.ascontiguousarray(sst_of, dtype=np.int32)
        mr = self._mr(n, max_rows)
        doff = np.zeros(n + 1, dtype=np.uint32)
        flat = []
        if date_ids is not None:
            for d in range(n):
                if date_ids[d]:
                    flat.extend(sorted(date_ids[d]))
                doff[d + 1] = len(flat)
        dids = np.asarray(flat if flat else [0], dtype=np.uint32)
        self._check(self.lib.rfq_sheets_parse(self.ctx, a.ptr, off.ctypes.data, ln.ctypes.data, n,
                                              kind.ctypes.data, sst_of.ctypes.data, mr.ctypes.data,
                                              doff.ctypes.data, dids.ctypes.data), "rfq_sheets_parse")
        return BatchResult(self)

    # -- mixed batch: text, CSV and sheet parts in one arena, one upload (rfq_mixed_parse)
    def mixed_parse_arena(self, arena: PinnedArena, offsets, lengths, kind, max_rows=1000, sst_of=None,
                          date_ids=None) -> BatchResult:
        n = len(offsets)
        kind = np.ascontiguousarray(kind, dtype=np.uint8)
        mr = self._mr(n, max_rows)
        so = np.full(n, -1, dtype=np.int32) if sst_of is None else np.ascontiguousarray(sst_of, dtype=np.int32)
        doff = np.zeros(n + 1, dtype=np.uint32)
        flat = []
        if date_ids is not None:
            for d in range(n):
                if date_ids[d]:
                    flat.extend(sorted(date_ids[d]))
                doff[d + 1] = len(flat)
        dids = np.asarray(flat if flat else [0], dtype=np.uint32)
        with self.lock:
            self._check(self.lib.rfq_mixed_parse(self.ctx, arena.ptr, offsets.ctypes.data, lengths.ctypes.data, n,
                                                 kind.ctypes.data, mr.ctypes.data, so.ctypes.data,
                                                 doff.ctypes.data, dids.ctypes.data), "rfq_mixed_parse")
            return BatchResult(self)

    def mixed_batch(self, docs: Sequence[bytes], kind, max_rows=1000, sst_of=None, date_ids=None) -> BatchResult:
        """docs[i] of kind[i] (_lib.KIND_*): one pack, one upload, the three pipelines over their subsets."""
        with self.lock:
            a, off, ln = self._pack(docs)
            return self.mixed_parse_arena(a, off, ln, kind, max_rows, sst_of, date_ids)

    # -- pre-packed pinned arena (no Python-side copies): what a high-throughput caller uses
    def _mr(self, n, max_rows):
        if np.isscalar(max_rows):
            return np.full(n, int(max_rows), dtype=np.int64)
        return np.ascontiguousarray(max_rows, dtype=np.int64)

    def csv_parse_arena(self, arena: PinnedArena, offsets, lengths, max_rows=1000) -> BatchResult:
        mr = self._mr(len(offsets), max_rows)
        with self.lock:
            self._check(self.lib.rfq_csv_parse(self.ctx, arena.ptr, offsets.ctypes.data,
                                               lengths.ctypes.data, len(offsets), mr.ctypes.data),
                        "rfq_csv_parse")
            return BatchResult(self)

    def text_parse_arena(self, arena: PinnedArena, offsets, lengths) -> BatchResult:
        with self.lock:
            self._check(self.lib.rfq_text_parse(self.ctx, arena.ptr, offsets.ctypes.data,
                                                lengths.ctypes.data, len(offsets)), "rfq_text_parse")
            return BatchResult(self)

    def csv_run(self, n, max_rows=1000):
        """Kernels only, on the staged (device-resident) batch."""
        mr = self._mr(n, max_rows)
        self._check(self.lib.rfq_csv_run(self.ctx, mr.ctypes.data), "rfq_csv_run")

    def text_run(self):
        self._check(self.lib.rfq_text_run(self.ctx), "rfq_text_run")

    def fetch(self) -> BatchResult:
        self._check(self.lib.rfq_fetch(self.ctx), "rfq_fetch")
        return BatchResult(self)

    def set_profile(self, on: bool):
        self.lib.rfq_set_profile(self.ctx, int(on))

    # -- measurement
    def last_run_ms(self) -> float:
        return float(self.lib.rfq_last_run_ms(self.ctx))

    def launch_count(self) -> int:
        return int(self.lib.rfq_launch_count(self.ctx))

    def profile(self) -> "list[tuple[str, float]]":
        k = self.lib.rfq_profile_count(self.ctx)
        return [(self.lib.rfq_profile_name(self.ctx, i).decode(), float(self.lib.rfq_profile_ms(self.ctx, i)))
                for i in range(k)]

    def close(self):
        if getattr(self, "ctx", None):
            if self._arena is not None:
                self._arena.free()
                self._arena = None
            self.lib.rfq_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = {}
_default_lock = threading.Lock()


def default_engine(device: int = 0) -> Engine:
    with _default_lock:
        if device not in _default:
            _default[device] = Engine(device)
        return _default[device]


class EnginePool:
    """`depth` engines (one ctx + one CUDA stream + own pinned output each) on ONE device.

    A single call is a chain H2D copy -> kernels -> D2H copy, and on a PCIe-attached GPU the two
    copies are most of its time.  PCIe is full duplex and the copy engines run beside the SMs, so
    two calls in flight overlap one call's upload with the other's kernels and download: the
    steady-state cost per call drops from (H2D + kernels + D2H) to about max(H2D, kernels, D2H).
    This is the GPU counterpart of the reference's 4-thread pool behind one shared FileParser
    (app/file_parser.py:33, app/main.py:47): `submit` may be called from any thread.
    """

    def __init__(self, device: int = 0, depth: int = 2):
        import queue
        from concurrent.futures import ThreadPoolExecutor
        self.engines = [Engine(device) for _ in range(depth)]
        self._free = queue.Queue()
        for e in self.engines:
            self._free.put(e)
        self._pool = ThreadPoolExecutor(max_workers=depth)
        self._rr, self._rr_lock = 0, threading.Lock()

    def _run(self, fn, args):
        e = self._free.get()
        try:
            return fn(e, *args)
        finally:
            self._free.put(e)

    def submit(self, fn, *args):
        """fn(engine, *args) on the next free engine; returns a Future.  A BatchResult returned by fn
        views that engine's pinned output and stays valid until the engine's next call: copy what
        you need inside fn."""
        return self._pool.submit(self._run, fn, args)

    def map(self, fn, items):
        return [f.result() for f in [self.submit(fn, it) for it in items]]

    def pick(self) -> Engine:
        """An engine that is idle right now, else the next one in turn (its caller then waits on that
        engine's lock): concurrent callers of one shared FileParser land on different streams.  Every
        parse_*_many call holds its engine's lock until its result has been turned into Python
        objects, so a busy engine handed out here is never rerun under a reader."""
        for e in self.engines:
            if e.lock.acquire(blocking=False):
                e.lock.release()
                return e
        with self._rr_lock:
            self._rr = (self._rr + 1) % len(self.engines)
            return self.engines[self._rr]

    def close(self):
        self._pool.shutdown(wait=True)


_default_pools: "dict[int, EnginePool]" = {}


def default_pool(device: int = 0, depth: int = 2) -> EnginePool:
    with _default_lock:
        if device not in _default_pools:
            _default_pools[device] = EnginePool(device, depth)
        return _default_pools[device]
