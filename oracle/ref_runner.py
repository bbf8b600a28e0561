"""Runs the reference's OWN `FileParser` (oracle/_ref/app/file_parser.py, put there unmodified by
oracle/build_ref.py) on the host cores.  TEST / BASELINE INFRASTRUCTURE ONLY (bench.py --impl reference,
tests): never imported by the product."""
from __future__ import annotations

import importlib.util
import logging
import os
import sys
import tempfile
from pathlib import Path

REF_DIR = Path(__file__).resolve().parent / "_ref"
_mod = None


def available() -> bool:
    return (REF_DIR / "app" / "file_parser.py").exists()


def module():
    """The reference module, loaded under a private name so it never shadows anything."""
    global _mod
    if _mod is None:
        stubs = str(REF_DIR / "stubs")
        if stubs not in sys.path:
            sys.path.append(stubs)          # `import docx` inside the reference finds the empty stand-in
        spec = importlib.util.spec_from_file_location("_rfq_reference_file_parser", REF_DIR / "app" / "file_parser.py")
        m = importlib.util.module_from_spec(spec)
        logging.getLogger("_rfq_reference_file_parser").setLevel(logging.CRITICAL)
        spec.loader.exec_module(m)
        m.logger.setLevel(logging.CRITICAL)
        _mod = m
    return _mod


def parse_csv_bytes(data: bytes, max_rows: int = 1000) -> str:
    """FileParser(max_excel_rows=max_rows)._parse_csv_sync on a temp file holding `data`."""
    m = module()
    d = "/dev/shm" if os.path.isdir("/dev/shm") else None
    fd, path = tempfile.mkstemp(suffix=".csv", dir=d)
    try:
        with os.fdopen(fd, "wb") as f:
            f.write(data)
        return m.FileParser(max_file_size_mb=1 << 20, max_excel_rows=max_rows)._parse_csv_sync(Path(path))
    finally:
        os.unlink(path)
