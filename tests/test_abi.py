"""No-GPU checks of the drop-in boundary: the library builds, loads, and exports exactly the
symbols include/rfq_parse.h declares; the host mirror keeps the reference's error behaviour."""
import asyncio
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="session")
def lib():
    from replisense_rfq_b200 import build, _lib
    build.build()
    return _lib.load()


def declared_symbols():
    text = (ROOT / "include" / "rfq_parse.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rfq_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from replisense_rfq_b200 import _lib
    syms = declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), s
    assert set(syms) == set(_lib.SIGNATURES), set(syms) ^ set(_lib.SIGNATURES)


def test_version_and_no_device_is_loud(lib):
    assert b"sm_100a" in lib.rfq_version()
    import torch
    if not torch.cuda.is_available():
        from replisense_rfq_b200 import Engine, EngineError
        with pytest.raises(EngineError):
            Engine(0)


def test_error_behaviour_matches_reference(tmp_path):
    """tests/test_file_parser.py:23-40 of the reference: unsupported type and missing file."""
    from replisense_rfq_b200 import FileParser, FileParsingError
    p = tmp_path / "unsupported.xyz"
    p.write_text("This is a test.")
    fp = FileParser()
    with pytest.raises(FileParsingError, match="Unsupported file type"):
        asyncio.run(fp.parse_file_async(str(p)))
    with pytest.raises(FileNotFoundError):
        asyncio.run(fp.parse_file_async("non_existent_file.txt"))
    big = tmp_path / "big.csv"
    big.write_bytes(b"a,b\n" + b"1,2\n" * 300000)
    with pytest.raises(FileParsingError, match="File too large"):
        asyncio.run(FileParser(max_file_size_mb=1).parse_file_async(str(big)))


def test_supported_extensions():
    from replisense_rfq_b200 import FileParser, get_supported_extensions
    assert get_supported_extensions() == {".txt", ".csv", ".xlsx", ".json"}
    assert FileParser().supported_extensions == {".txt", ".csv", ".xlsx", ".json"}
