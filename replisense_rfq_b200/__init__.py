"""B200-native document→text engine behind the `FileParser` interface of Replisense-rfq."""
from .file_parser import (FileParser, FileParsingError, HAS_PYMUPDF, get_supported_extensions,
                          parse_file_async)
from .engine import Engine, EngineError, default_engine
from .rfq_agent import RFQFieldGenerator, StubLLM, extract_json_many

__all__ = ["FileParser", "FileParsingError", "HAS_PYMUPDF", "get_supported_extensions",
           "parse_file_async", "Engine", "EngineError", "default_engine", "RFQFieldGenerator", "StubLLM",
           "extract_json_many"]
