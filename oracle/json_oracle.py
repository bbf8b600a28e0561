"""TEST INFRASTRUCTURE ONLY (never imported by the product path): the reference's .json parser restated.

`_parse_json_async` (/root/reference/app/file_parser.py:315-331):

    async with aiofiles.open(file_path, 'r', encoding='utf-8') as f:      # text mode: strict UTF-8,
        content = await f.read()                                          # universal newlines
    data = json.loads(content)
    formatted_json = json.dumps(data, indent=2, ensure_ascii=False)
    return f"=== JSON Data ===\\n{formatted_json}"
    except json.JSONDecodeError as e:  raise FileParsingError(f"Invalid JSON format: {str(e)}")
    except Exception as e:             raise FileParsingError(f"JSON parsing failed: {str(e)}")

The engine behind it is CPython's own `json` module, which is what runs here too.  Pinned by
tests/golden/ref_json.json: cases run through the reference's unmodified FileParser.parse_file_async
(generator tests/golden/make_json_goldens.py).
"""
from __future__ import annotations

import io
import json


class OracleParsingError(Exception):
    pass


def read_text(data: bytes) -> str:
    """open(path, 'r', encoding='utf-8').read() on these bytes."""
    return io.TextIOWrapper(io.BytesIO(data), encoding="utf-8", newline=None).read()


def parse_json_bytes(data: bytes) -> str:
    try:
        content = read_text(data)
        obj = json.loads(content)
        return "=== JSON Data ===\n" + json.dumps(obj, indent=2, ensure_ascii=False)
    except json.JSONDecodeError as e:
        raise OracleParsingError(f"Invalid JSON format: {str(e)}")
    except Exception as e:
        raise OracleParsingError(f"JSON parsing failed: {str(e)}")
