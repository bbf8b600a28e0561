"""Golden fixtures for the .json path: inputs (base64) and what the reference's UNMODIFIED
FileParser._parse_json_async (/root/reference/app/file_parser.py:315-331) returns or raises for them.

Run in the AUTHORING container only (needs /root/reference, through oracle/_ref built by oracle/build_ref.py):
    python tests/golden/make_json_goldens.py
Writes tests/golden/ref_json.json.  Nothing of the reference is copied: only its outputs.
"""
from __future__ import annotations

import asyncio
import base64
import json
import os
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))

from oracle import build_ref, ref_runner  # noqa: E402
from tests import corpus  # noqa: E402

HAND = ['', ' ', '{', '{"a"', '{"a":', '{"a":1', '{"a":1,', '{"a":1,}', '[1,]', '[', '[1', '[1 2]', '{"a" 1}', '{1:2}',
        '{"a":1 "b":2}', 'nul', 'tru', '-', '-x', '1.', '1.e5', '01', '1e', '1e+', '"abc', '"a\\', '"a\\x"', '"a\\u12"',
        '"a\\u12G4"', '"a\tb"', '"\\ud83d\\ude00"', '﻿{}', '{} x', '[1,,2]', '{"a":}', '[}', '{]', 'NaN', '-Infinity',
        'Infinity', '-NaN', '+1', '.5', '1e5', '1E5', '-0', '-0.0', '1e999', '-1e999', '1e-999', '"\\u00e9\\n\\/"',
        '[1.0, 2.50, 1e15, 1e16, 1.5e-5, 0.0001, 123456789012345678, 0.1e1]', '{"a":{"b":[]},"c":{}}', '"\x7f"',
        '[nul]', '"a\nb"', '[1,\n 2 x]', '{"é":1 x}',
        '[0.30000000000000004, 0.1000000000000000055511151231257827, 1.7976931348623157e308, 5e-324, '
        '2.2250738585072014e-308, 123456789.123456789, 9007199254740993.0]',
        '{"a":\r\n [1,\r 2, x]}', '"\\u0000\\u001f\\b\\f"',
        '{"rfq": {"part_number": "RJF544", "quantity": 1200, "target_price": 1250.75, "currency": "₹", '
        '"lines": [{"sl": 1, "desc": "Bearing 6204-2RS", "qty": 40}, {"sl": 2, "desc": "Seal \\"V\\" 25mm", "qty": 12.50}]}}']


def main():
    assert build_ref.build(), "reference not available"
    m = ref_runner.module()
    fp = m.FileParser()
    cases = [h.encode("utf-8") for h in HAND]
    cases += [corpus.json_doc(s) for s in range(120)] + [corpus.json_doc_broken(s) for s in range(200, 320)]
    out = []
    d = "/dev/shm" if os.path.isdir("/dev/shm") else None
    for data in cases:
        fd, path = tempfile.mkstemp(suffix=".json", dir=d)
        try:
            with os.fdopen(fd, "wb") as f:
                f.write(data)
            try:
                out.append({"in": base64.b64encode(data).decode(), "text": asyncio.run(fp._parse_json_async(Path(path)))})
            except m.FileParsingError as e:
                out.append({"in": base64.b64encode(data).decode(), "error": str(e)})
        finally:
            os.unlink(path)
    (HERE / "ref_json.json").write_text(json.dumps(out, ensure_ascii=True, indent=0))
    print(len(out), "cases,", sum("error" in c for c in out), "errors")


if __name__ == "__main__":
    main()
