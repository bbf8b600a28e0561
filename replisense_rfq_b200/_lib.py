"""ctypes binding of the C ABI declared in include/rfq_parse.h.

There is no CPU fallback: if the CUDA library is missing or no device is present, loading or
creating an engine raises — callers never get a silently slower path.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "librfqparse.so"

RFQ_OK = 0
DOC_OK, DOC_CSV_UNPARSEABLE, DOC_UNSUPPORTED, DOC_EMPTY_SHEET, DOC_EMPTY_TEXT, DOC_VALUE_ERROR, DOC_CORRUPT = 0, 1, 2, 3, 4, 5, 6
DOC_JSON_INVALID = 7
# RFQ_JSON_E_*: the messages of json.JSONDecodeError (Lib/json/decoder.py, Modules/_json.c)
JSON_ERRORS = {1: "Expecting value", 2: "Expecting property name enclosed in double quotes", 3: "Expecting ':' delimiter",
               4: "Expecting ',' delimiter", 5: "Extra data", 6: "Unterminated string starting at",
               7: "Invalid control character at", 8: "Invalid \\escape", 9: "Invalid \\uXXXX escape",
               10: "Unexpected UTF-8 BOM (decode using utf-8-sig)"}

_lib = None

# name -> (restype, argtypes); mirrors include/rfq_parse.h one to one
SIGNATURES = {
    "rfq_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "rfq_destroy": (None, [C.c_void_p]),
    "rfq_last_error": (C.c_char_p, [C.c_void_p]),
    "rfq_version": (C.c_char_p, []),
    "rfq_device_count": (C.c_int, []),
    "rfq_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rfq_get_stream": (C.c_void_p, [C.c_void_p]),
    "rfq_host_alloc": (C.c_void_p, [C.c_void_p, C.c_uint64]),
    "rfq_host_free": (None, [C.c_void_p, C.c_void_p]),
    "rfq_stage": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "rfq_text_run": (C.c_int, [C.c_void_p]),
    "rfq_csv_run": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rfq_fetch": (C.c_int, [C.c_void_p]),
    "rfq_md5_run": (C.c_int, [C.c_void_p]),
    "rfq_numeric_run": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rfq_json_spans_run": (C.c_int, [C.c_void_p]),
    "rfq_json_run": (C.c_int, [C.c_void_p]),
    "rfq_out_plan_many": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "rfq_out_widen_many": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "rfq_frame_run": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "rfq_inflate_run": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "rfq_restage_output": (C.c_int, [C.c_void_p]),
    "rfq_fetch_tables": (C.c_int, [C.c_void_p]),
    "rfq_prompt_run": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_char_p,
                                 C.c_uint32, C.c_uint32]),
    "rfq_prompt_build": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_char_p,
                                   C.c_uint32, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_uint32]),
    "rfq_sheets_run": (C.c_int, [C.c_void_p] * 6),
    "rfq_sheets_parse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32] + [C.c_void_p] * 5),
    "rfq_text_parse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "rfq_csv_parse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                C.c_void_p]),
    "rfq_csv_prompt_run": (C.c_int, [C.c_void_p, C.c_void_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32,
                                     C.c_char_p, C.c_uint32, C.c_uint32]),
    "rfq_csv_prompt_parse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                       C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32,
                                       C.c_uint32]),
    "rfq_out_raw_bytes": (C.c_void_p, [C.c_void_p]),
    "rfq_out_raw_chars": (C.c_void_p, [C.c_void_p]),
    "rfq_out_chars": (C.c_void_p, [C.c_void_p]),
    "rfq_out_maxchar": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]),
    "rfq_out_widen": (C.c_int, [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_uint64]),
    "rfq_mixed_run": (C.c_int, [C.c_void_p] * 6),
    "rfq_mixed_parse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32] + [C.c_void_p] * 5),
    "rfq_out_count": (C.c_uint32, [C.c_void_p]),
    "rfq_out_text": (C.c_void_p, [C.c_void_p]),
    "rfq_out_offsets": (C.c_void_p, [C.c_void_p]),
    "rfq_out_lengths": (C.c_void_p, [C.c_void_p]),
    "rfq_out_status": (C.c_void_p, [C.c_void_p]),
    "rfq_out_total_bytes": (C.c_uint64, [C.c_void_p]),
    "rfq_out_rows": (C.c_void_p, [C.c_void_p]),
    "rfq_out_cols": (C.c_void_p, [C.c_void_p]),
    "rfq_out_attempt": (C.c_void_p, [C.c_void_p]),
    "rfq_last_run_ms": (C.c_float, [C.c_void_p]),
    "rfq_launch_count": (C.c_uint64, [C.c_void_p]),
    "rfq_set_profile": (C.c_int, [C.c_void_p, C.c_int]),
    "rfq_profile_count": (C.c_int, [C.c_void_p]),
    "rfq_profile_name": (C.c_char_p, [C.c_void_p, C.c_int]),
    "rfq_profile_ms": (C.c_float, [C.c_void_p, C.c_int]),
    "rfq_text_normalise_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                             C.c_void_p]),
}
PART_ALWAYS, PART_IF_TEXT, PART_FALLBACK = 0, 1, 2
KIND_TEXT, KIND_CSV, KIND_SHEET, KIND_SHEET_1904, KIND_SST = 0, 1, 2, 3, 4
NOALLOC_MAX_BYTES = 4 << 20
ERR_TOO_LARGE = -4


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m replisense_rfq_b200.build` "
            "(there is no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export the symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
