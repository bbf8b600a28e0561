"""ctypes access to the TEST-ONLY host build of the device core headers (tests/hostcore)."""
import ctypes as C
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent


class HcCell(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("is_na", "int_status", "uint_status", "uint_neg", "uint_gt",
                                         "float_ok", "bool_tok", "inf_tok")] + [
        ("ival", C.c_int64), ("uval", C.c_uint64), ("fval", C.c_double),
        ("nbytes", C.c_uint32), ("ncp", C.c_uint32), ("nesc", C.c_uint32),
        ("utf8_ok", C.c_int32), ("has_fchar", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        so = HERE / "libhostcore.so"
        srcs = [HERE / "hostcore.cpp"] + sorted((HERE.parent.parent / "replisense_rfq_b200/csrc").glob("*.cuh")) + sorted((HERE.parent.parent / "replisense_rfq_b200/csrc").glob("*.h"))
        if not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=off",
                                   "-o", str(so), str(HERE / "hostcore.cpp")])
        _lib = C.CDLL(str(so))
        _lib.hc_txt.restype = C.c_int64
        _lib.hc_txt.argtypes = [C.c_char_p, C.c_int64, C.c_char_p]
        _lib.hc_fixed.argtypes = [C.c_double, C.c_int, C.c_int, C.c_char_p]
        _lib.hc_objfloat.argtypes = [C.c_double, C.c_char_p]
        _lib.hc_sci.argtypes = [C.c_double, C.c_char_p]
        _lib.hc_fixed_info.argtypes = [C.c_double] + [C.POINTER(C.c_int32)] * 3
    return _lib


def parse_cell(b: bytes, csv: bool = True) -> HcCell:
    o = HcCell()
    lib().hc_parse_cell(b, len(b), int(csv), C.byref(o))
    return o


def fixed(v: float, dec: int, space: bool = False) -> str:
    buf = C.create_string_buffer(400)
    n = lib().hc_fixed(v, dec, int(space), buf)
    return buf.raw[:n].decode()


def objfloat(v: float) -> str:
    buf = C.create_string_buffer(400)
    n = lib().hc_objfloat(v, buf)
    assert n > 0, "objfloat_len mismatch"
    return buf.raw[:n].decode()


def sci(v: float) -> str:
    buf = C.create_string_buffer(64)
    n = lib().hc_sci(v, buf)
    assert n > 0, "sci_len mismatch"
    return buf.raw[:n].decode()


def fixed_info(v: float):
    a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
    lib().hc_fixed_info(v, C.byref(a), C.byref(b), C.byref(c))
    return a.value, b.value, c.value


def txt(b: bytes) -> bytes:
    buf = C.create_string_buffer(len(b) + 16)
    n = lib().hc_txt(b, len(b), buf)
    return buf.raw[:n]


def csv_parse(data: bytes, nrows: int = 1000):
    """Sequential model of the GPU CSV pipeline. Returns text, None (unparseable) or 'UNSUPPORTED'."""
    cap = len(data) * 12 + 65536
    buf = C.create_string_buffer(cap)
    n = C.c_uint64()
    L = lib()
    L.hc_csv_parse.argtypes = [C.c_char_p, C.c_uint64, C.c_int64, C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64)]
    st = L.hc_csv_parse(data, len(data), nrows, buf, cap, C.byref(n))
    if st == 0:
        return buf.raw[:n.value].decode("utf-8")
    if st == 1:
        return None
    if st == 2:
        return "UNSUPPORTED"
    raise RuntimeError("hostcore buffer too small")


def sheet_parse(sheet_xml: bytes, sst_xml, date_ids=(), nrows: int = 1000, td_ids=(), date1904=False):
    """Sequential model of the GPU sheet pipeline: (status, text). status 0 ok, 2 unsupported,
    3 empty sheet."""
    import numpy as np
    cap = len(sheet_xml) * 8 + 65536
    buf = C.create_string_buffer(cap)
    n = C.c_uint64()
    L = lib()
    L.hc_sheet_parse.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                 C.c_int, C.c_int64, C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64)]
    d = np.asarray(sorted(int(i) | (0x80000000 if i in td_ids else 0) for i in date_ids), dtype=np.uint32)
    st = L.hc_sheet_parse(sheet_xml, len(sheet_xml), sst_xml, len(sst_xml) if sst_xml else 0,
                          d.ctypes.data if len(d) else None, len(d), 1 if date1904 else 0, nrows, buf, cap, C.byref(n))
    return st, (buf.raw[:n.value].decode("utf-8") if st == 0 else None)


def set_fused(on: bool) -> None:
    """Switch the sequential CSV model between the two-sweep form (classify, then measure under the
    decided type) and the one-sweep candidates the GPU classify kernel reduces."""
    lib().hc_set_fused(int(on))


def fused_used() -> int:
    L = lib()
    L.hc_fused_used.restype = C.c_uint64
    return int(L.hc_fused_used())


def set_walk(on: bool) -> None:
    """Cells found by the row walker (csv2_rows.cuh scan_field) instead of the field index; the hot
    kernels' per-cell shortcut is cross-checked against the general functions while it runs."""
    lib().hc_set_walk(int(on))


def walk_stats():
    """(walker/index disagreements, hot/general typing disagreements, cells typed hot, cells left cold,
    hot/general printing disagreements, cells printed hot, cells left to the general formatters)"""
    import numpy as np
    o = np.zeros(7, dtype=np.uint64)
    lib().hc_walk_stats(C.c_void_p(o.ctypes.data))
    return tuple(int(x) for x in o)


def float_repr(v: float) -> str:
    """py_float_repr of header_names.h (labels of sheet columns)."""
    L = lib()
    L.hc_float_repr.argtypes = [C.c_double, C.c_char_p]
    buf = C.create_string_buffer(64)
    n = L.hc_float_repr(v, buf)
    return buf.raw[:n].decode()


def timedelta_us(v: float) -> int:
    L = lib()
    L.hc_timedelta_us.argtypes = [C.c_double]
    L.hc_timedelta_us.restype = C.c_longlong
    return L.hc_timedelta_us(v)


def timedelta_str(us: int) -> str:
    L = lib()
    L.hc_timedelta_str.argtypes = [C.c_longlong, C.c_char_p]
    buf = C.create_string_buffer(64)
    n = L.hc_timedelta_str(us, buf)
    return buf.raw[:n].decode()


JSON_ERRORS = {1: "Expecting value", 2: "Expecting property name enclosed in double quotes", 3: "Expecting ':' delimiter",
               4: "Expecting ',' delimiter", 5: "Extra data", 6: "Unterminated string starting at",
               7: "Invalid control character at", 8: "Invalid \\escape", 9: "Invalid \\uXXXX escape",
               10: "Unexpected UTF-8 BOM (decode using utf-8-sig)"}


def json_pretty(doc: bytes):
    """(status, text or error message): json_pretty of json_pretty.cuh, measure pass + write pass."""
    L = lib()
    L.hc_json_pretty.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
    cap = 6 * len(doc) + 64
    buf = C.create_string_buffer(cap)
    n = C.c_uint64(0)
    info = (C.c_uint32 * 4)()
    st = L.hc_json_pretty(doc, len(doc), buf, cap, C.byref(n), info)
    if st == 0:
        return 0, buf.raw[:n.value].decode("utf-8")
    if st == 1:
        return 1, "%s: line %d column %d (char %d)" % (JSON_ERRORS[info[0]], info[2], info[3], info[1])
    return st, None


def json_nested(on: bool) -> None:
    """Select the nested-loop formulation of json_pretty.cuh (default: the flat state machine the GPU runs)."""
    L = lib()
    L.hc_json_nested.argtypes = [C.c_int]
    L.hc_json_nested(1 if on else 0)
