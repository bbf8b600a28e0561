"""The .json path (app/file_parser.py:315-331): oracle pinned by the reference's own
outputs, the host model of json_pretty.cuh against the oracle, then the CUDA path through the C ABI."""
import base64
import json
from pathlib import Path

import pytest

from oracle import json_oracle as J
from tests import corpus

GOLD = json.loads((Path(__file__).parent / "golden" / "ref_json.json").read_text())


def _oracle(data: bytes):
    try:
        return J.parse_json_bytes(data), None
    except J.OracleParsingError as e:
        return None, str(e)


def test_oracle_reproduces_the_reference():
    """302 documents through the reference's unmodified FileParser._parse_json_async (make_json_goldens.py)."""
    assert len(GOLD) > 250
    for c in GOLD:
        text, err = _oracle(base64.b64decode(c["in"]))
        assert text == c.get("text") and err == c.get("error"), c["in"]


def _model(data: bytes):
    from tests.hostcore import binding as H
    st, t = H.json_pretty(data)
    if st == 0:
        return "=== JSON Data ===\n" + t, None
    if st == 1:
        return None, "Invalid JSON format: " + t
    return None, "declined"


@pytest.mark.parametrize("nested", [False, True])
def test_host_model_matches_goldens_and_fuzz(nested):
    """json_pretty.cuh compiled for the host — the flat state machine the GPU runs, and the nested-loop
    formulation beside it: every golden, 3000 valid and 3000 broken documents; declines only what DESIGN.md
    §4 lists (not UTF-8, duplicate keys, lone surrogates)."""
    from tests.hostcore import binding as H
    H.json_nested(nested)
    try:
        _check_model()
    finally:
        H.json_nested(False)


def _check_model():
    docs = [base64.b64decode(c["in"]) for c in GOLD]
    docs += [corpus.json_doc(s) for s in range(1000, 4000)] + [corpus.json_doc_broken(s) for s in range(1000, 4000)]
    declined = 0
    for d in docs:
        got = _model(d)
        want = _oracle(d)
        if got[1] == "declined":
            declined += 1
            assert want[1] is None or want[1].startswith("JSON parsing failed: 'utf-8' codec"), (d[:80], want[1])
            continue
        assert got == want, (d[:120], got, want)
    assert declined < 0.01 * len(docs)


def test_float_repr_is_shortest_round_trip():
    """Floats written with 16+ digits or extreme exponents go through the exact shortest-digits search."""
    import random
    import struct
    r = random.Random(3)
    vals = [struct.unpack("<d", struct.pack("<Q", r.getrandbits(64)))[0] for _ in range(1500)]
    vals = [v for v in vals if v == v and abs(v) != float("inf")]
    vals += [2.0 ** k for k in range(-1074, 1024, 37)] + [2.0 ** k * (1 - 2 ** -53) for k in range(-1000, 1000, 91)]
    # the magnitudes the 128-bit search covers, with the lopsided intervals at powers of two and their neighbours
    import math
    vals += [r.uniform(0, 1) * 10 ** r.randint(-6, 22) for _ in range(6000)]
    vals += [r.randint(1, 10 ** 7) / 10 ** r.randint(0, 7) for _ in range(3000)]
    for k in range(-20, 72):
        vals += [2.0 ** k, math.nextafter(2.0 ** k, 0.0), math.nextafter(2.0 ** k, math.inf)]
    vals += [9.999999999999999e22, 1e23, 5e-5, 9.5367431640625e-07, 123456789012345680.0, 0.1 + 0.2, 1 / 3, 2 / 3, 1e21, 1e16]
    texts = ["%.17g" % v for v in vals] + ["%.25e" % v for v in vals[:300]] + ["%.16g" % v for v in vals[:300]]
    doc = ("[" + ",".join(t for t in texts if "n" not in t) + "]").encode()
    assert _model(doc) == _oracle(doc)


@pytest.mark.gpu
def test_json_on_gpu_matches_reference_and_oracle(tmp_path):
    import asyncio
    from replisense_rfq_b200 import Engine, FileParser, FileParsingError, build
    build.build()
    eng = Engine(0)
    fp = FileParser(engine=eng)
    docs = [base64.b64decode(c["in"]) for c in GOLD]
    docs += [corpus.json_doc(s) for s in range(5000, 6500)] + [corpus.json_doc_broken(s) for s in range(5000, 6500)]
    big = ("[" + ",".join(corpus.json_doc(s).decode() for s in range(7000, 7400)) + "]").encode()   # one large document
    docs.append(big)
    got = fp.parse_many(docs, ["json"] * len(docs))
    declined = 0
    for i, (g, d) in enumerate(zip(got, docs)):
        text, err = _oracle(d)
        if isinstance(g, FileParsingError) and str(g).startswith("JSON parsing failed: JSON document not decoded on the GPU path"):
            declined += 1
            assert err is None or err.startswith("JSON parsing failed: 'utf-8' codec"), (i, d[:80])
            continue
        if isinstance(g, Exception):
            assert str(g) == err, (i, d[:80], str(g), err)
        else:
            assert g == text, (i, d[:80])
    assert declined < 0.01 * len(docs)
    # and through the file-level entry point, envelope and all
    p = tmp_path / "rfq.json"
    p.write_bytes(docs[len(GOLD) - 1])
    res = asyncio.run(fp.parse_file_async(str(p)))
    assert res["raw_text"] == GOLD[-1].get("text", res["raw_text"]) and res["parsing_method"] == "async_json"
    bad = tmp_path / "bad.json"
    bad.write_bytes(b'{"a": 1,}')
    with pytest.raises(FileParsingError) as ei:
        asyncio.run(fp.parse_file_async(str(bad)))
    assert str(ei.value) == ("Failed to parse bad.json: Invalid JSON format: Expecting property name enclosed in double "
                             "quotes: line 1 column 9 (char 8)")
