"""GPU parity: text path vs the oracle / the reference's committed outputs, through the C ABI."""
import base64
import json
from pathlib import Path

import pytest

from oracle import file_parser_oracle as O
from tests import corpus

pytestmark = pytest.mark.gpu
G = Path(__file__).parent / "golden"


@pytest.fixture(scope="module")
def eng():
    from replisense_rfq_b200 import Engine, build
    build.build()
    return Engine(0)


def test_txt_reference_goldens(eng):
    cases = json.loads((G / "ref_txt.json").read_text())
    docs = [base64.b64decode(c["data"]) for c in cases]
    got = eng.parse_text_many(docs)
    bad = [c["name"] for c, g in zip(cases, got) if g != c["text"]]
    assert not bad, bad


def test_txt_cache_bodies_roundtrip(eng):
    rows = json.loads((G / "cache_rows.json").read_text())
    docs = [rows[k].encode("utf-8") for k in ("2", "3", "5", "6", "7", "13")]
    docs += [b"\r\n  " + d.replace(b"\n", b"\r\n") + b" \t\r\n" for d in docs]
    got = eng.parse_text_many(docs)
    assert got == [O.parse_txt_bytes(d) for d in docs]


def test_txt_fuzz_batch(eng):
    docs = [corpus.text_body(s) for s in range(2000, 4000)]
    got = eng.parse_text_many(docs)
    want = [O.parse_txt_bytes(d) for d in docs]
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, bad[:10]


def test_txt_edge_sizes(eng):
    docs = [b"", b"x", b" ", b"a" * 511, b"b" * 512, b"c" * 513, b" " * 512 + b"z" + b" " * 1024,
            ("é" * 300).encode(), b"\xe2\x82" * 400, b"\r\n" * 700,
            b"x" + b"\r" + b"\xff" * 600 + b"\ny"]
    docs.append((" \n" * 5000 + "body ₹ line\r\n" * 50000 + "\t " * 3000).encode())
    got = eng.parse_text_many(docs)
    want = [O.parse_txt_bytes(d) for d in docs]
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, bad


def test_txt_idempotent_large(eng):
    """Size-independent property at scale: normalising a normalised body changes nothing."""
    docs = [corpus.text_body(s) for s in range(20000)]
    once = eng.parse_text_many(docs)
    twice = eng.parse_text_many([t.encode("utf-8") for t in once])
    assert once == twice
