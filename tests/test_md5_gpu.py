"""Batched MD5 file hashes (app/file_parser.py:122-130) and the batched envelope."""
import asyncio
import hashlib
import random

import pytest

from tests import corpus

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from replisense_rfq_b200 import Engine, build
    build.build()
    return Engine(0)


def test_md5_matches_hashlib(eng):
    r = random.Random(3)
    docs = [b"", b"a", b"abc", b"message digest", b"a" * 55, b"a" * 56, b"a" * 57, b"a" * 63, b"a" * 64,
            b"a" * 65, b"a" * 119, b"a" * 120, b"a" * 128, bytes(range(256)) * 40]
    docs += [r.randbytes(r.randint(0, 5000)) for _ in range(3000)]
    docs += [r.randbytes(300000)]
    got = eng.md5_many(docs)
    want = [hashlib.md5(d).hexdigest() for d in docs]
    assert got == want
    # RFC 1321 test suite
    assert got[0] == "d41d8cd98f00b204e9800998ecf8427e" and got[2] == "900150983cd24fb0d6963f7d28e17f72"


def test_envelopes_match_parse_file_async(eng, tmp_path):
    from replisense_rfq_b200 import FileParser, FileParsingError
    fp = FileParser(engine=eng)
    docs = [corpus.rfq_csv(20, seed=1), "Dear supplier,\r\n qty: 750 \n".encode(), b"a;b\n1;2\n", b"junk", b"x,y\n1,2\n"]
    names = ["rfq.csv", "mail.txt", "semi.CSV", "notes.docx", "dir/rfq2.csv"]
    got = fp.parse_many_envelopes(docs, names)
    for d, nm, g in zip(docs, names, got):
        p = tmp_path / nm.split("/")[-1]
        p.write_bytes(d)
        try:
            want = asyncio.run(FileParser(engine=eng).parse_file_async(str(p)))
        except FileParsingError as e:
            assert isinstance(g, FileParsingError) and str(g) == str(e), (nm, g, e)
            continue
        assert g == want, nm
