"""GPU parity: device-side DEFLATE (the zip step of app/file_parser.py:225) against
zlib, member by member; then real .xlsx containers end to end through FileParser."""
import io
import random
import warnings
import zipfile
import zlib

import pytest

from oracle import file_parser_oracle as O
from oracle import xlsx_reader as X
from tests import corpus

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from replisense_rfq_b200 import Engine, build
    build.build()
    return Engine(0)


def _raw_deflate(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
    return c.compress(data) + c.flush()


def test_members_inflate_like_zlib(eng):
    """Stored, fixed-Huffman and dynamic-Huffman streams, every compression level, empty and
    incompressible members, multi-block members; corrupted streams, sizes and CRCs are refused."""
    from replisense_rfq_b200.xlsx_host import Member
    r = random.Random(2)
    plain = [b"", b"a", b"hello hello hello hello", bytes(r.getrandbits(8) for _ in range(70000)),
             corpus.rfq_csv(3000, seed=1), corpus.sheet_xml_doc(3, 50, 400)[0], corpus.sheet_xml_doc(4)[1],
             b"\x00" * 200000, ("é₹日" * 5000).encode(), corpus.text_body(5) * 40]
    members, want = [], []
    for p in plain:
        for level, strat in ((0, zlib.Z_DEFAULT_STRATEGY), (1, zlib.Z_DEFAULT_STRATEGY), (6, zlib.Z_DEFAULT_STRATEGY),
                             (9, zlib.Z_DEFAULT_STRATEGY), (6, zlib.Z_FIXED), (6, zlib.Z_HUFFMAN_ONLY), (6, zlib.Z_RLE)):
            members.append(Member(_raw_deflate(p, level, strat), 8, len(p), zlib.crc32(p)))
            want.append(p)
        members.append(Member(p, 0, len(p), zlib.crc32(p)))      # stored in the zip sense
        want.append(p)
    got = eng.inflate_many(members)
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, bad[:10]
    # refused, never guessed
    p = plain[4]
    good = _raw_deflate(p)
    flipped = bytearray(good); flipped[len(good) // 2] ^= 0x40
    broken = [Member(bytes(flipped), 8, len(p), zlib.crc32(p)), Member(good[:-7], 8, len(p), zlib.crc32(p)),
              Member(good, 8, len(p) + 1, zlib.crc32(p)), Member(good, 8, len(p) - 1, zlib.crc32(p)),
              Member(good, 8, len(p), zlib.crc32(p) ^ 1), Member(good, 12, len(p), zlib.crc32(p)),
              Member(b"\x07", 8, 5, 0), Member(good, 8, len(p), zlib.crc32(p))]
    got = eng.inflate_many(broken)
    assert got[:-1] == [None] * (len(broken) - 1) and got[-1] == p


def test_xlsx_containers_inflated_on_the_device(eng, tmp_path):
    """Whole .xlsx files (zip + DEFLATE at several levels, one stored): FileParser output equals the
    oracle's over the host-inflated parts; every member's inflated bytes equal zipfile's."""
    import asyncio
    from replisense_rfq_b200 import FileParser, xlsx_host
    fp = FileParser(engine=eng)
    books = []
    for s in range(40):
        sheets, ssts = [], None
        sx, ss = corpus.sheet_xml_doc(s, 5, 200)
        books.append(corpus.make_xlsx([sx], ss, level=[1, 6, 9, 0][s % 4]))
    got = fp.parse_many(books, ["xlsx"] * len(books))
    for i, (g, b) in enumerate(zip(got, books)):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            want = O.parse_excel_parts(X.open_workbook(b), 1000)
        assert (str(g) if isinstance(g, Exception) else g) == want, i
        wb = xlsx_host.scan_workbook(b)
        z = zipfile.ZipFile(io.BytesIO(b))
        inflated = eng.inflate_many(wb.sheets + [wb.sst])
        assert inflated[0] == z.read("xl/worksheets/sheet1.xml") and inflated[1] == z.read("xl/sharedStrings.xml")
    p = tmp_path / "book.xlsx"
    p.write_bytes(books[1])
    res = asyncio.run(fp.parse_file_async(str(p)))
    assert res["raw_text"] == got[1] and res["parsing_method"] == "async_xlsx"
    # a damaged container is an error, as zipfile's BadZipFile / zlib.error is in the reference
    bad = bytearray(books[2])
    wb = xlsx_host.scan_workbook(bytes(bad))
    off = bytes(bad).find(bytes(wb.sheets[0].raw[:16]))
    bad[off + len(wb.sheets[0].raw) // 2] ^= 0xFF
    g = fp.parse_many([bytes(bad)], ["xlsx"])[0]
    assert isinstance(g, Exception) and str(g).startswith("Excel parsing failed:")
