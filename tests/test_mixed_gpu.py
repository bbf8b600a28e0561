"""GPU parity: mixed batches (text + CSV + .xlsx in one arena, one upload) through FileParser.parse_many
and rfq_mixed_parse, against the oracle document by document."""
import warnings

import pytest

from oracle import file_parser_oracle as O
from oracle import xlsx_reader as X
from tests import corpus

pytestmark = pytest.mark.gpu


def _want(doc, kind, nrows=1000):
    if kind == "txt":
        return O.parse_txt_bytes(doc)
    if kind == "csv":
        try:
            return O.parse_csv_bytes(doc, nrows)
        except O.OracleParsingError:
            return None
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return O.parse_excel_parts(X.open_workbook(doc), nrows)


def test_mixed_batch_matches_oracle():
    from replisense_rfq_b200 import Engine, FileParser, build
    build.build()
    eng = Engine(0)
    fp = FileParser(engine=eng)
    docs, kinds = corpus.mixed_batch(240, seed=3)
    docs += [b"a\n1\n", b"", b"not a zip"]
    kinds += ["csv", "txt", "xlsx"]
    got = fp.parse_many(docs, kinds)
    assert len(got) == len(docs)
    n_text = 0
    for i, (g, d, k) in enumerate(zip(got, docs, kinds)):
        if d == b"not a zip":
            assert isinstance(g, Exception) and str(g).startswith("Excel parsing failed:")
            continue
        w = _want(d, k)
        if isinstance(g, Exception):
            if k == "csv":
                assert w is None, (i, str(g))
            else:
                assert "not decoded on the GPU path" in str(g), (i, str(g))   # declined sheet construct: loud
            continue
        assert g == w, (i, k)
        n_text += 1
    assert n_text > 230
    # the same batch one kind at a time gives the same texts
    for k in ("txt", "csv", "xlsx"):
        idx = [i for i, kk in enumerate(kinds) if kk == k]
        again = fp.parse_many([docs[i] for i in idx], [k] * len(idx))
        for i, a in zip(idx, again):
            assert (str(a) if isinstance(a, Exception) else a) == (str(got[i]) if isinstance(got[i], Exception) else got[i])
