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


def test_large_mixed_batch_sheets_on_the_second_context():
    """From 256 sheet parts on, rfq_mixed_run gives the sheet documents to a second context on a host thread
    beside the CSV run (mixed_api.cu): one arena with text, CSV, sharedStrings and sheet XML documents through
    rfq_mixed_parse — every text equals the oracle's, and a second call gives the same again."""
    import random
    from replisense_rfq_b200 import Engine, build, _lib
    build.build()
    eng = Engine(0)
    r = random.Random(21)
    docs, kind, sst_of, want = [], [], [], []
    for i in range(1500):
        x = r.random()
        if x < 0.4:
            d = corpus.rfq_csv(r.choice([10, 25, 60, 150, 400]), seed=7000 + i)
            docs.append(d); kind.append(_lib.KIND_CSV); sst_of.append(-1)
            want.append(O.parse_csv_bytes(d, 1000))
        elif x < 0.7:
            sx, ss = corpus.sheet_xml_doc(7000 + i) if i % 4 else corpus.sheet_xml_fuzz(7000 + i)
            docs += [ss, sx]; kind += [_lib.KIND_SST, _lib.KIND_SHEET]; sst_of += [-1, len(docs) - 2]
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                df = X.table_to_frame(X.sheet_table(sx, X.read_string_table(ss), nrows=1000), 1000)
            want += [None, None if df.empty else O.frame_to_text(df)]
        else:
            d = corpus.text_body(7000 + i)
            docs.append(d); kind.append(_lib.KIND_TEXT); sst_of.append(-1)
            want.append(O.parse_txt_bytes(d))
    assert sum(k in (_lib.KIND_SST, _lib.KIND_SHEET) for k in kind) >= 600
    texts = []
    for rep in range(2):
        with eng.lock:
            res = eng.mixed_batch(docs, kind, 1000, sst_of)
            got = [(int(res.status[d]), res.doc_str(d) if int(res.status[d]) == 0 else None) for d in range(res.n)]
        texts.append(got)
    assert texts[0] == texts[1]
    declined = 0
    for d, ((st, g), w, k) in enumerate(zip(texts[0], want, kind)):
        if k == _lib.KIND_SST:
            continue
        if st == _lib.DOC_UNSUPPORTED:
            declined += 1
            continue
        if k == _lib.KIND_SHEET and w is None:
            assert st == _lib.DOC_EMPTY_SHEET, (d, st)
            continue
        assert st == 0 and g == w, (d, k, st)
    assert declined <= 0.02 * len(docs)
