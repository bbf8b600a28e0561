"""GPU parity: CSV path vs the reference's committed outputs and vs the oracle (pandas), through
the C ABI.  Bit-exact or it fails."""
import base64
import json
from pathlib import Path

import pytest

from oracle import file_parser_oracle as O
from tests import corpus

pytestmark = pytest.mark.gpu
G = Path(__file__).parent / "golden"

UNSUPPORTED_GOLDENS = set()


@pytest.fixture(scope="module")
def eng():
    from replisense_rfq_b200 import Engine, build
    build.build()
    return Engine(0)


def oracle(data, nrows=1000):
    try:
        return O.parse_csv_bytes(data, nrows)
    except O.OracleParsingError:
        return None


def test_csv_reference_goldens(eng):
    cases = [c for c in json.loads((G / "ref_csv.json").read_text())
             if c["name"] not in UNSUPPORTED_GOLDENS]
    docs = [base64.b64decode(c["data"]) for c in cases]
    got = eng.parse_csv_many(docs, [c["nrows"] for c in cases], on_unsupported="return")
    bad, declined = [], []
    for c, g, d in zip(cases, got, docs):
        if isinstance(g, Exception):
            # the one tokenizer case that is declined instead of reproduced: a blank-first line after a
            # bare CR (pandas re-reads earlier rows there); anything else declined is a failure
            (declined if corpus.bare_cr_blank_first(d) else bad).append(c["name"])
        elif g != c.get("text"):
            bad.append(c["name"])
    assert not bad, bad
    assert len(declined) <= 8, declined


def test_csv_unsupported_is_loud(eng):
    from replisense_rfq_b200 import EngineError
    wide = (",".join("c%d" % i for i in range(1100)) + "\n" + ",".join("1" for _ in range(1100)) + "\n").encode()
    with pytest.raises(EngineError):   # > 1024 columns: reported, never routed to a CPU path
        eng.parse_csv_many([wide], 1000)


def test_csv_big_integer_columns(eng):
    """int64 overflow -> uint64 pass -> Python-int objects / strings, in pandas' order-dependent way."""
    cols = [["18446744073709551616", "1"], ["1", "18446744073709551616"], ["18446744073709551616", "NA", "2"],
            ["18446744073709551616", " 7 ", "+3", "-0", "007"], ["18446744073709551616", "1.5"],
            ["-9223372036854775809", "5"], ["-9223372036854775809", "9223372036854775808"],
            ["99999999999999999999999999999999999999", "-5"], ["1_000", "18446744073709551616"],
            ["18446744073709551616", "1_000"], ["9223372036854775808", "-1", "18446744073709551616"],
            ["9223372036854775808", "abc", "18446744073709551616"], ["9223372036854775808", "1"],
            ["9223372036854775808", ""], ["18446744073709551616", "1__0"]]
    docs = [("a,b\n" + "".join(f"{v},x\n" for v in c)).encode() for c in cols]
    got = eng.parse_csv_many(docs, 1000)
    bad = [c for c, g, d in zip(cols, got, docs) if g != oracle(d)]
    assert not bad, bad


def test_csv_fuzz_batch(eng):
    docs = [corpus.fuzz_csv(s) for s in range(30000, 31500)]
    got = eng.parse_csv_many(docs, 1000, on_unsupported="return")
    bad, declined = [], 0
    for i, (g, d) in enumerate(zip(got, docs)):
        if isinstance(g, Exception) and corpus.bare_cr_blank_first(d):
            declined += 1
        elif g != oracle(d):
            bad.append(30000 + i)
    assert not bad, bad[:20]
    assert declined <= 45, declined    # < 3 % of the corpus: bare-CR files with a blank-first line


def test_csv_one_by_one_equals_batched(eng):
    docs = [corpus.fuzz_csv(s) for s in range(40000, 40060)]
    norm = lambda xs: ["declined" if isinstance(x, Exception) else x for x in xs]
    batched = norm(eng.parse_csv_many(docs, 1000, on_unsupported="return"))
    single = norm([eng.parse_csv_many([d], 1000, on_unsupported="return")[0] for d in docs])
    assert batched == single


def test_csv_rfq_sizes(eng):
    for n, cap in [(1, 1000), (7, 1000), (300, 1000), (2500, 1000), (2500, 10 ** 6), (70000, 10 ** 6)]:
        d = corpus.rfq_csv(n, seed=n)
        assert eng.parse_csv_many([d], cap)[0] == oracle(d, cap), (n, cap)


def test_csv_low_memory_chunk_boundary(eng):
    """A non-numeric token after the first buffer_lines rows: earlier chunks were already
    converted ('1.50' prints 1.5), SURVEY.md App. C."""
    for ncols, k in [(2, 300000), (8, 65536), (8, 65535), (16, 40000)]:
        rows = ["1.50" + ",x" * (ncols - 1)] * (k + 3)
        rows[k] = "abc" + ",x" * (ncols - 1)
        d = (",".join("c%d" % i for i in range(ncols)) + "\n" + "\n".join(rows) + "\n").encode()
        assert eng.parse_csv_many([d], 10 ** 7)[0] == oracle(d, 10 ** 7), (ncols, k)


def test_csv_crlf_and_wide(eng):
    d1 = corpus.rfq_csv(5000, seed=5, newline="\r\n")
    wide = (",".join("c%d" % i for i in range(300)) + "\n" +
            "\n".join(",".join(str((r * 7 + c) % 1000 / 4) for c in range(300)) for r in range(50)) + "\n").encode()
    for d in (d1, wide):
        assert eng.parse_csv_many([d], 10 ** 6)[0] == oracle(d, 10 ** 6)


def test_csv_widest_tables(eng):
    """1000 and 1024 columns (the one-sweep kernel needs its > 48 KB shared-memory opt-in there), mixed
    column types, next to a narrow table in the same batch; 1025 columns are declined loudly."""
    def table(w, nrows):
        kinds = ["int", "float", "str", "bool", "na"]
        def cell(r, c):
            k = kinds[c % 5]
            return {"int": str(r * c % 977), "float": "%.3f" % (r * 1.25 + c), "str": "p%d-%d" % (r, c),
                    "bool": "True" if (r + c) % 3 else "False", "na": "" if (r + c) % 2 else "7"}[k]
        return (",".join("c%d" % i for i in range(w)) + "\n" +
                "\n".join(",".join(cell(r, c) for c in range(w)) for r in range(nrows)) + "\n").encode()
    docs = [table(1000, 40), corpus.rfq_csv(50, seed=2), table(1024, 3)]
    got = eng.parse_csv_many(docs, 1000)
    for g, d in zip(got, docs):
        assert g == oracle(d, 1000)
    from replisense_rfq_b200 import EngineError
    with pytest.raises(EngineError):
        eng.parse_csv_many([table(1025, 2)], 1000)


def test_csv_file_parser_api(eng, tmp_path):
    from replisense_rfq_b200 import FileParser, FileParsingError
    import asyncio
    fp = FileParser(engine=eng)
    p = tmp_path / "ok.csv"
    data = corpus.rfq_csv(50, seed=9)
    p.write_bytes(data)
    res = asyncio.run(fp.parse_file_async(str(p)))
    assert res["raw_text"] == oracle(data) and res["parsing_method"] == "async_csv"
    assert res["source_file"] == "ok.csv" and res["file_size"] == len(data) and len(res["file_hash"]) == 32
    bad = tmp_path / "bad.csv"
    bad.write_bytes(b"a\n1\n2\n")
    with pytest.raises(FileParsingError, match="Failed to parse bad.csv: CSV parsing failed: Unable to parse CSV"):
        asyncio.run(fp.parse_file_async(str(bad)))


def test_csv_c2_full_size_bit_exact(eng):
    """BASELINE config C2 at its full size (1 M rows x 8 columns, 16 low_memory chunks): the GPU
    text equals the oracle's byte for byte (about a minute of pandas on one host core)."""
    d = corpus.rfq_csv(1_000_000, seed=0)
    got = eng.parse_csv_many([d], 10 ** 6)[0]
    want = oracle(d, 10 ** 6)
    assert len(got) == len(want)
    assert got == want


def test_engine_pool_concurrent_calls_keep_their_results():
    """The reference shares ONE FileParser between all requests behind a 4-thread pool
    (app/main.py:47,59, app/file_parser.py:33).  Many threads on a small engine pool, documents from a
    few hundred bytes to several MB, many rounds: every call must get its own document's text (a result
    is turned into Python objects while its engine is still locked, whatever overlaps with it)."""
    from concurrent.futures import ThreadPoolExecutor
    from replisense_rfq_b200 import FileParser, build
    from replisense_rfq_b200 import engine as E
    build.build()
    docs = [corpus.rfq_csv(n, seed=i) for i, n in enumerate([40, 60000, 300, 25000, 7, 45000, 1500, 90])]
    docs += [corpus.fuzz_csv(s) for s in range(16)]
    want = [oracle(d, 10 ** 6) for d in docs]
    norm = lambda x: "declined" if isinstance(x, Exception) and not isinstance(x, AssertionError) else x
    pool = E.EnginePool(0, depth=3)
    try:
        for _ in range(2):
            got = pool.map(lambda e, d: norm(e.parse_csv_many([d], 10 ** 6, on_unsupported="return")[0]), docs)
            assert [g for g in got if g != "declined"] == [w for g, w in zip(got, want) if g != "declined"]
    finally:
        pool.close()
    # the shared parser: 16 threads over a pool of depth 2
    E._default_pools.pop(0, None)
    E.default_pool(0, depth=2)
    fp = FileParser(max_file_size_mb=4096, max_excel_rows=10 ** 6)
    order = [i for r in range(6) for i in range(len(docs))]
    with ThreadPoolExecutor(16) as tp:
        got = list(tp.map(lambda i: fp.parse_many([docs[i]], ["csv"])[0], order))
    n_declined = 0
    for i, g in zip(order, got):
        if isinstance(g, Exception) and "unsupported construct" in str(g):
            n_declined += 1
            assert corpus.bare_cr_blank_first(docs[i])
            continue
        assert (None if isinstance(g, Exception) else g) == want[i], i
    assert n_declined <= 6 * 2


def test_csv_decimal_literals(eng):
    """Short decimal literals (prices) take a digits-only path in the typing sweep and in emit: every
    shape around its limits (8 / 9 integer digits, 6 / 7 decimals, 1e6, 1e-5, signs, zeros) against
    pandas, in fixed, object-float and mixed columns."""
    import random
    r = random.Random(17)
    docs = []
    for t in range(120):
        rows = []
        for i in range(r.randint(2, 40)):
            ip = r.choice([0, 0, 1, 9, 12, 999, 123456, 999999, 1000000, 1000001, 99999999, 123456789])
            nd = r.randint(1, 7)
            fr = "".join(r.choice("0123456789") for _ in range(nd))
            if r.random() < 0.2:
                fr = "0" * nd
            tok = "%s%d.%s" % (r.choice(["", "", "-", "+"]), ip, fr)
            other = r.choice(["1.5", "", "2", "NA", "0.000001", "0.00001", "1e3", "7.25", "x"])
            rows.append("%s,%s,%s" % (tok, other, r.choice(["a", "b", ""])))
        docs.append(("p,q,s\n" + "\n".join(rows) + "\n").encode())
    got = eng.parse_csv_many(docs, 1000)
    bad = [i for i, (g, d) in enumerate(zip(got, docs)) if g != oracle(d, 1000)]
    assert not bad, bad[:10]


def test_big_texts_become_python_strings_at_every_width(eng):
    """Texts above 1 MB are built as str objects of their final width (1, 2 or 4 bytes per code point)
    and filled by the library's threads instead of CPython's one-thread UTF-8 decoder: ASCII, Latin-1
    range, BMP and astral documents must come out equal to the oracle's strings."""
    for word in ("plain", "Würth é", "₹ 日本語", "ok 😀"):
        rows = "\n".join('P-%06d,"%s %d",%d.%02d' % (i, word, i % 97, i % 5000, i % 100) for i in range(60000))
        d = ("part,desc,price\n" + rows + "\n").encode("utf-8")
        got = eng.parse_csv_many([d], 10 ** 6)[0]
        want = oracle(d, 10 ** 6)
        assert len(got) == len(want) > (1 << 20) and got == want, word


def test_one_launch_scan_gives_the_same_texts(eng):
    """RFQ2_SCAN_SPLIT=0 selects the one-launch row scan (count look-back inside the kernel) instead of the
    marks / prefix / write launches: both must print the same bytes.  The switch is read once per process,
    so the other variant runs in a child."""
    import hashlib
    import os
    import subprocess
    import sys
    docs = [base64.b64decode(c["data"]) for c in json.loads((G / "ref_csv.json").read_text())]
    docs += [corpus.fuzz_csv(s) for s in range(41000, 41400)] + [corpus.rfq_csv(30000, seed=3)]

    def digest(texts):
        h = hashlib.sha256()
        for t in texts:
            h.update(b"\x00" if isinstance(t, Exception) else b"\x02" if t is None else b"\x01" + t.encode("utf-8"))
        return h.hexdigest()

    here = digest(eng.parse_csv_many(docs, 100000, on_unsupported="return"))
    child = (
        "import base64, json, hashlib, sys\n"
        "from pathlib import Path\n"
        "sys.path.insert(0, %r)\n"
        "from tests import corpus\n"
        "from replisense_rfq_b200 import Engine\n"
        "G = Path(%r)\n"
        "docs = [base64.b64decode(c['data']) for c in json.loads((G / 'ref_csv.json').read_text())]\n"
        "docs += [corpus.fuzz_csv(s) for s in range(41000, 41400)] + [corpus.rfq_csv(30000, seed=3)]\n"
        "h = hashlib.sha256()\n"
        "for t in Engine(0).parse_csv_many(docs, 100000, on_unsupported='return'):\n"
        "    h.update(b'\\x00' if isinstance(t, Exception) else b'\\x02' if t is None else b'\\x01' + t.encode('utf-8'))\n"
        "print('DIGEST', h.hexdigest())\n"
    ) % (str(Path(__file__).resolve().parent.parent), str(G))
    env = dict(os.environ, RFQ2_SCAN_SPLIT="0")
    out = subprocess.run([sys.executable, "-c", child], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [ln for ln in out.stdout.splitlines() if ln.startswith("DIGEST ")]
    assert line and line[-1].split()[1] == here
