"""PDF text framing (SURVEY.md §8 f4, app/file_parser.py:166-212): the oracle against the framed text
the reference stored in its cache.db (row 13), and the CUDA path against the oracle."""
import json
import random
from pathlib import Path

import pytest

from oracle.framing_oracle import frame_pdf

G = Path(__file__).parent / "golden"
ROW13 = json.loads((G / "cache_rows.json").read_text())["13"]


def _pages_of_row13():
    parts = ROW13.split("\n\n=== Page 2 ===\n")
    return [(parts[0][len("=== Page 1 ===\n"):], [], 0), (parts[1], [], 0)]


def test_oracle_reproduces_the_cached_pdf_text():
    pages = _pages_of_row13()
    assert frame_pdf(pages) == ROW13
    # what fitz hands over carries surrounding blank space: strip() removes it
    assert frame_pdf([("\n  " + t + " \n\n", tb, im) for t, tb, im in pages]) == ROW13
    assert frame_pdf([]) == frame_pdf([("  \n", [], 0)]) == "PDF appears to be empty or contains no extractable text"


def _corpus():
    r = random.Random(4)
    from tests import corpus
    docs = [_pages_of_row13(), [], [("", [], 0)], [(" \n\t　", [], 0), ("x", [], 0)], [("", [], 2)],
            [("　é₹ 日本語\n", [[["a", None, "b"], [None, ""], ["", "q"], []], [], [["only"]]], 1)]]
    for s in range(120):
        pages = []
        for _ in range(r.randint(0, 5)):
            text = corpus.text_body(r.randint(0, 10 ** 6)).decode("utf-8", "ignore")
            if r.random() < 0.2:
                text = r.choice(["", "   ", "\n\n", "  "])
            if r.random() < 0.3:
                text = r.choice([" ", "\n", "\t\r\n", "　"]) + text + r.choice([" ", "\n\n", "\x1f"])
            tables = []
            for _ in range(r.randint(0, 2)):
                tables.append([[r.choice(["PRT-1", "", None, "12", "é", 5, 0, 2.5]) for _ in range(r.randint(0, 4))]
                               for _ in range(r.randint(0, 4))])
            pages.append((text, tables, r.choice([0, 0, 1, 3])))
        docs.append(pages)
    return docs


@pytest.mark.gpu
def test_framing_on_gpu_matches_oracle():
    from replisense_rfq_b200 import Engine, build
    from replisense_rfq_b200.framing import Page, frame_pdf_texts
    build.build()
    eng = Engine(0)
    docs = _corpus()
    got = frame_pdf_texts(eng, [[Page(t, tb, im) for t, tb, im in pages] for pages in docs])
    want = [frame_pdf(pages) for pages in docs]
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, (bad[:5], got[bad[0]][:200], want[bad[0]][:200])
    assert got[0] == ROW13
