"""Prompt assembly (SURVEY.md §8 f1, app/rfq_agent.py:138-151): oracle vs the prompts stored in the
reference's cache.db, and the CUDA path vs the oracle."""
import hashlib
import json
import random
from pathlib import Path

import pytest

from oracle.prompt_oracle import build_prompt
from replisense_rfq_b200 import prompt as P

G = Path(__file__).parent / "golden"
GOLD = json.loads((G / "cache_prompts.json").read_text())


def _raw(prompt):
    return prompt.split('\n"""\n', 1)[1].rsplit('\n"""', 1)[0]


def test_template_is_the_references():
    assert hashlib.sha256(P.EXTRACTION_PROMPT_TEMPLATE.encode()).hexdigest() == GOLD["template_sha256"]


def test_oracle_reproduces_cached_prompts():
    assert len(GOLD["prompts"]) >= 4
    for rid, prompt in GOLD["prompts"].items():
        assert build_prompt(_raw(prompt), P.EXTRACTION_PROMPT_TEMPLATE) == prompt, rid


def test_oracle_truncation_and_blank():
    t = P.EXTRACTION_PROMPT_TEMPLATE
    assert build_prompt("", t) is None and build_prompt(" \n\t ", t) is None
    long = "é" * 8001
    p = build_prompt(long, t)
    assert p.endswith("é... [truncated]\n\"\"\"") and p.count("é") == 8000
    assert build_prompt("x" * 8000, t).count("truncated") == 0


def _texts():
    r = random.Random(5)
    cache = json.loads((G / "cache_rows.json").read_text())
    out = [_raw(p) for p in GOLD["prompts"].values()] + list(cache.values())
    out += ["", " ", "\n\t \r\n", "  　", "​", "x", " x ", "　é"]
    alphabet = ["a", "Z", " ", "\n", "é", "₹", "日", "😀", "\t", "0", ",", " "]
    for n in (1, 15, 16, 17, 4095, 4096, 4097, 7999, 8000, 8001, 8002, 12000, 40000):
        out.append("".join(r.choice(alphabet) for _ in range(n)))
        out.append("日" * n)
        out.append(" " * n + "x")
        out.append(" " * n)
    out.append("😀" * 7999 + "ab")
    out.append("😀" * 8000 + "é")
    out.append("a" * 200000)
    return out


@pytest.mark.gpu
def test_prompts_match_oracle_on_gpu():
    from replisense_rfq_b200 import Engine, build
    build.build()
    eng = Engine(0)
    texts = _texts()
    got = P.build_prompts(eng, texts)
    want = [build_prompt(t, P.EXTRACTION_PROMPT_TEMPLATE) for t in texts]
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, (bad[:10], len(texts[bad[0]]))
    for rid, prompt in GOLD["prompts"].items():
        assert P.build_prompts(eng, [_raw(prompt)]) == [prompt], rid


@pytest.mark.gpu
def test_prompts_after_parser_on_gpu(tmp_path):
    """parse -> prompt, the reference's /upload/ order (app/main.py:59 then rfq_agent)."""
    from replisense_rfq_b200 import Engine, build
    from tests import corpus
    build.build()
    eng = Engine(0)
    docs = [corpus.rfq_csv(n, seed=n) for n in (3, 40, 400)]
    raw = eng.parse_csv_many(docs, 1000)
    got = P.build_prompts(eng, raw)
    assert got == [build_prompt(t, P.EXTRACTION_PROMPT_TEMPLATE) for t in raw]
    assert got[2].endswith('... [truncated]\n"""')


@pytest.mark.gpu
def test_fused_csv_to_prompt_on_gpu():
    """rfq_csv_prompt_parse: parse and prompt assembly back to back on the device (raw_text is never
    downloaded) give the reference's prompt and text_length for every document, and the parse stage's
    verdict where it fails."""
    from oracle import file_parser_oracle as O
    from replisense_rfq_b200 import Engine, build, _lib
    from tests import corpus
    build.build()
    eng = Engine(0)
    docs = [corpus.rfq_csv(n, seed=n) for n in (3, 40, 400, 30000)] + [corpus.fuzz_csv(s) for s in range(70000, 70300)]
    docs += [b"a\n1\n", b"", "näme,₹\n日本語,₹500\nWürth,1\n".encode()]
    got = P.csv_to_prompts(eng, docs, 1000)
    n_ok = 0
    for d, (prompt, chars, status) in zip(docs, got):
        try:
            raw = O.parse_csv_bytes(d, 1000)
        except O.OracleParsingError:
            raw = None
        if status == _lib.DOC_UNSUPPORTED:
            assert corpus.bare_cr_blank_first(d)
            continue
        if raw is None:
            assert status == _lib.DOC_CSV_UNPARSEABLE and prompt is None
            continue
        assert status == _lib.DOC_OK
        assert prompt == build_prompt(raw, P.EXTRACTION_PROMPT_TEMPLATE)
        assert chars == len(raw) or chars == 2 ** 64 - 1   # ~0: the three-sweep path does not count
        n_ok += chars == len(raw)
    assert n_ok > 200
