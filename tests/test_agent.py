"""The rows after the parser the reference owns itself: JSON extraction from the
LLM's reply, the LineItem validators, response / fallback / error dictionaries (app/rfq_agent.py:185-267),
with the remote model replaced by the deterministic stand-in.  CPU tests pin the oracle on the replies
stored in the reference's own cache.db; GPU tests compare the CUDA-backed mirror with the oracle."""
import json
import random
import warnings
from pathlib import Path

import pytest

from oracle import agent_oracle as A
from oracle.prompt_oracle import build_prompt

G = Path(__file__).parent / "golden"
REPLIES = json.loads((G / "cache_replies.json").read_text())
PROMPTS = json.loads((G / "cache_prompts.json").read_text())["prompts"]


def _reply_variants():
    r = random.Random(3)
    base = list(REPLIES.values())
    out = list(base)
    for b in base[:8]:
        out += ["```json\n" + b + "\n```", "```json" + b + "```", "Sure! Here you go:\n" + b + "\nHope this helps {not json}",
                " \n\t" + b + " 　", "```json\n" + b, b + "\n```", "prefix " + b, b[:-1], b.replace('"', "'", 1),
                "``` json\n" + b + "\n```", " ```json\n" + b + "\n``` "]
    out += ["", " ", "{}", "[]", "[1, 2]", "42", "null", "{", "}", "}{", "no json here", "{\"a\": {\"b\": 1}} trailing }",
            "```json```", "```json\n```", "``````", "{\"x\": \"é₹日\"}", "text {\"x\": \"é₹日\"} more é", "﻿{\"a\": 1}",
            "{\"a\": 1}\x00", "{'a': 1}", "{\"a\": NaN}", "{\"a\": 1,}"]
    alphabet = ["{", "}", "`", "json", "\"a\"", ":", "1", ",", " ", "\n", "x", "[", "]", " "]
    for _ in range(400):
        out.append("".join(r.choice(alphabet) for _ in range(r.randint(0, 14))))
    return out


def _oracle_extract(s):
    try:
        return A.extract_json_from_string(s)
    except ValueError as e:
        return "ValueError: " + str(e)


def test_oracle_extraction_on_the_references_cached_replies():
    assert len(REPLIES) == 14
    for rid, reply in REPLIES.items():
        assert A.extract_json_from_string(reply) == json.loads(reply), rid
    assert A.extract_json_from_string("```json\n{\"a\": 1}\n```") == {"a": 1}
    assert A.extract_json_from_string("noise {\"a\": [1, 2]} noise") == {"a": [1, 2]}
    with pytest.raises(ValueError, match="Unable to parse JSON from LLM response"):
        A.extract_json_from_string("no json")


def test_oracle_validation_on_cached_replies():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        d13 = A.parse_and_validate_response(REPLIES["13"], "a.pdf")
        d4 = A.parse_and_validate_response(REPLIES["4"], "b.txt")
    assert d13["success"] and d13["message"] == "RFQ processed from a.pdf"
    assert [li["quantity"] for li in d13["line_items"]] == [2000, 1000, 800]          # "2,000" -> int
    assert [li["target_price"] for li in d13["line_items"]] == [48.0, 175.0, 1280.0]    # "1,280.00" -> float
    assert d4["message"].startswith("RFQ processed with validation warnings: 6 validation errors for RFQResponse")
    assert "invalid literal for int() with base 10: '2000 pcs'" in d4["message"]


def test_stub_is_deterministic_and_covers_the_strategies():
    from replisense_rfq_b200.rfq_agent import StubLLM
    from replisense_rfq_b200.prompt import EXTRACTION_PROMPT_TEMPLATE as T
    from tests import corpus
    stub = StubLLM({PROMPTS[k]: REPLIES[k] for k in PROMPTS})
    assert stub.reply(PROMPTS["12"]) == REPLIES["12"]
    shapes = set()
    for s in range(40):
        text = corpus.rfq_csv(5, seed=s).decode()
        p = build_prompt(text, T)
        a, b = stub.reply(p), stub.reply(p)
        assert a == b
        shapes.add(a[:1])
        d = A.extract_json_from_string(a)
        assert len(d["line_items"]) == 5 and d["line_items"][0]["part_number"].startswith("PRT-")
    assert len(shapes) >= 3


@pytest.mark.gpu
def test_extract_json_many_on_gpu():
    from replisense_rfq_b200 import Engine, build, extract_json_many
    build.build()
    eng = Engine(0)
    replies = _reply_variants()
    got = extract_json_many(eng, replies)
    for s, g in zip(replies, got):
        g = "ValueError: " + str(g) if isinstance(g, Exception) else g
        w = _oracle_extract(s)
        assert g == w or (g != g and w != w), repr(s)[:120]


@pytest.mark.gpu
def test_parse_prompt_stub_extract_validate_chain_on_gpu():
    """parse -> prompt -> (stand-in) reply -> JSON extraction -> validators, as one chain, against the oracle
    chain on the same texts: the reference's cached prompts and replies, CSV documents, blank text."""
    from replisense_rfq_b200 import Engine, RFQFieldGenerator, StubLLM, build
    from tests import corpus
    build.build()
    eng = Engine(0)
    stub = StubLLM({PROMPTS[k]: REPLIES[k] for k in PROMPTS})
    gen = RFQFieldGenerator(engine=eng, llm=stub.reply)
    texts = [PROMPTS[k].split('\n"""\n', 1)[1].rsplit('\n"""', 1)[0] for k in PROMPTS]
    texts += eng.parse_csv_many([corpus.rfq_csv(n, seed=n) for n in (4, 30, 300)], 1000)
    texts += ["", "   \n", "no parts here", "PRT-1-A qty 1,200 price 1,250.75 USD\nMP-1001 needs 2,000 pcs at 3.50"]
    files = ["f%d.csv" % i for i in range(len(texts))]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        got = gen.generate_many(texts, files)
        want = [A.generate(t, stub.reply, f) for t, f in zip(texts, files)]
    assert got == want
    assert got[0]["success"] and got[-4]["error"] == "Empty or invalid input text"
    # replies that are not valid, and numeric fields Python rejects
    bad = StubLLM()
    bad.reply = lambda p: "I could not find any JSON"    # noqa: E731
    g2 = RFQFieldGenerator(engine=eng, llm=bad.reply).generate("x", "y.txt")
    assert g2 == A.generate("x", bad.reply, "y.txt") and not g2["success"]
    weird = lambda p: json.dumps({"line_items": [{"quantity": "12 pcs", "target_price": "1,2,3.5"}, {"quantity": "007", "target_price": "1e3"}]})  # noqa: E731
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert RFQFieldGenerator(engine=eng, llm=weird).generate("x", "z") == A.generate("x", weird, "z")
