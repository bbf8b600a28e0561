"""Host-side mirror of the reference's `RFQFieldGenerator` (app/rfq_agent.py:107-270) for the rows after
the parser, batched: blank check + truncation + prompt wrap (GPU,
prompt_kernels.cu), the LLM call replaced by a deterministic stand-in (`StubLLM`: the remote model is
network-bound and non-deterministic, out of scope by the north star), locating the JSON inside every reply
(GPU, json_kernels.cu; json.loads stays CPython's parser as in the reference), the quantity / price
validators of `LineItem` (GPU, numeric_kernels.cu), and the reference's response / fallback / error
dictionaries.  Same method names, argument meaning and messages as the reference.
"""
from __future__ import annotations

import hashlib
import json
import re
import threading
from typing import Any, Callable, Dict, Optional, Sequence

import numpy as np
from pydantic import BaseModel, ValidationError, field_validator

from . import prompt as P
from . import validators as V

_NONE = 0xFFFFFFFF
_numeric = threading.local()   # results of the batched GPU validators, consulted by the model validators


def _looked_up(kind: str, v: str):
    table = getattr(_numeric, "table", None)
    r = table.get((kind, v)) if table is not None else None
    if r is None:
        raise RuntimeError("LineItem validated outside RFQFieldGenerator.generate_many: its numeric fields go "
                           "through the GPU validators there (there is no CPU path)")
    if isinstance(r, Exception):   # ValueError as Python raises it -> pydantic's validation error; a field the GPU
        raise r                    # validators decline (EngineError) surfaces as the request's error response
    return r


class LineItem(BaseModel):                                   # app/rfq_agent.py:19-39
    part_number: Optional[str] = None
    description: Optional[str] = None
    quantity: Optional[int] = None
    target_price: Optional[float] = None
    currency: Optional[str] = None

    @field_validator("quantity", mode="before")
    @classmethod
    def parse_quantity(cls, v):
        if isinstance(v, str):
            return _looked_up("q", v)
        return v

    @field_validator("target_price", mode="before")
    @classmethod
    def parse_target_price(cls, v):
        if isinstance(v, str):
            return _looked_up("p", v)
        return v


class RFQResponse(BaseModel):                                # app/rfq_agent.py:41-59
    title: Optional[str] = None
    client_name: Optional[str] = None
    client_email: Optional[str] = None
    client_contact: Optional[str] = None
    client_phone: Optional[str] = None
    rfq_to: Optional[str] = None
    delivery_location: Optional[str] = None
    delivery_deadline: Optional[str] = None
    response_due_date: Optional[str] = None
    description: Optional[str] = None
    line_items: list[LineItem] = []
    requested_documents: list[str] = []
    confidence_score: float = 0.0
    missing_fields: list[str] = []
    requires_review: bool = True
    source_file: str = "email-body"
    success: bool = True
    message: str = ""


# ------------------------------------------------------------------ the deterministic stand-in for the LLM
class StubLLM:
    """Deterministic replacement of the remote model (`self.agent.generate_reply`, app/rfq_agent.py:166).

    A prompt found in `cached` gets the stored reply (the reference itself answers from its autogen
    cache, `.cache/42/cache.db`, when the prompt is known: `cache_seed: 42`).  Any other prompt gets a
    reply synthesised from the text inside it: the line items it can see (part number, quantity, price,
    currency), echoed as JSON, in one of the shapes `_extract_json_from_string` has strategies for —
    bare JSON, a ```json fence, or JSON inside a sentence — chosen by a hash of the text."""

    _PART = re.compile(r"\b(?=[A-Z0-9/\-]*\d)(?=[A-Z0-9/\-]*[A-Z])[A-Z0-9]{2,}(?:[-/][A-Z0-9]+)+\b")
    _NUM = re.compile(r"(?<![\w.])\d{1,3}(?:,\d{3})+(?:\.\d+)?|(?<![\w.])\d+(?:\.\d+)?")
    _CUR = re.compile(r"\b(USD|EUR|INR|GBP)\b|[$€£₹]")

    def __init__(self, cached: Optional[Dict[str, str]] = None):
        self.cached = dict(cached or {})

    @staticmethod
    def raw_text_of(prompt: str) -> str:
        return prompt.split('\n"""\n', 1)[1].rsplit('\n"""', 1)[0] if '\n"""\n' in prompt else prompt

    def reply(self, prompt: str) -> str:
        if prompt in self.cached:
            return self.cached[prompt]
        text = self.raw_text_of(prompt)
        items = []
        for line in text.split("\n"):
            m = self._PART.search(line)
            if not m:
                continue
            rest = line[m.end():]
            nums = self._NUM.findall(rest)
            cur = self._CUR.search(rest)
            ints = [x for x in nums if "." not in x]
            decs = [x for x in nums if "." in x]
            items.append({"part_number": m.group(), "description": None,
                          "quantity": ints[0] if ints else None,
                          "target_price": decs[0] if decs else (ints[1] if len(ints) > 1 else None),
                          "currency": cur.group() if cur else None})
            if len(items) == 25:
                break
        body = {"title": text.strip().split("\n", 1)[0][:80] or None, "client_name": None, "client_email": None,
                "client_contact": None, "client_phone": None, "rfq_to": None, "delivery_location": None,
                "delivery_deadline": None, "response_due_date": None, "description": None, "line_items": items,
                "requested_documents": [], "confidence_score": 0.9 if items else 0.4,
                "missing_fields": ["client_name", "client_email"], "requires_review": not items}
        js = json.dumps(body, ensure_ascii=False, indent=1)
        h = hashlib.md5(text.encode("utf-8")).digest()[0] % 4
        if h == 0:
            return js
        if h == 1:
            return "```json\n" + js + "\n```"
        if h == 2:
            return "Here is the extracted data:\n" + js + "\nLet me know if you need anything else."
        return "  \n" + js + "\n\t "


# ------------------------------------------------------------------ JSON extraction, batched
def extract_json_many(engine, replies: Sequence[str]) -> "list[Any]":
    """`_extract_json_from_string` (app/rfq_agent.py:208-236) for every reply: the parsed value, or the
    ValueError the reference raises.  Strategy 1 is json.loads of the whole string; for the replies it
    rejects, the spans of strategies 2 and 3 are found on the GPU in one batch."""
    out: "list[Any]" = [None] * len(replies)
    todo = []
    for i, s in enumerate(replies):
        try:
            out[i] = json.loads(s)
        except json.JSONDecodeError:
            todo.append(i)
    if not todo:
        return out
    raw = [replies[i].encode("utf-8") for i in todo]
    spans = engine.json_spans(raw)
    for k, i in enumerate(todo):
        first, last, c0, c1 = (int(x) for x in spans[k])
        b = raw[k]
        if first != _NONE and last != _NONE and first <= last:      # re.search(r'\{.*\}', s, re.DOTALL)
            try:
                out[i] = json.loads(b[first:last + 1].decode("utf-8"))
                continue
            except json.JSONDecodeError:
                pass
        try:
            out[i] = json.loads(b[c0:c1].decode("utf-8"))
        except json.JSONDecodeError as e:
            out[i] = ValueError(f"Unable to parse JSON from LLM response: {str(e)}")
    return out


def _create_error_response(error_message: str) -> Dict[str, Any]:          # app/rfq_agent.py:238-246
    return {"success": False, "error": error_message, "confidence_score": 0.0, "requires_review": True,
            "missing_fields": ["all"], "source_file": "unknown"}


def _create_fallback_response(raw_reply: dict, source_file: str, validation_error: str) -> Dict[str, Any]:   # :248-267
    fallback = RFQResponse()
    for field_name, field_value in raw_reply.items():
        if hasattr(fallback, field_name):
            try:
                setattr(fallback, field_name, field_value)
            except Exception:   # noqa: BLE001
                continue
    fallback.source_file = source_file
    fallback.success = True
    fallback.message = f"RFQ processed with validation warnings: {validation_error}"
    fallback.requires_review = True
    fallback.confidence_score = max(0.3, fallback.confidence_score)
    return fallback.model_dump()


class RFQFieldGenerator:
    """Drop-in for app/rfq_agent.py:107 with the remote model replaced by `llm` (default: StubLLM())."""

    def __init__(self, engine=None, llm: Optional[Callable[[str], str]] = None, device: int = 0):
        from .engine import default_engine
        self.engine = engine if engine is not None else default_engine(device)
        stub = StubLLM()
        self.llm = llm if llm is not None else stub.reply

    def generate(self, raw_text: str, source_file: str = "email-body") -> Dict[str, Any]:
        return self.generate_many([raw_text], [source_file])[0]

    async def generate_async(self, raw_text: str, source_file: str = "email-body") -> Dict[str, Any]:
        import asyncio
        return await asyncio.get_running_loop().run_in_executor(None, self.generate, raw_text, source_file)

    def generate_many(self, raw_texts: Sequence[str], source_files: Optional[Sequence[str]] = None) -> "list[Dict[str, Any]]":
        """`_generate_sync` (app/rfq_agent.py:136-182) for a batch: three GPU batches (prompts, JSON
        spans, numeric fields) around the stand-in model."""
        n = len(raw_texts)
        source_files = list(source_files) if source_files is not None else ["email-body"] * n
        prompts = P.build_prompts(self.engine, raw_texts)
        out: "list[Any]" = [None] * n
        live = [i for i in range(n) if prompts[i] is not None]
        for i in range(n):
            if prompts[i] is None:
                out[i] = _create_error_response(P.EMPTY_TEXT_ERROR)
        replies, failed = {}, {}
        for i in live:
            try:
                replies[i] = self.llm(prompts[i])
            except Exception as e:  # noqa: BLE001   (:178-182)
                failed[i] = _create_error_response(str(e))
        idx = [i for i in live if i in replies]
        strs = [i for i in idx if isinstance(replies[i], str)]
        parsed = dict(zip(strs, extract_json_many(self.engine, [replies[i] for i in strs])))
        # every string quantity / price of every reply through the GPU validators in one batch each
        qs, ps = [], []
        for i in idx:
            rep = parsed.get(i, replies[i])
            if isinstance(rep, dict) and isinstance(rep.get("line_items"), list):
                for li in rep["line_items"]:
                    if isinstance(li, dict):
                        if isinstance(li.get("quantity"), str):
                            qs.append(li["quantity"])
                        if isinstance(li.get("target_price"), str):
                            ps.append(li["target_price"])
        table = {}
        uq, up = sorted(set(qs)), sorted(set(ps))
        for v, r in zip(uq, V.parse_quantities(self.engine, uq) if uq else []):
            table[("q", v)] = r
        for v, r in zip(up, V.parse_target_prices(self.engine, up) if up else []):
            table[("p", v)] = r
        _numeric.table = table
        try:
            for i in idx:
                rep = parsed.get(i, replies[i])
                try:
                    if isinstance(rep, Exception):
                        raise rep
                    if not isinstance(rep, dict):
                        raise ValueError("LLM did not return a valid dictionary structure")
                    try:
                        validated = RFQResponse(**rep)
                        validated.source_file = source_files[i]
                        validated.success = True
                        validated.message = f"RFQ processed from {source_files[i]}"
                        out[i] = validated.model_dump()
                    except ValidationError as e:
                        out[i] = _create_fallback_response(rep, source_files[i], str(e))
                except Exception as e:  # noqa: BLE001   (:178-182)
                    out[i] = _create_error_response(str(e))
        finally:
            _numeric.table = None
        for i, r in failed.items():
            out[i] = r
        return out
