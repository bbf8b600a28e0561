"""TEST INFRASTRUCTURE ONLY — CPU oracle for the rows after the parser that the reference owns itself
(SURVEY.md §8 f2): what RFQFieldGenerator does with the LLM's reply (app/rfq_agent.py:185-236, models
:19-58, fallback :248-267).  The module cannot be imported here (`autogen` is absent), so its lines are
restated over the same engines it calls: CPython's `json`, `re`, `int()`, `float()` and pydantic.
Pinned by the 14 replies the reference stored as values of its own .cache/42/cache.db
(tests/golden/cache_replies.json): every one must come out as CPython's json.loads makes it.
Only tests/, __graft_entry__.smoke() and bench.py may import this module.
"""
from __future__ import annotations

import json
import re
from typing import Any, Dict, Optional

from pydantic import BaseModel, ValidationError, field_validator


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
            return int(v.replace(",", ""))
        return v

    @field_validator("target_price", mode="before")
    @classmethod
    def parse_target_price(cls, v):
        if isinstance(v, str):
            return float(v.replace(",", ""))
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


def extract_json_from_string(response_str: str) -> dict:
    """app/rfq_agent.py:208-236, strategy by strategy."""
    try:
        return json.loads(response_str)                      # :212-215
    except json.JSONDecodeError:
        pass
    json_match = re.search(r'\{.*\}', response_str, re.DOTALL)   # :219
    if json_match:
        try:
            return json.loads(json_match.group())            # :221-224
        except json.JSONDecodeError:
            pass
    cleaned = response_str.strip()                           # :227-231
    if cleaned.startswith('```json'):
        cleaned = cleaned[7:]
    if cleaned.endswith('```'):
        cleaned = cleaned[:-3]
    try:
        return json.loads(cleaned.strip())                   # :233-236
    except json.JSONDecodeError as e:
        raise ValueError(f"Unable to parse JSON from LLM response: {str(e)}")


def create_error_response(error_message: str) -> Dict[str, Any]:   # :238-246
    return {"success": False, "error": error_message, "confidence_score": 0.0, "requires_review": True,
            "missing_fields": ["all"], "source_file": "unknown"}


def create_fallback_response(raw_reply: dict, source_file: str, validation_error: str) -> Dict[str, Any]:   # :248-267
    fallback = RFQResponse()
    for field_name, field_value in raw_reply.items():
        if hasattr(fallback, field_name):
            try:
                setattr(fallback, field_name, field_value)
            except Exception:   # noqa: BLE001  (the reference: bare except)
                continue
    fallback.source_file = source_file
    fallback.success = True
    fallback.message = f"RFQ processed with validation warnings: {validation_error}"
    fallback.requires_review = True
    fallback.confidence_score = max(0.3, fallback.confidence_score)
    return fallback.model_dump()


def parse_and_validate_response(reply: Any, source_file: str) -> Dict[str, Any]:   # :185-206
    if isinstance(reply, str):
        reply = extract_json_from_string(reply)
    if not isinstance(reply, dict):
        raise ValueError("LLM did not return a valid dictionary structure")
    try:
        validated = RFQResponse(**reply)
        validated.source_file = source_file
        validated.success = True
        validated.message = f"RFQ processed from {source_file}"
        return validated.model_dump()
    except ValidationError as e:
        return create_fallback_response(reply, source_file, str(e))


def generate(raw_text: str, reply_for, source_file: str = "email-body") -> Dict[str, Any]:
    """_generate_sync (:136-182) with the LLM call replaced by `reply_for(prompt)`."""
    from oracle.prompt_oracle import build_prompt
    from replisense_rfq_b200.prompt import EXTRACTION_PROMPT_TEMPLATE   # a constant (the template text), no product code
    prompt = build_prompt(raw_text, EXTRACTION_PROMPT_TEMPLATE)
    if prompt is None:
        return create_error_response("Empty or invalid input text")
    try:
        return parse_and_validate_response(reply_for(prompt), source_file)
    except Exception as e:  # noqa: BLE001
        return create_error_response(str(e))
