"""Prompt assembly on the GPU: the text handling RFQFieldGenerator._generate_sync
applies to raw_text before the LLM call (app/rfq_agent.py:138-151) — blank check, truncation to
8000 code points + "... [truncated]", and wrapping in the extraction prompt.  The LLM call itself
is out of scope (network-bound, non-deterministic); `build_prompts` returns exactly the string the
reference would send.

The template is the reference's prompt text (app/rfq_agent.py:75-105): it is data the output is
defined by, so it is reproduced verbatim; tests/test_prompt.py pins it against the prompts stored
in the reference's own .cache/42/cache.db.
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np

EXTRACTION_PROMPT_TEMPLATE = """Extract the following fields from the RFQ text below and return ONLY valid JSON:
Required fields:
- title: Document title or subject
- client_name: Client/company name
- client_email: Contact email address
- client_contact: Contact person name
- client_phone: Phone number
- rfq_to: Who the RFQ is addressed to
- delivery_location: Where items should be delivered
- delivery_deadline: When delivery is needed
- response_due_date: When response is due
- description: Brief description of requirements
- line_items: Array of objects with part_number, description, quantity, target_price (numeric value only), currency (currency symbol or word as found in text)
- requested_documents: Array of required document types
- confidence_score: Float 0.0-1.0 indicating extraction confidence
- missing_fields: Array of field names that couldn't be extracted
- requires_review: Boolean indicating if human review is needed

IMPORTANT: For line_items, separate price and currency:
- target_price: Extract only the numeric value (e.g., "100", "50.75", "1000")
- currency: Extract currency as found in text - can be symbol ($, €, £, ₹) or word (dollars, euros, pounds, rupees) or code (USD, EUR, GBP, INR)
- Examples:
  * "$100" → target_price: "100", currency: "$"
  * "50 euros" → target_price: "50", currency: "euros"  
  * "75 GBP" → target_price: "75", currency: "GBP"
  * "₹500" → target_price: "500", currency: "₹"
  * "25" (no currency) → target_price: "25", currency: null

Return only valid JSON. Do not include any explanatory text.
Text to analyze:
"""

MAX_PROMPT_CHARS = 8000                      # app/rfq_agent.py:143
TRUNCATION_MARK = "... [truncated]"          # app/rfq_agent.py:144
PROMPT_HEAD = EXTRACTION_PROMPT_TEMPLATE + '\n"""\n'   # app/rfq_agent.py:147
PROMPT_TAIL = '\n"""'
EMPTY_TEXT_ERROR = "Empty or invalid input text"       # app/rfq_agent.py:140


def build_prompts(engine, texts: Sequence[str], template: str = EXTRACTION_PROMPT_TEMPLATE,
                  max_chars: int = MAX_PROMPT_CHARS) -> "list[Optional[str]]":
    """The prompt the reference sends for every raw_text; None where it answers
    EMPTY_TEXT_ERROR instead (empty or whitespace-only text).  One GPU batch."""
    docs = [t.encode("utf-8") for t in texts]
    if not docs:
        return []
    with engine.lock:   # the pinned output is ours until the strings have been built
        r = engine.prompt_batch(docs, (template + '\n"""\n').encode("utf-8"), PROMPT_TAIL.encode("utf-8"),
                                TRUNCATION_MARK.encode("utf-8"), max_chars)
        st = np.asarray(r.status)
        return [None if int(st[d]) else r.doc_str(d) for d in range(r.n)]


def csv_to_prompts(engine, docs: Sequence[bytes], max_rows: int = 1000, template: str = EXTRACTION_PROMPT_TEMPLATE,
                   max_chars: int = MAX_PROMPT_CHARS) -> "list[tuple[Optional[str], int, int]]":
    """What /upload/ needs of a CSV file (app/main.py:243-273): the prompt the reference would send and
    `text_length`, in ONE fused GPU call — raw_text is parsed and wrapped on the device and never
    downloaded.  Returns (prompt | None, len(raw_text) in code points, RFQ_DOC_* status) per document."""
    return engine.csv_prompts_many(docs, max_rows, (template + '\n"""\n').encode("utf-8"), PROMPT_TAIL.encode("utf-8"),
                                   TRUNCATION_MARK.encode("utf-8"), max_chars)
