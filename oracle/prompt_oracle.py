"""TEST INFRASTRUCTURE ONLY — CPU oracle for the prompt-assembly row.

Restates the text handling of RFQFieldGenerator._generate_sync (app/rfq_agent.py:138-151); the
class itself cannot be imported here (`autogen` is absent).  Pinned by the prompts stored as keys
of the reference's own .cache/42/cache.db (tests/golden/cache_prompts.json: rows 11-14 carry the
current template); the truncation branch has no golden there (longest stored text: 1917 chars) and
is pinned only by the reference's two lines it restates.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
from __future__ import annotations

from typing import Optional


def build_prompt(raw_text: str, template: str) -> Optional[str]:
    """None where the reference returns its "Empty or invalid input text" error response."""
    if not raw_text or not raw_text.strip():                 # app/rfq_agent.py:138
        return None
    if len(raw_text) > 8000:                                 # :143
        raw_text = raw_text[:8000] + "... [truncated]"       # :144
    return f"{template}\n\"\"\"\n{raw_text}\n\"\"\""         # :147
