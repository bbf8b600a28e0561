"""Quantity / price validators on the GPU: LineItem.parse_quantity and
LineItem.parse_target_price (app/rfq_agent.py:27-39) for a whole batch of field strings — what the
reference's pydantic model does to every line item the LLM returns:

    quantity      int(v.replace(",", ""))      if isinstance(v, str) else v
    target_price  float(v.replace(",", ""))    if isinstance(v, str) else v

Each result is the Python value, or the ValueError Python would raise; fields this build does not
decide on the GPU (non-ASCII digits / blanks, integers beyond int64) raise EngineError instead of
guessing.
"""
from __future__ import annotations

from typing import Any, Sequence

import numpy as np

from . import _lib
from .engine import EngineError


def _run(engine, values: Sequence[Any], kind: int) -> "list[Any]":
    idx = [i for i, v in enumerate(values) if isinstance(v, str)]
    out = list(values)                      # non-strings pass through (`return v`)
    if not idx:
        return out
    raw, st = engine.numeric_batch([values[i].encode("utf-8") for i in idx], np.full(len(idx), kind, np.uint8))
    vals = raw.view(np.float64) if kind else raw.view(np.int64)
    for k, i in enumerate(idx):
        s = int(st[k])
        if s == _lib.DOC_OK:
            out[i] = float(vals[k]) if kind else int(vals[k])
        elif s == _lib.DOC_VALUE_ERROR:
            cleaned = values[i].replace(",", "")
            out[i] = ValueError(f"could not convert string to float: {cleaned!r}" if kind else
                                f"invalid literal for int() with base 10: {cleaned!r}")
        else:
            out[i] = EngineError(f"numeric field not decided on the GPU: {values[i]!r}")
    return out


def parse_quantities(engine, values: Sequence[Any]) -> "list[Any]":
    """LineItem.parse_quantity over a batch (app/rfq_agent.py:27-32)."""
    return _run(engine, values, 0)


def parse_target_prices(engine, values: Sequence[Any]) -> "list[Any]":
    """LineItem.parse_target_price over a batch (app/rfq_agent.py:34-39)."""
    return _run(engine, values, 1)
