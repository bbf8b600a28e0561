"""Quantity / price validators (app/rfq_agent.py:27-39) vs Python's own int() / float()."""
import math
import random

import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from replisense_rfq_b200 import Engine, build
    build.build()
    return Engine(0)


def _fields():
    r = random.Random(11)
    out = ["750", "1,200", "1,000,000", " 42 ", "+7", "-3", "007", "1_000", "1__0", "_1", "1_", "", " ", ",", "1.5",
           "12 000", "0x10", "9223372036854775807", "-9223372036854775808", "1e3", "abc", "1,2,3", "\t5\n",
           "125.50", "1,250.75", ".5", "5.", ".", "1e5", "1E-3", "1e", "e5", "+.5e+2", "inf", "-Infinity", "NaN",
           "nan", "infinit", "1_0.5_0", "1_.5", "1._5", "1e1_0", "0.1", "0.30000000000000004",
           "179769313486231570000000000000000000000", "1e400", "1e-400", "4.9e-324", "2.4703282292062328e-324",
           "9007199254740993", "123456789012345678901234567890.5", "₹500", "１２", "5 ", "--5", "+-5", "5-",
           "1\x1f", "\x1c2", "1\x002", "$100"]
    alpha = "0123456789" * 3 + ",,._eE+- \t" + "xinfa"
    for _ in range(6000):
        out.append("".join(r.choice(alpha) for _ in range(r.randint(0, 12))))
    for _ in range(2000):
        out.append(repr(r.uniform(-1e6, 1e6)))
        out.append("{:,.2f}".format(r.uniform(0, 1e7)))
        out.append("{:,}".format(r.randint(0, 10 ** 9)))
        out.append("%.17g" % (r.random() * 10 ** r.randint(-300, 300)))
    return out


def _py(f, kind):
    try:
        return (float if kind else int)(f.replace(",", ""))
    except ValueError:
        return ValueError


def test_validators_match_python(eng):
    from replisense_rfq_b200 import validators as V
    from replisense_rfq_b200.engine import EngineError
    fields = _fields()
    for kind, fn in ((0, V.parse_quantities), (1, V.parse_target_prices)):
        got = fn(eng, fields)
        bad, declined = [], 0
        for f, g in zip(fields, got):
            w = _py(f, kind)
            if isinstance(g, EngineError):
                declined += 1
                assert not f.isascii() or (kind == 0 and w is not ValueError and abs(w) >= 2 ** 63) or len(f) > 160, f
                continue
            if w is ValueError:
                ok = isinstance(g, ValueError)
            elif isinstance(g, Exception):
                ok = False
            elif kind and math.isnan(w):
                ok = math.isnan(g)
            else:
                ok = g == w and type(g) is type(w) and (not kind or math.copysign(1, g) == math.copysign(1, w))
            if not ok:
                bad.append((f, g, w))
        assert not bad, bad[:10]
        assert declined < 20
    assert V.parse_quantities(eng, [5, None, "7"]) == [5, None, 7]
    e = V.parse_quantities(eng, ["1.5"])[0]
    assert isinstance(e, ValueError)
    try:
        int("1.5")
    except ValueError as w:
        assert str(e) == str(w)
