"""Run on the GPU box: throughput of the .json path on many small documents and on a few large ones."""
import json
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from replisense_rfq_b200 import Engine  # noqa: E402
from tests import corpus  # noqa: E402


def main():
    eng = Engine(0)
    eng.set_profile(True)
    small = [corpus.json_doc(s) for s in range(20000)]
    big = [("[" + ",".join(corpus.json_doc(s).decode() for s in range(k * 300, k * 300 + 300)) + "]").encode() for k in range(64)]
    import random
    r = random.Random(1)

    def rfq(seed):
        r.seed(seed)
        items = [{"sl_no": i + 1, "part_number": "RJF%03d-%d" % (r.randint(0, 999), r.randint(1, 9)),
                  "description": r.choice(["Bearing 6204-2RS", "Seal \"V\" 25mm", "Würth DIN 912 M8x40", "O-ring 12x2 NBR"]),
                  "quantity": r.randint(1, 5000), "target_price": round(r.uniform(0.5, 9000), 2), "currency": r.choice(["INR", "EUR", "USD"]),
                  "delivery_date": "2025-%02d-%02d" % (r.randint(1, 12), r.randint(1, 28)), "notes": r.choice(["", "urgent", None])}
                 for i in range(r.randint(1, 30))]
        return json.dumps({"rfq_id": "RFQ-%06d" % seed, "customer": "ACME GmbH", "line_items": items}, ensure_ascii=False).encode()
    rfqs = [rfq(s) for s in range(50000)]
    for name, docs in (("rfq-like", rfqs), ("fuzz small", small), ("fuzz large", big)):
        nb = sum(len(d) for d in docs)
        eng.parse_json_many(docs[:8])
        t0 = time.perf_counter()
        r = eng.parse_json_many(docs)
        t1 = time.perf_counter()
        print("%s: %d docs, %.1f MB, call %.1f ms, device %.2f ms, %s" % (name, len(docs), nb / 1e6, (t1 - t0) * 1e3,
              eng.last_run_ms(), [(n, round(ms, 3)) for n, ms in eng.profile()]), flush=True)
        t0 = time.perf_counter()
        k = max(1, len(docs) // 5)
        for d in docs[:k]:
            json.dumps(json.loads(d.decode()), indent=2, ensure_ascii=False)
        print("   cpython json on %d docs: %.1f ms" % (k, (time.perf_counter() - t0) * 1e3), flush=True)


if __name__ == "__main__":
    main()
