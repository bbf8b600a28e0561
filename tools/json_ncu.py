import json, random, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from replisense_rfq_b200 import Engine
r = random.Random(1)
def rfq(seed):
    r.seed(seed)
    items = [{"sl_no": i + 1, "part_number": "RJF%03d-%d" % (r.randint(0, 999), r.randint(1, 9)), "description": r.choice(["Bearing 6204-2RS", "O-ring 12x2 NBR"]),
              "quantity": r.randint(1, 5000), "target_price": round(r.uniform(0.5, 9000), 2), "currency": "INR", "notes": None} for i in range(r.randint(1, 30))]
    return json.dumps({"rfq_id": "RFQ-%06d" % seed, "line_items": items}).encode()
docs = [rfq(s) for s in range(8000)]
eng = Engine(0)
eng.parse_json_many(docs)
print("ok", eng.last_run_ms())
