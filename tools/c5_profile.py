"""Stage times of one rfq_mixed_parse over a C5-style batch (tools/bench_configs.run_c5 corpus)."""
import sys, time, random
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from replisense_rfq_b200 import Engine, build, _lib
from replisense_rfq_b200.engine import PinnedArena
from tests import corpus

build.build()
eng = Engine(0)
n_total = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
only = sys.argv[2] if len(sys.argv) > 2 else ""
r0 = random.Random(5)
distinct = []
for j in range(1000):
    x = r0.random()
    if x < 0.4:
        distinct.append(("csv", corpus.rfq_csv(r0.choice([10, 25, 60, 150, 400, 1000, 2000]), seed=j)))
    elif x < 0.7:
        distinct.append(("sheet", corpus.sheet_xml_doc(j)))
    else:
        distinct.append(("txt", corpus.text_body(j)))
docs, kind, sst_of = [], [], []
for g in range(n_total):
    k, p = distinct[g % 1000]
    if only and k != only:
        continue
    if k == "sheet":
        docs += [p[1], p[0]]; kind += [_lib.KIND_SST, _lib.KIND_SHEET]; sst_of += [-1, len(docs) - 2]
    else:
        docs.append(p); kind.append(_lib.KIND_CSV if k == "csv" else _lib.KIND_TEXT); sst_of.append(-1)
arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
off, ln, _ = arena.pack(docs)
for _ in range(2):
    eng.mixed_parse_arena(arena, off, ln, kind, 1000, sst_of)
eng.set_profile(True)
t0 = time.perf_counter()
r = eng.mixed_parse_arena(arena, off, ln, kind, 1000, sst_of)
t1 = time.perf_counter()
print("docs", len(docs), "in MB", sum(map(len, docs)) / 1e6, "out MB", r.total / 1e6, "call ms", (t1 - t0) * 1e3, "run ms", eng.last_run_ms())
agg = {}
for k, v in eng.profile():
    agg[k] = agg.get(k, 0) + v
for k, v in sorted(agg.items(), key=lambda x: -x[1])[:25]:
    print("  %-32s %8.3f ms" % (k, v))
