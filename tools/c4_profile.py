"""Stage times of the text path on a C4-style batch (1 M e-mail / PDF-text bodies)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from replisense_rfq_b200 import Engine
from replisense_rfq_b200.engine import PinnedArena
from tests import corpus

n_total = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
distinct = [corpus.text_body(s) for s in range(20000)]
docs = [distinct[g % 20000] for g in range(n_total)]
eng = Engine(0)
arena = PinnedArena(eng, sum((len(d) + 15) & ~15 for d in docs) + 4096)
off, ln, _ = arena.pack(docs)
eng.stage(arena, off, ln)
for _ in range(3):
    eng.text_run()
eng.set_profile(True)
eng.text_run()
eng.fetch()
import numpy as np
print("docs", n_total, "in MB", float(np.sum(ln)) / 1e6, "run ms", eng.last_run_ms())
for k, v in eng.profile():
    print("  %-28s %8.3f ms" % (k, v))
