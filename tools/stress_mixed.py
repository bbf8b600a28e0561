import sys, threading, random, warnings
sys.path.insert(0, ".")
from replisense_rfq_b200 import _lib
from replisense_rfq_b200.engine import EnginePool
from tests import corpus
r = random.Random(3)
docs, kind, sst_of = [], [], []
for i in range(900):
    x = r.random()
    if x < 0.4:
        docs.append(corpus.rfq_csv(r.choice([10, 60, 400]), seed=i)); kind.append(_lib.KIND_CSV); sst_of.append(-1)
    elif x < 0.75:
        sx, ss = corpus.sheet_xml_doc(i)
        docs += [ss, sx]; kind += [_lib.KIND_SST, _lib.KIND_SHEET]; sst_of += [-1, len(docs) - 2]
    else:
        docs.append(corpus.text_body(i)); kind.append(_lib.KIND_TEXT); sst_of.append(-1)
pool = EnginePool(0, depth=2)
def one(eng, _):
    with eng.lock:
        res = eng.mixed_batch(docs, kind, 1000, sst_of)
        return [(int(res.status[d]), res.doc_bytes(d)) for d in range(res.n)]
ref = one(pool.engines[0], 0)
outs = pool.map(one, range(16))
print("sheet parts", sum(k in (_lib.KIND_SST, _lib.KIND_SHEET) for k in kind), "all equal:", all(o == ref for o in outs))
pool.close()
